"""Timeline of CTA 0 of the tcgen05 decoder (library built by tools/build_variant.py with -DASB_DEC_TRACE)."""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), os.environ.get("ASB_TRACE_LIB", "_var_dec_trace.so"))
from artspeech_b200 import models, synthetic

net = models.ArtSpeech(45, 10, precision="fast").cuda().eval()
x = synthetic.tokens(256, 100, 45, seed=5).cuda()
L = [100] * 256
with torch.no_grad():
    for _ in range(3):
        net(x, L)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (8 * 4096))()
fn = getattr(_lib.lib(), os.environ.get("ASB_TRACE_FN", "asb_debug_decoder16_trace"))   # tf32 kernel: asb_debug_decoder_trace
fn.argtypes = [ctypes.POINTER(ctypes.c_longlong)]
assert fn(buf) == 0
ev = []
for role in range(8):
    for i in range(2047):
        tag, t = buf[role * 4096 + 2 * i], buf[role * 4096 + 2 * i + 1]
        if t == 0:
            break
        ev.append((t, role, tag))
ev.sort()
t0 = ev[0][0]
names = {99: "commit D", 200: "D ready seen", 201: "chunk signalled", 300: "heads begin", 301: "heads end"}
limit = int(sys.argv[1]) if len(sys.argv) > 1 else 260
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for t, role, tag in ev[skip:skip + limit]:
    if tag >= 600:
        what = {600: "heads D seen", 603: "heads D read (freed)", 604: "contours stored", 605: "columns 0-31 out", 606: "columns 32-63 out", 607: "columns 64-95 out"}[tag]
    elif tag >= 500:
        what = f"weights in slot {tag - 500} seen"
    elif tag >= 400:
        what = f"copy into slot {tag - 400} issued"
    else:
        what = names.get(tag, f"gemm N={tag - 100} begin" if tag >= 100 else f"chunk {tag - 50} MMAs issued" if tag >= 50
                         else f"A chunk {tag - 30} in TMEM seen" if tag >= 30 else f"A chunk {tag - 10} ready (weights too)")
    print(f"{t - t0:8d}  {['MMA ', 'EPI0', 'EPI1', 'PROD', 'OUT0', 'OUT1', 'OUT2', 'OUT3'][role]}  {what}")
