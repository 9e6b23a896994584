"""Per-step timeline of CTA (0, 0) of the tcgen05 recurrent kernel (library built with -DASB_RNN_TRACE)."""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from artspeech_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_var_rnn_trace.so")
from artspeech_b200 import models, synthetic

net = models.ArtSpeech(45, 10, precision="fast").cuda().eval()
x = synthetic.tokens(256, 100, 45, seed=5).cuda()
with torch.no_grad():
    for _ in range(2):
        net(x, [100] * 256)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (2 * 4096))()
fn = _lib.lib().asb_debug_rnn_trace
fn.argtypes = [ctypes.POINTER(ctypes.c_longlong)]
assert fn(buf) == 0
ev = []
for role in range(2):
    for i in range(2047):
        tag, t = buf[role * 4096 + 2 * i], buf[role * 4096 + 2 * i + 1]
        if t == 0:
            break
        ev.append((t, role, tag))
ev.sort()
names = {15: "EPI: step begins (out stored)", 16: "EPI: gi of this step in registers", 1: "MMA: h seen", 2: "MMA: issued + committed", 10: "EPI: gi prefetched, waiting", 11: "EPI: D seen",
         12: "EPI: D loaded", 13: "EPI: gates done, h stored", 14: "EPI: h handed over"}
t0 = ev[0][0]
for t, role, tag in ev[200:260]:
    print(f"{t - t0:8d}  {names[tag]}")
