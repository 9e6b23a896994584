"""Build tools/_var_<name>.so: the library with ONE source recompiled with extra -D flags.
usage: python tools/build_variant.py <name> <source.cu> -DFOO=1 ..."""
import os
import subprocess
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from artspeech_b200 import build

name, src, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
build.build_library()
here = os.path.dirname(os.path.abspath(__file__))
obj = os.path.join(here, f"_var_{name}.o")
src_path = os.path.join(build.CSRC, src)
subprocess.check_call([build._nvcc(), *build.NVCC_FLAGS, *flags, "-c", src_path, "-o", obj])
objs = [os.path.join(build.OBJ_DIR, os.path.basename(s)[:-3] + ".o") for s in build.sources() if os.path.basename(s) != src]
out = os.path.join(here, f"_var_{name}.so")
subprocess.check_call([build._nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out, obj, *objs, "-lcudart"])
os.remove(obj)
print(out)
