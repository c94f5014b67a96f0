"""Dev tool: build tuning variants of the library into build_tune/ (git-ignored; travels with gpurun).
usage: python tools/build_tune.py name1="-DRMB_UNROLL=4" name2="-DRMB_TPB=64 -DRMB_MINB=2" ..."""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rigidmultiblobswall_b200 import _build
os.makedirs(os.path.join(ROOT, "build_tune"), exist_ok=True)

def one(arg):
    name, flags = arg.split("=", 1)
    so = os.path.join(ROOT, "build_tune", name + ".so")
    srcs = [os.path.join(_build.CSRC, s) for s in _build.SOURCES]
    subprocess.check_call([_build._nvcc()] + _build.NVCC_FLAGS + flags.split() + ["-o", so] + srcs, cwd=_build.CSRC)
    return so

with ThreadPoolExecutor(4) as ex:
    for so in ex.map(one, sys.argv[1:]):
        print("built", so)
