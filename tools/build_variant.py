"""Build an A/B variant of the kernel library with extra -D defines (diagnostic).
usage: python tools/build_variant.py NAME -DLLIE_ROLL_PF=2 ...  ->  lowlight_image_enhancement_b200/_lib/variants/NAME.so"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lowlight_image_enhancement_b200 import build as B
from concurrent.futures import ThreadPoolExecutor

name, defs = sys.argv[1], sys.argv[2:]
out_dir = os.path.join(B.LIBDIR, "variants", name)
os.makedirs(out_dir, exist_ok=True)
objs = []
def one(src):
    obj = os.path.join(out_dir, os.path.basename(src)[:-3] + ".o")
    r = subprocess.run([B._nvcc(), *B.NVCC_FLAGS, *defs, "-I", B.INCLUDE, "-c", src, "-o", obj], capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr); raise SystemExit(1)
    return obj
with ThreadPoolExecutor(8) as ex:
    objs = list(ex.map(one, B.sources()))
lib = os.path.join(B.LIBDIR, "variants", name + ".so")
subprocess.check_call([B._nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib, *objs])
print(lib)
