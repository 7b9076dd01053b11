"""Build the library with extra compile-time defines into stereo_matching_b200/_variants/<name>.so (for A/B runs on
one GPU box: `STEREO_B200_LIB=stereo_matching_b200/_variants/<name>.so python ...`).
`python scripts/build_variant.py <name> -DSTEREO_MBAR_HINT=4096 ...`"""
import concurrent.futures
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereo_matching_b200 import build as b          # noqa: E402

name, defs = sys.argv[1], sys.argv[2:]
out_dir = os.path.join(b.HERE, "_variants")
obj_dir = os.path.join(out_dir, "_obj_" + name)
os.makedirs(obj_dir, exist_ok=True)
nvcc = b._nvcc()
srcs = sorted(glob.glob(os.path.join(b.CSRC, "*.cu")))


def one(src):
    obj = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".o")
    p = subprocess.run([nvcc] + b.NVCC_FLAGS + defs + ["-c", src, "-o", obj], capture_output=True, text=True)
    if p.returncode:
        sys.exit(p.stdout + p.stderr)
    return obj


with concurrent.futures.ThreadPoolExecutor(8) as ex:
    objs = list(ex.map(one, srcs))
lib = os.path.join(out_dir, name + ".so")
subprocess.run([nvcc, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcuda"], check=True)
print(lib)
