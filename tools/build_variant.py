"""Development helper: build variants/libd3ga_<name>.so with extra nvcc flags for ONE source file (the other objects are
reused from the regular build).  Load it with D3GA_LIB=variants/libd3ga_<name>.so.
    python tools/build_variant.py <name> <file.cu> <nvcc flags...>"""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from d3ga_b200 import build as B  # noqa: E402

name, src, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
B.build_library()
root = os.path.dirname(B.HERE)
os.makedirs(os.path.join(root, "variants"), exist_ok=True)
objs = []
for n, extra in B.SOURCES:                      # <file.cu> = ALL rebuilds every source with the flags
    if src in (n, "ALL"):
        obj = os.path.join(root, "variants", f"{name}_{n.replace('.cu', '.o')}")
        subprocess.check_call([B.NVCC] + B.COMMON + extra + flags + ["-c", os.path.join(B.CSRC, n), "-o", obj])
    else:
        obj = os.path.join(B.OBJ, n.replace(".cu", ".o"))
    objs.append(obj)
out = os.path.join(root, "variants", f"libd3ga_{name}.so")
subprocess.check_call([B.NVCC, "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs)
print(out)
