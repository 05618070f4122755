"""Development: per-(pass, tile) phase timeline of the depth sort (library built with -DD3GA_SORT_TRACE).
   D3GA_LIB=variants/libd3ga_sorttrace.so python tools/sort_trace.py metric [cam]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import gpu_util  # noqa: E402
import parity_util as util  # noqa: E402
from d3ga_b200 import _lib, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "metric"
ci = int(sys.argv[2]) if len(sys.argv) > 2 else 20
sc = synth.make_workload(name)
cam = synth.camera_ring(sc.W, sc.H, 160)[ci]
means, cov6 = util.scene_geometry(sc)
g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
lib = _lib.load()
for _ in range(3):
    got = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1])
buf = np.zeros((4, 512, 8), np.uint64)
lib.d3ga_debug_sort_trace.argtypes = [C.c_void_p]
assert lib.d3ga_debug_sort_trace(buf.ctypes.data_as(C.c_void_p)) == 0
t = buf.astype(np.int64)
names = ["ticket", "load", "rank", "lookback", "scan+reorder", "store"]
for p in range(4):
    live = t[p, :, 6] > 0
    if not live.any():
        print("pass", p, "skipped")
        continue
    tt = t[p, live]
    t0 = tt[:, 0].min()
    print(f"pass {p}: tiles {int(live.sum())}  span {(tt[:, 6].max() - t0) / 1e3:.1f} us   first start..last start {(tt[:, 0].max() - t0) / 1e3:.1f} us")
    for i, n in enumerate(names):
        d = (tt[:, i + 1] - tt[:, i]) / 1e3
        print(f"    {n:13s} median {np.median(d):5.2f}  p90 {np.percentile(d, 90):5.2f}  max {d.max():5.2f} us")
    end = (tt[:, 6] - t0) / 1e3
    print("    end of tile: p10/50/90/100", np.round(np.percentile(end, [10, 50, 90, 100]), 1))
    order = np.argsort(np.where(live)[0])
    lb = (tt[:, 4] - tt[:, 3]) / 1e3
    print("    look-back by tile index (every 32nd):", np.round(lb[::32], 2))
