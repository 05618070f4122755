"""Development: per-(row, warp) timeline of the direct tile scatter (library built with -DD3GA_SCATTER_TRACE).
   D3GA_LIB=variants/libd3ga_sctrace.so python tools/scatter_trace.py metric [cam] [views_in_flight]"""
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
if len(sys.argv) > 3:
    lib.d3ga_set_views_in_flight(int(sys.argv[3]))
for _ in range(3):
    got = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1])
R = 8192
buf = np.zeros((R, 4, 4), np.uint64)
lib.d3ga_debug_scatter_trace.argtypes = [C.c_void_p, C.c_int]
assert lib.d3ga_debug_scatter_trace(buf.ctypes.data_as(C.c_void_p), R) == 0
t0 = buf[..., 0].astype(np.int64); t1 = buf[..., 1].astype(np.int64); t2 = buf[..., 2].astype(np.int64)
live = t2 > 0
rows = int(live.any(axis=1).sum())
start = t0[live].min()
smid = (buf[..., 3] & np.uint64(0xffffffff)).astype(np.int64)
batches = (buf[..., 3] >> np.uint64(32)).astype(np.int64)
print("rows", rows, "kernel span us", (t2[live].max() - start) / 1e3, "D", got["num_rendered"])
s0 = (t0 - start) / 1e3; setup = (t1 - t0) / 1e3; body = (t2 - t1) / 1e3; end = (t2 - start) / 1e3
q = lambda a: np.round(np.percentile(a[live], [0, 10, 50, 90, 100]), 1)  # noqa: E731
print("start  us p0/10/50/90/100", q(s0))
print("set-up us", q(setup))
print("body   us", q(body))
print("end    us", q(end))
print("batches/warp", q(batches), " ns/batch", np.round(np.median(body[live] * 1e3 / np.maximum(1, batches[live]))))
per_warp = [np.round(np.median(body[:rows, w]), 1) for w in range(4)]
print("median body per warp index", per_warp, "median batches", [int(np.median(batches[:rows, w])) for w in range(4)])
per_sm = np.bincount(smid[:rows, 0], minlength=148)
print("CTAs per SM: min/max", per_sm.min(), per_sm.max(), "SMs used", int((per_sm > 0).sum()))
edges = np.arange(0, end[live].max() + 5, 5)
print("warps in flight per 5us:", [int((s0[live] < e + 5).sum() - (end[live] < e).sum()) for e in edges])
o = np.argsort(-body[:rows].max(axis=1))[:8]
for r in o:
    print("row", int(r), "start", np.round(s0[r, 0], 1), "body per warp", np.round(body[r], 1), "batches", batches[r], "sm", smid[r, 0])
o = np.argsort(body[:rows].max(axis=1))[:4]
for r in o:
    print("row", int(r), "start", np.round(s0[r, 0], 1), "body per warp", np.round(body[r], 1), "batches", batches[r], "sm", smid[r, 0])
