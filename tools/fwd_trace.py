"""Development: per-(tile, warp) timeline of the forward blend (library built with -DD3GA_BLEND_TRACE).
   D3GA_LIB=variants/libd3ga_trace.so python tools/fwd_trace.py metric [cam]"""
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
ci = int(sys.argv[2]) if len(sys.argv) > 2 else 0
sc = synth.make_workload(name)
cam = synth.camera_ring(sc.W, sc.H, 160)[ci]
means, cov6 = util.scene_geometry(sc)
g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
for _ in range(3):
    got = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1])
n = ((sc.W + 15) // 16) * ((sc.H + 15) // 16)
buf = np.zeros((n, 8, 4), np.uint64)
lib = _lib.load()
assert lib.d3ga_debug_fwd_trace(buf.ctypes.data_as(C.c_void_p), n) == 0
t0 = buf[..., 0].astype(np.int64); t1 = buf[..., 1].astype(np.int64)
live = (t1 > 0) & (t0 > t0.max() - 2_000_000)   # the last launch only
start = t0[live].min()
dur = (t1 - t0) / 1e3
end = (t1 - start) / 1e3
chunks = (buf[..., 2] & np.uint64(0xffffffff)).astype(np.int64)
tile = (buf[..., 2] >> np.uint64(32)).astype(np.int64)
count = (buf[..., 3] >> np.uint64(32)).astype(np.int64)
surv = (buf[..., 3] & np.uint64(0xffffffff)).astype(np.int64)
print("kernel span us", end[live].max(), " warps", int(live.sum()))
order = np.argsort(-dur.reshape(-1))[:12]
for o in order:
    b, w = divmod(int(o), 8)
    print(f"slot {b:5d} warp {w} tile {tile[b, w]:5d} list {count[b, w]:6d} chunks {chunks[b, w]:4d} survivors {surv[b, w]:6d} "
          f"start {(t0[b, w] - start) / 1e3:7.1f} dur {dur[b, w]:7.1f} us  -> {dur[b, w] * 1e3 / max(1, surv[b, w]):.1f} ns/survivor "
          f"{dur[b, w] * 1e3 / max(1, chunks[b, w]):.0f} ns/chunk")
# occupancy over time: warps running per 10 us bucket
edges = np.arange(0, end[live].max() + 10, 10)
run = [(int(((t0[live] - start) / 1e3 < e + 10).sum() - (end[live] < e).sum())) for e in edges]
print("warps in flight per 10us:", run)
tot_s = surv[live].sum(); tot_c = chunks[live].sum()
print("total survivors", int(tot_s), "chunks", int(tot_c), "survivors/chunk", tot_s / tot_c)
