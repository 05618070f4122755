"""Development: per-(tile, warp) timeline of the backward blend (library built with -DD3GA_BLEND_TRACE).
   D3GA_LIB=variants/libd3ga_trace.so python tools/bwd_trace.py metric [cam]"""
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
w = np.random.default_rng(9).uniform(-1, 1, size=(3, sc.H, sc.W)).astype(np.float32)
for _ in range(3):
    got = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1], invdepth=True)
    gg = gpu_util.backward(got, sc.W, sc.H, cam, w)
n = ((sc.W + 15) // 16) * ((sc.H + 15) // 16)
buf = np.zeros((n, 8, 4), np.uint64)
lib = _lib.load()
lib.d3ga_debug_bwd_trace.argtypes = [C.c_void_p, C.c_int]
assert lib.d3ga_debug_bwd_trace(buf.ctypes.data_as(C.c_void_p), n) == 0
t0 = buf[..., 0].astype(np.int64); t1 = buf[..., 1].astype(np.int64)
live = (t1 > 0) & (t0 > t0.max() - 2_000_000)   # the last launch only (a slot's warp may have had no work this time)
start = t0[live].min()
dur = np.where(live, (t1 - t0) / 1e3, 0.0)
end = (t1 - start) / 1e3
nproc = (buf[..., 2] & np.uint64(0xffffffff)).astype(np.int64)
tile = (buf[..., 2] >> np.uint64(32)).astype(np.int64)
surv = buf[..., 3].astype(np.int64)
print("kernel span us", end[live].max(), " warps that worked", int(live.sum()), "of", n * 8)
cta_live = live.any(axis=1)
cta_max = dur.max(axis=1)[cta_live]; cta_sum = dur.sum(axis=1)[cta_live]
print("CTAs with work", int(cta_live.sum()), " warp-time / (8 x slowest warp) per CTA: mean", round(float((cta_sum / (8 * cta_max)).mean()), 3),
      " weighted by CTA time", round(float(cta_sum.sum() / (8 * cta_max.sum())), 3))
print("sum of warp-us", round(float(dur.sum())), " sum of 8 x CTA-us", round(float(8 * cta_max.sum())))
order = np.argsort(-dur.reshape(-1))[:10]
for o in order:
    b, w_ = divmod(int(o), 8)
    print(f"slot {b:5d} warp {w_} tile {tile[b, w_]:5d} walked {nproc[b, w_]:6d} survivors {surv[b, w_]:6d} start {(t0[b, w_] - start) / 1e3:7.1f} "
          f"dur {dur[b, w_]:7.1f} us -> {dur[b, w_] * 1e3 / max(1, surv[b, w_]):.1f} ns/survivor; CTA warps {np.round(dur[b], 0)}")
edges = np.arange(0, end[live].max() + 10, 10)
print("warps in flight per 10us:", [int(((t0[live] - start) / 1e3 < e + 10).sum() - (end[live] < e).sum()) for e in edges])
print("total survivors", int(surv[live].sum()), " ns/survivor overall (warp-time)", round(float(dur.sum() * 1e3 / surv[live].sum()), 1))
