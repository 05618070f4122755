"""Development helper: run N views of a workload with a library built with -DD3GA_BLEND_STATS and print the counters.
    D3GA_LIB=variants/libd3ga_stats.so python tools/blend_stats.py [workload] [views]"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from d3ga_b200 import _lib, synth, training  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "metric"
views = int(sys.argv[2]) if len(sys.argv) > 2 else 4
dev = torch.device("cuda", 0)
lib = _lib.load()
scene = synth.make_workload(name)
cams = synth.camera_ring(scene.W, scene.H, 160)
avatar = training.Avatar(scene, dev)
bank = training.CameraBank(cams, dev)
bg = torch.ones(3, device=dev)
w = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (3, scene.H, scene.W)).astype(np.float32)).to(dev)
training.multi_view_step(avatar, bank, [0], w, bg)
torch.cuda.synchronize()
out = (ctypes.c_ulonglong * 8)()
lib.d3ga_debug_blend_stats(out, 1)
training.multi_view_step(avatar, bank, list(range(1, 1 + views)), w, bg)
torch.cuda.synchronize()
lib.d3ga_debug_blend_stats(out, 1)
from d3ga_b200 import diff_gaussian_rasterization as dgr
D = dgr.last_stats.get("num_rendered", 0)
v = [x / views for x in out]
print(f"{name}: D={D} per view: fwd warp-steps {v[0]:.3g} (x32 = {32*v[0]:.3g} entries looked at, {32*v[0]/max(D,1):.2f} x D) survivors {v[1]:.3g} "
      f"({v[1]/max(32*v[0],1):.1%}) groups {v[2]:.3g} passing pairs {v[3]:.3g} ({v[3]/max(v[1]*32,1):.1%} of survivor x 32 px)")
print(f"          bwd warp-steps {v[4]:.3g} (x32 = {32*v[4]:.3g}, {32*v[4]/max(D,1):.2f} x D) survivors {v[5]:.3g} ({v[5]/max(32*v[4],1):.1%}) "
      f"groups {v[6]:.3g} passing pairs {v[7]:.3g} ({v[7]/max(v[5]*32,1):.1%})")

# per-(tile, warp) trace of the forward of ONE view
tiles = ((scene.W + 15) // 16) * ((scene.H + 15) // 16)
buf = torch.zeros(tiles * 8 * 4, dtype=torch.int64, device=dev)
lib.d3ga_debug_blend_trace.argtypes = [ctypes.c_void_p]
lib.d3ga_debug_blend_trace(buf.data_ptr())
from d3ga_b200 import diff_gaussian_rasterization as dgr2
training.multi_view_step(avatar, bank, [5], w, bg)
torch.cuda.synchronize()
lib.d3ga_debug_blend_trace(None)
t = buf.cpu().numpy().reshape(-1, 4)
t = t[t[:, 0] > 0]
t0, t1 = t[:, 0].min(), t[:, 1].max()
dur = (t[:, 1] - t[:, 0]).astype(np.float64)
steps = (t[:, 2] & 0xffffffff).astype(np.float64)
surv = (t[:, 3] & 0xffffffff).astype(np.float64)
sm = t[:, 2] >> 32
print(f"forward trace: span {(t1 - t0) / 1e3:.1f} us, {len(t)} warp work items, sum of warp times {dur.sum() / 1e3:.0f} us "
      f"= {dur.sum() / (t1 - t0) / (148 * 32):.1%} of 148x32 warp slots; longest warp {dur.max() / 1e3:.1f} us "
      f"({steps[dur.argmax()]:.0f} steps, {surv[dur.argmax()]:.0f} survivors)")
order = np.argsort(-dur)[:8]
print("  longest:", [(round(dur[i] / 1e3, 1), int(steps[i]), int(surv[i]), round((t[i, 0] - t0) / 1e3, 1)) for i in order], "(us, steps, survivors, start us)")
ns_per_step = dur[steps > 20] / steps[steps > 20]
print(f"  ns per step (warps with > 20 steps): median {np.median(ns_per_step):.0f}, p10 {np.percentile(ns_per_step, 10):.0f}, p90 {np.percentile(ns_per_step, 90):.0f}; "
      f"survivors per step median {np.median(surv[steps > 20] / steps[steps > 20]):.1f}")
# how much of the span has few warps running?
ev = np.concatenate([np.stack([t[:, 0], np.ones(len(t))], 1), np.stack([t[:, 1], -np.ones(len(t))], 1)])
ev = ev[np.argsort(ev[:, 0])]
run = np.cumsum(ev[:, 1])
dt = np.diff(ev[:, 0], append=ev[-1, 0])
for thr in (148 * 32 * 0.5, 148 * 32 * 0.25, 148 * 4, 148):
    print(f"  time with fewer than {thr:.0f} warps running: {dt[run < thr].sum() / 1e3:.1f} us")
