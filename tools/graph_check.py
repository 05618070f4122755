"""Development helper: graphed vs eager multi-view step -- same gradients, time per view."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from d3ga_b200 import synth, training  # noqa: E402

dev = torch.device("cuda", 0)
scene = synth.make_workload(sys.argv[1] if len(sys.argv) > 1 else "metric")
cams = synth.camera_ring(scene.W, scene.H, 160)
avatar = training.Avatar(scene, dev)
bank = training.CameraBank(cams, dev)
bg = torch.ones(3, device=dev)
wh = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (3, scene.H, scene.W)).astype(np.float32)).pin_memory()
w = wh.to(dev)
loss_h = torch.zeros((), dtype=torch.float32).pin_memory()
feed = training.HostFeed(wh, dev, cam_host=bank.host, tetpoints_host=avatar.tetpoints_host)
ids = list(range(16))
tot_e = training.multi_view_step(avatar, bank, ids, w, bg)
g_e = avatar.grads.flat.clone()
gv = training.GraphedViewStep(avatar, cams[0], bg)
tot_g = training.multi_view_step_graphed(avatar, bank, ids, gv, weights=w).clone()
g_g = avatar.grads.flat.clone()
torch.cuda.synchronize()
rel = float((g_e - g_g).norm() / g_e.norm())
print(f"loss eager {float(tot_e):.6e} graphed {float(tot_g):.6e}; grad rel-L2 diff {rel:.2e}; captures {gv.captures}, "
      f"launches/replay {gv.launches_per_replay}")
for name, fn in (("eager", lambda: training.multi_view_step(avatar, bank, ids, w, bg)),
                 ("graphed", lambda: training.multi_view_step_graphed(avatar, bank, ids, gv, weights=w)),
                 ("eager e2e", lambda: training.multi_view_step(avatar, bank, ids, w, bg, from_host=True, feed=feed, loss_host=loss_h)),
                 ("graphed e2e", lambda: training.multi_view_step_graphed(avatar, bank, ids, gv, from_host=True, feed=feed, loss_host=loss_h))):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    print(f"{name}: {(time.perf_counter() - t0) / 64 * 1e6:.0f} us/view (captures so far {gv.captures})")
# upload bandwidth of this box (pinned -> device, one 3xHxW image)
buf = torch.empty_like(w)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    buf.copy_(wh, non_blocking=True)
b.record()
torch.cuda.synchronize()
print(f"H2D: {wh.numel() * 4 * 10 / (a.elapsed_time(b) * 1e-3) / 1e9:.1f} GB/s for {wh.numel() * 4 / 1e6:.1f} MB images")
# two lanes
LANES = int(os.environ.get("LANES", "2"))
pv = training.PipelinedViews(avatar, cams[0], bg, lanes=LANES)
feed4 = training.HostFeed(wh, dev, cam_host=bank.host, tetpoints_host=avatar.tetpoints_host, slots=2 * LANES)
tot_p = training.multi_view_step_pipelined(avatar, bank, ids, pv, weights=w).clone()
torch.cuda.synchronize()
print(f"pipelined: loss {float(tot_p):.6e}, grad rel-L2 diff vs eager {float((g_e - avatar.grads.flat).norm() / g_e.norm()):.2e}, captures {pv.captures}")
for name, fn in (("pipelined", lambda: training.multi_view_step_pipelined(avatar, bank, ids, pv, weights=w)),
                 ("pipelined e2e", lambda: training.multi_view_step_pipelined(avatar, bank, ids, pv, from_host=True, feed=feed4, loss_host=loss_h))):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    print(f"{name}: {(time.perf_counter() - t0) / 64 * 1e6:.0f} us/view (captures so far {pv.captures})")
