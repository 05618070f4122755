"""Development helper: is the per-view loop host-bound?  Compares the time the host needs to ENQUEUE n views with the
time until the GPU has finished them, for the device-resident loop and the host-fed (e2e) loop."""
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
feed = training.HostFeed(wh, dev)
n = 16
for from_host in (False, True):
    kw = dict(from_host=True, feed=feed, loss_host=loss_h) if from_host else {}
    for _ in range(3):
        training.multi_view_step(avatar, bank, list(range(n)), w, bg, **kw)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for rep in range(4):
        training.multi_view_step(avatar, bank, list(range(n)), w, bg, **kw)
        t_enq = time.perf_counter()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    print(f"from_host={from_host}: {4 * n} views, enqueue done after {(t_enq - t0) * 1e3:.1f} ms, GPU done after {(t1 - t0) * 1e3:.1f} ms "
          f"-> {(t1 - t0) / (4 * n) * 1e6:.0f} us/view")
