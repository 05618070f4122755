"""Development: per-kernel event times of a training frame, fused pair render against the two-call form.
   python tools/pair_probe.py [workload]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from d3ga_b200 import _lib, synth, training  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "metric"
dev = torch.device("cuda", 0)
scene = synth.make_workload(name)
cams = synth.camera_ring(scene.W, scene.H, 8)
avatar = training.Avatar(scene, dev)
bank = training.CameraBank(cams, dev)
bg = torch.ones(3, device=dev)
w = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (3, scene.H, scene.W)).astype(np.float32)).to(dev)
for fused in (True, False):
    for i in range(3):
        training.frame_step(avatar, bank, i, w, bg, fused_pair=fused)
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    for i in range(4):
        training.frame_step(avatar, bank, 3 + i, w, bg, fused_pair=fused)
    torch.cuda.synchronize()
    prof = _lib.profile_collect()
    _lib.profile_enable(False)
    print("fused" if fused else "two calls", {k: (round(v[0] / v[1] * 1e3, 1), v[1] // 4) for k, v in prof.items() if v[1]})
