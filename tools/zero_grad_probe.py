"""Development: which share of the Gaussians receives an exactly-zero gradient from one view (culled, or never blended:
behind the saturated surface)?   python tools/zero_grad_probe.py [workload] [cam]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import gpu_util  # noqa: E402
import parity_util as util  # noqa: E402
from d3ga_b200 import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "metric"
sc = synth.make_workload(name)
means, cov6 = util.scene_geometry(sc)
g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
w = np.random.default_rng(9).uniform(-1, 1, size=(3, sc.H, sc.W)).astype(np.float32)
for ci in ([int(sys.argv[2])] if len(sys.argv) > 2 else [0, 20, 45, 90]):
    cam = synth.camera_ring(sc.W, sc.H, 160)[ci]
    got = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1])
    gg = gpu_util.backward(got, sc.W, sc.H, cam, w)
    dm, dc = gg["means3D"], gg["cov3D"]
    zero = (np.abs(dm).max(axis=1) == 0) & (np.abs(dc).max(axis=1) == 0)
    print(f"camera {ci}: {zero.mean() * 100:.1f} % of {sc.N} Gaussians get an exactly-zero (means3D, cov3D) gradient; "
          f"culled {np.mean(got['radii'] <= 0) * 100:.1f} %")
