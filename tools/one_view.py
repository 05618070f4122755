"""Development helper: N views of a workload through the C-ABI (forward stages + backward), nothing else -- the
command ncu is pointed at.   python tools/one_view.py metric [views] [cam]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import gpu_util  # noqa: E402
import parity_util as util  # noqa: E402
from d3ga_b200 import _lib, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "metric"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ci = int(sys.argv[3]) if len(sys.argv) > 3 else 0
sc = synth.make_workload(name)
cam = synth.camera_ring(sc.W, sc.H, 160)[ci]
means, cov6 = util.scene_geometry(sc)
g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
w = np.random.default_rng(9).uniform(-1, 1, size=(3, sc.H, sc.W)).astype(np.float32)
for i in range(n):
    if i == n - 1:
        _lib.profile_enable(True)
    got = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1], invdepth=True)
    gg = gpu_util.backward(got, sc.W, sc.H, cam, w)
print({k: round(v[0] / v[1] * 1e3, 1) for k, v in _lib.profile_collect().items() if v[1]}, "D", got["num_rendered"])
