"""Development: the deformation launch pair at a workload's size (what ncu is pointed at).  python tools/deform_probe.py metric"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from d3ga_b200 import _lib, synth  # noqa: E402
from d3ga_b200.tetra_sampler import deform_gaussians  # noqa: E402

sc = synth.make_workload(sys.argv[1] if len(sys.argv) > 1 else "metric")
if len(sys.argv) > 2 and sys.argv[2] == "sorted":
    sc = synth.sort_scene_by_tet(sc)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
tp, ba, scl, rot = (t(a).requires_grad_(True) for a in (sc.tetpoints, sc.barys, np.log(sc.scales), sc.rotation))
tets, tid, cg = t(sc.tets), t(sc.tetra_id), t(sc.canonical_gradient)
gm, gc = torch.randn(sc.N, 3, device="cuda"), torch.randn(sc.N, 6, device="cuda")
for i in range(4):
    if i == 3:
        _lib.profile_enable(True)
    m, c = deform_gaussians(tp, tets, tid, ba, cg, scl, rot, log_scales=True)
    torch.autograd.backward([m, c], [gm, gc])
torch.cuda.synchronize()
print({k: round(v[0] / v[1] * 1e3, 1) for k, v in _lib.profile_collect().items() if v[1]})
