"""Development: the deformation launch pair at a workload's size (what ncu is pointed at).
    python tools/deform_probe.py metric [sorted] [full] [sparse]      full = with the three deltas and the cage energy, like bench.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from d3ga_b200 import _lib, synth  # noqa: E402
from d3ga_b200.tetra_sampler import deform_gaussians  # noqa: E402

sc = synth.make_workload(sys.argv[1] if len(sys.argv) > 1 else "metric")
if "sorted" in sys.argv[2:]:
    sc = synth.sort_scene_by_tet(sc)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
tp, ba, scl, rot = (t(a).requires_grad_(True) for a in (sc.tetpoints, sc.barys, np.log(sc.scales), sc.rotation))
tets, tid, cg = t(sc.tets), t(sc.tetra_id), t(sc.canonical_gradient)
gm, gc = torch.randn(sc.N, 3, device="cuda"), torch.randn(sc.N, 6, device="cuda")
if "sparse" in sys.argv[2:]:      # like a rendered view: ~90 % of the Gaussians get an exactly-zero incoming gradient
    dead = torch.rand(sc.N, device="cuda") < 0.9
    gm[dead] = 0
    gc[dead] = 0
full = "full" in sys.argv[2:]
if full:
    from oracle import deform_oracle as do
    dn_inv = torch.linalg.inv(do.tet_gradient(torch.from_numpy(sc.canon_points)[torch.from_numpy(sc.tets)])).cuda().contiguous()
    d4, d3, dr = (torch.zeros(sc.N, k, device="cuda").requires_grad_(True) for k in (4, 3, 4))
    ge = torch.ones(sc.tets.shape[0], device="cuda")
for i in range(4):
    if i == 3:
        _lib.profile_enable(True)
    if full:
        m, c, e = deform_gaussians(tp, tets, tid, ba, cg, scl, rot, delta_bary=d4, delta_scale=d3, delta_rot=dr,
                                   log_scales=True, dn_inv=dn_inv)
        torch.autograd.backward([m, c, e], [gm, gc, ge])
    else:
        m, c = deform_gaussians(tp, tets, tid, ba, cg, scl, rot, log_scales=True)
        torch.autograd.backward([m, c], [gm, gc])
torch.cuda.synchronize()
print({k: round(v[0] / v[1] * 1e3, 1) for k, v in _lib.profile_collect().items() if v[1]})
