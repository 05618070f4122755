"""Generates tests/golden/deform_golden.npz by executing the REFERENCE's own Python functions on CPU tensors.

The reference modules hard-code device="cuda" (utils/general_utils.py:25,63,82) and tetra.py imports meshio/trimesh at
module scope, so this script (a) registers empty stub modules for the missing imports and (b) strips the device
argument from torch.zeros while the reference code runs.  The functions executed are the reference's, unmodified:
  utils/general_utils.py: build_rotation, build_scaling_rotation, strip_symmetric
  submodules/tetra-sampler/tetra_sampler/tetra.py: Tetra.gradient
The composition (models/cage_net.py:218-230, lib/cage.py:329,339-342) is written out below with those functions.

    python tests/golden/make_deform_golden.py      (needs /root/reference)
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.environ.get("D3GA_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
from d3ga_b200 import synth  # noqa: E402

for name in ("meshio", "trimesh"):
    sys.modules.setdefault(name, types.ModuleType(name))


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


gu = load(os.path.join(REF, "utils", "general_utils.py"), "ref_general_utils")
tetra = load(os.path.join(REF, "submodules", "tetra-sampler", "tetra_sampler", "tetra.py"), "ref_tetra")

_zeros = torch.zeros


def zeros_cpu(*a, **k):
    k.pop("device", None)
    return _zeros(*a, **k)


torch.zeros = zeros_cpu
try:
    sc = synth.make_scene("golden", 257, 96, 64, 64, seed=3)
    t = torch.from_numpy
    tetpoints, tets, tid = t(sc.tetpoints), t(sc.tets), t(sc.tetra_id)
    canon = t(sc.canon_points)
    barys, scales, rot = t(sc.barys), t(sc.scales), t(sc.rotation) * 1.7   # un-normalised on purpose
    gradient = lambda X: tetra.Tetra.gradient(None, X)  # noqa: E731
    canonical_gradient = torch.linalg.inv(gradient(canon[tets][tid]))          # lib/cage.py:329
    deformed_tetras = tetpoints[tets]                                          # cage_net.py:218
    J = gradient(deformed_tetras[tid]) @ canonical_gradient                    # lib/cage.py:340-341
    L = gu.build_scaling_rotation(scales, rot)                                 # cage_net.py:162
    cov = L @ L.transpose(1, 2)                                                # cage_net.py:163
    cov3D = J @ cov @ J.transpose(1, 2)                                        # cage_net.py:222
    cov6 = gu.strip_symmetric(cov3D)                                           # cage_net.py:223
    means3D = torch.einsum("ikj,ik->ij", deformed_tetras[tid], barys)          # cage_net.py:228-230
    R = gu.build_rotation(rot)
finally:
    torch.zeros = _zeros

np.savez(os.path.join(ROOT, "tests", "golden", "deform_golden.npz"), tetpoints=sc.tetpoints, canon=sc.canon_points,
         tets=sc.tets, tetra_id=sc.tetra_id, barys=sc.barys, scales=sc.scales, rotation=rot.numpy(),
         canonical_gradient=canonical_gradient.numpy(), R=R.numpy(), L=L.numpy(), J=J.numpy(), cov6=cov6.numpy(),
         means3D=means3D.numpy())
print("wrote deform_golden.npz", means3D.shape, cov6.shape)
