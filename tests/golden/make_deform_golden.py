"""Generates tests/golden/deform_golden.npz by executing the REFERENCE's own Python functions on CPU tensors.

The reference modules hard-code device="cuda" (utils/general_utils.py:25,63,82) and tetra.py imports meshio/trimesh at
module scope, so this script (a) registers empty stub modules for the missing imports and (b) strips the device
argument from torch.zeros while the reference code runs.  The functions executed are the reference's, unmodified:
  utils/general_utils.py: build_rotation, build_scaling_rotation, strip_symmetric
  submodules/tetra-sampler/tetra_sampler/tetra.py: Tetra.gradient
The composition (models/cage_net.py:218-230, lib/cage.py:329,339-342) is written out below with those functions.
The delta / activation lines (models/cage_net.py:213-215, with the activations of :79-81) and CageBase.fem_energy
(lib/cage.py:348-361) sit inside classes whose modules cannot be imported here; their SOURCE LINES are read from the
reference files and executed as they stand, against a stand-in `self` that holds the tensors.

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

    # ---- the fused form: cage_net.py:213-215 executed from the reference's own source lines --------------------------
    import textwrap
    import types as _types

    import torch as th   # the reference's name for it  # noqa: F401

    def ref_lines(path, first, last):
        with open(os.path.join(REF, path)) as f:
            return textwrap.dedent("".join(f.readlines()[first - 1:last]))

    g = torch.Generator().manual_seed(5)
    N = barys.shape[0]
    scaling = torch.log(scales)                                   # the pre-activation parameter (cage_net.py:66,70)
    delta_bary = 0.05 * torch.randn(N, 4, generator=g)
    delta_scale = 0.3 * torch.randn(N, 3, generator=g)
    delta_rot = 0.2 * torch.randn(N, 4, generator=g)
    self = _types.SimpleNamespace(geometry=_types.SimpleNamespace(barys=barys), scaling=scaling, rotation=rot)
    ns = {"self": self, "th": torch, "delta_bary": delta_bary, "delta_scale": delta_scale, "delta_rot": delta_rot}
    exec(ref_lines("models/cage_net.py", 79, 81), ns)              # scaling_activation = exp, rotation_activation = normalize
    exec(ref_lines("models/cage_net.py", 213, 215), ns)            # canon_barys, scales, rotation
    f_barys, f_scales, f_rot = ns["canon_barys"], ns["scales"], ns["rotation"]
    f_L = gu.build_scaling_rotation(f_scales, f_rot)
    f_cov6 = gu.strip_symmetric(J @ (f_L @ f_L.transpose(1, 2)) @ J.transpose(1, 2))
    f_means3D = torch.einsum("ikj,ik->ij", deformed_tetras[tid], f_barys)

    # ---- CageBase.fem_energy (lib/cage.py:348-361), its own source, on the same cage --------------------------------
    ns2 = {"torch": torch}
    exec(ref_lines("lib/cage.py", 348, 361), ns2)
    Dn_inv = torch.linalg.inv(gradient(canon[tets]))               # lib/cage.py:312-313
    cage = _types.SimpleNamespace(gradient=gradient, tetras=tets)
    fm = ns2["fem_energy"](_types.SimpleNamespace(cage=cage, Dn_inv=Dn_inv), tetpoints)
finally:
    torch.zeros = _zeros

np.savez(os.path.join(ROOT, "tests", "golden", "deform_golden.npz"), tetpoints=sc.tetpoints, canon=sc.canon_points,
         tets=sc.tets, tetra_id=sc.tetra_id, barys=sc.barys, scales=sc.scales, rotation=rot.numpy(),
         canonical_gradient=canonical_gradient.numpy(), R=R.numpy(), L=L.numpy(), J=J.numpy(), cov6=cov6.numpy(),
         means3D=means3D.numpy(), scaling=scaling.numpy(), delta_bary=delta_bary.numpy(), delta_scale=delta_scale.numpy(),
         delta_rot=delta_rot.numpy(), fused_barys=f_barys.numpy(), fused_scales=f_scales.numpy(), fused_rotation=f_rot.numpy(),
         fused_cov6=f_cov6.numpy(), fused_means3D=f_means3D.numpy(), Dn_inv=Dn_inv.numpy(), fem_energy=fm.numpy())
print("wrote deform_golden.npz", means3D.shape, cov6.shape)
