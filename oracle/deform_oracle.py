"""oracle/deform_oracle.py -- TEST INFRASTRUCTURE ONLY (torch, CPU; autograd gives the backward).

Device-parametrised restatement of D3GA's per-frame cage -> Gaussian deformation.  The reference modules cannot be
imported in the authoring container (kornia / trimesh / pytorch3d / meshio absent) and hard-code ``device="cuda"``
(utils/general_utils.py:25,63,82), so each function below restates the cited lines op for op.  Pinned against the
reference's own functions where they *can* be executed: ``tests/golden/make_deform_golden.py`` imports
``utils/general_utils.py`` (build_rotation / build_scaling_rotation / strip_symmetric) and
``tetra_sampler/tetra.py`` (Tetra.gradient) from /root/reference with stub modules and stores their outputs in
``tests/golden/deform_golden.npz``; ``tests/test_oracle_deform.py`` checks this file against them.
"""
from __future__ import annotations

import torch as th


def tet_gradient(tetras: th.Tensor) -> th.Tensor:
    """submodules/tetra-sampler/tetra_sampler/tetra.py:74-80 -- edges (v3-v0, v2-v0, v1-v0) stacked on dim=2 (columns)."""
    v0, v1, v2, v3 = tetras[:, 0], tetras[:, 1], tetras[:, 2], tetras[:, 3]
    return th.stack([v3 - v0, v2 - v0, v1 - v0], dim=2)


def build_rotation(r: th.Tensor) -> th.Tensor:
    """utils/general_utils.py:58-79 -- re-normalises the quaternion (r,x,y,z) and builds R."""
    norm = th.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = th.zeros((q.size(0), 3, 3), dtype=r.dtype, device=r.device)
    r_, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r_ * z)
    R[:, 0, 2] = 2 * (x * z + r_ * y)
    R[:, 1, 0] = 2 * (x * y + r_ * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r_ * x)
    R[:, 2, 0] = 2 * (x * z - r_ * y)
    R[:, 2, 1] = 2 * (y * z + r_ * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def build_scaling_rotation(s: th.Tensor, r: th.Tensor) -> th.Tensor:
    """utils/general_utils.py:81-90 -- L = R @ diag(s)."""
    L = th.zeros((s.shape[0], 3, 3), dtype=s.dtype, device=s.device)
    R = build_rotation(r)
    L[:, 0, 0] = s[:, 0]
    L[:, 1, 1] = s[:, 1]
    L[:, 2, 2] = s[:, 2]
    return R @ L


def strip_symmetric(sym: th.Tensor) -> th.Tensor:
    """utils/general_utils.py:24-34,54-55 -- (xx, xy, xz, yy, yz, zz)."""
    return th.stack([sym[:, 0, 0], sym[:, 0, 1], sym[:, 0, 2], sym[:, 1, 1], sym[:, 1, 2], sym[:, 2, 2]], dim=1)


def canonical_gradient(canon_points: th.Tensor, tets: th.Tensor, tetra_id: th.Tensor) -> th.Tensor:
    """lib/cage.py:329 -- inv(G(canonical_tetras[tetra_id]))."""
    return th.linalg.inv(tet_gradient(canon_points[tets][tetra_id]))


def deform_gaussians(tetpoints, tets, tetra_id, canon_barys, canonical_gradient, scales, rotation):
    """models/cage_net.py:218-230 with lib/cage.py:339-342 and models/cage_net.py:161-164.

    ``scales`` is post-activation (cage_net.py:214) and ``rotation`` post-F.normalize (cage_net.py:215); the
    barycentrics are NOT renormalised (cage_net.py:213).  Returns (means3D [N,3], cov3D_precomp [N,6]).
    """
    deformed_tetras = tetpoints[tets]                                  # cage_net.py:218
    deformed = deformed_tetras[tetra_id]                               # cage.py:340 / cage_net.py:228
    J = tet_gradient(deformed) @ canonical_gradient                    # cage.py:340-341
    L = build_scaling_rotation(scales, rotation)                       # cage_net.py:162
    actual_covariance = L @ L.transpose(1, 2)                          # cage_net.py:163
    cov3D = J @ actual_covariance @ J.transpose(1, 2)                  # cage_net.py:222
    means3D = th.einsum("ikj,ik->ij", deformed, canon_barys)           # cage_net.py:230
    return means3D, strip_symmetric(cov3D)                             # cage_net.py:223


def deform_block(tetpoints, tets, tetra_id, barys, delta_bary, canonical_gradient, scaling, delta_scale, rotation,
                 delta_rot):
    """models/cage_net.py:213-215 in front of deform_gaussians: the canonical-field offsets and the activations
    (scaling_activation = exp, rotation_activation = F.normalize, cage_net.py:79-81)."""
    canon_barys = barys + delta_bary                                   # cage_net.py:213
    scales = th.exp(scaling + delta_scale)                             # cage_net.py:214
    rot = th.nn.functional.normalize(rotation + delta_rot)             # cage_net.py:215
    return deform_gaussians(tetpoints, tets, tetra_id, canon_barys, canonical_gradient, scales, rot)


def fem_energy(tetpoints, tets, Dn_inv):
    """lib/cage.py:349-361 -- stable neo-Hookean energy per tet (l = m = 0.5)."""
    Ds = tet_gradient(tetpoints[tets])
    F = Ds @ Dn_inv
    vol_loss = th.pow(th.linalg.det(F) - 1, 2)
    FtF = F.transpose(1, 2) @ F
    trace = th.einsum("ijj->i", FtF)
    return 0.5 * vol_loss + 0.5 * (trace - 3)
