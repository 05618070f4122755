"""Import surface of the reference's `tetra_sampler.lbs` that D3GA uses: `batch_rodrigues`
(submodules/tetra-sampler/tetra_sampler/lbs.py:272-302, imported at lib/smplman.py:16 and called at :167, :203).

Pure PyTorch on whatever device the input lives on -- it turns one global orientation per frame into a matrix, far from
the render hot path.  Written from the formula, pinned to the reference's own outputs (tests/golden/lbs_golden.npz).
"""
from __future__ import annotations

import torch


def batch_rodrigues(rot_vecs: torch.Tensor, epsilon: float = 1e-8, dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """Axis-angle vectors [N, 3] -> rotation matrices [N, 3, 3]:  R = I + sin(a) K + (1 - cos(a)) K^2.

    Like the reference, the angle is the norm of the vector with 1e-8 added to every component (so the zero vector
    gives the identity without a division by zero), and `epsilon` is accepted and not used."""
    v = rot_vecs
    angle = (v + 1e-8).norm(dim=1, keepdim=True)                 # [N, 1]
    x, y, z = (v / angle).unbind(dim=1)
    o = torch.zeros_like(x)
    K = torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=1).view(-1, 3, 3).to(dtype)
    s, c = torch.sin(angle).to(dtype)[:, :, None], torch.cos(angle).to(dtype)[:, :, None]
    eye = torch.eye(3, dtype=dtype, device=v.device)[None]
    return eye + s * K + (1 - c) * (K @ K)
