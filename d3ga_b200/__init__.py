"""d3ga_b200 -- the render hot path of D3GA (cage deformation + Gaussian-splat rasterizer) as hand-written sm_100a
CUDA kernels behind the reference's own Python operator API.  See DESIGN.md and INTEGRATION.md."""
import sys


def install_dropin(lbs: bool = False) -> None:
    """Make ``import diff_gaussian_rasterization`` / ``import tetra_sampler`` (renderer.py:13-16, lib/cage.py:17)
    resolve to this package's implementations.  ``lbs=True`` also replaces ``lbsmodel`` / ``lbsmodel.body_model``
    (lib/blueman.py:9) by the fused cage-skinning module."""
    from d3ga_b200 import diff_gaussian_rasterization as dgr
    from d3ga_b200 import tetra_sampler as ts

    sys.modules["diff_gaussian_rasterization"] = dgr
    sys.modules["tetra_sampler"] = ts
    if lbs:
        from d3ga_b200 import lbsmodel

        sys.modules["lbsmodel"] = lbsmodel
        sys.modules["lbsmodel.body_model"] = lbsmodel.body_model
