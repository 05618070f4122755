"""-m gpu test of d3ga_raster_backward_noalloc, the backward for hosts that cannot allocate device memory (a C/C++
frame loop without a caching allocator, or a process whose device memory was all handed out at start-up): its
gradient scratch is a fixed array inside the library instead of a caller buffer.

Called through the C-ABI twice in a row -- with different image gradients, so the second call would show stale sums
if the shared array were not cleared -- and compared with d3ga_raster_backward on the same forward state.
Collected last on purpose (file name), and self-contained so it also passes alone in a fresh process."""
import numpy as np
import pytest
import torch

import gpu_util
import parity_util as util
from d3ga_b200 import _lib, synth

pytestmark = pytest.mark.gpu
GRAD_TOL = 1e-3   # the relative-L2 bar of the other gradient tests (atomics reorder sums)
MAX_GAUSSIANS = 262144   # D3GA_NOALLOC_MAX_GAUSSIANS in include/d3ga_b200.h


def backward_noalloc(out, W, H, cam, dL_dcolor):
    lib = _lib.load()
    t = out["tensors"]
    P = out["P"]
    gcol = gpu_util.dev(dL_dcolor)
    nan = lambda *s: torch.full(s, float("nan"), dtype=torch.float32, device="cuda")  # noqa: E731
    res = dict(means2D=nan(P, 3), opacity=nan(P), colors=nan(P, 3), means3D=nan(P, 3), cov3D=nan(P, 6))
    rc = lib.d3ga_raster_backward_noalloc(
        P, 0, 0, W, H, t["bg"].data_ptr(), _lib.ptr(t["means"]), _lib.ptr(t["col"]), _lib.ptr(t["op"]), None, 1.0, None,
        _lib.ptr(t["cov"]), None, t["view"].data_ptr(), t["proj"].data_ptr(), t["campos"].data_ptr(), cam.tanfovx,
        cam.tanfovy, 0, t["radii"].data_ptr(), t["geom"].data_ptr(), t["binning"].data_ptr(), out["capacity"],
        t["image"].data_ptr(), gcol.data_ptr(), None, _lib.ptr(res["means2D"]), _lib.ptr(res["opacity"]),
        _lib.ptr(res["colors"]), _lib.ptr(res["means3D"]), _lib.ptr(res["cov3D"]), None, None, None,
        torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "raster_backward_noalloc")
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in res.items()}


def test_noalloc_backward_matches_regular_backward_twice(cuda_lib):
    sc = synth.make_workload("tiny")
    means, cov6 = util.scene_geometry(sc)
    g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
    cam = synth.make_camera(sc.W, sc.H, 25, 10)
    W, H = sc.W, sc.H
    fwd = gpu_util.forward_stages(g, cam, W, H, [0.2, 0.5, 0.8])
    assert fwd["num_rendered"] > 0
    rng = np.random.default_rng(12)
    for call in range(2):   # two calls in a row share the static array
        w = rng.uniform(-1, 1, size=(3, H, W)).astype(np.float32) * (1.0 + 3.0 * call)
        ref = gpu_util.backward(fwd, W, H, cam, w)
        got = backward_noalloc(fwd, W, H, cam, w)
        for k in ("means3D", "means2D", "opacity", "colors", "cov3D"):
            assert np.isfinite(got[k]).all(), (call, k)
            assert util.rel_l2(got[k], ref[k]) < GRAD_TOL, (call, k, util.rel_l2(got[k], ref[k]))


def test_noalloc_backward_rejects_sizes_beyond_its_scratch(cuda_lib):
    lib = _lib.load()
    rc = lib.d3ga_raster_backward_noalloc(
        MAX_GAUSSIANS + 1, 0, 0, 64, 64, *([None] * 5), 1.0, *([None] * 6), 1.0, 1.0, 0, None, None, None, 0,
        *([None] * 12))
    assert rc == -5
    assert b"fixed scratch" in lib.d3ga_error_string(rc)
    with pytest.raises(_lib.D3GAError, match="fixed scratch"):
        _lib.check(rc, "raster_backward_noalloc")
