"""-m gpu parity tests of the rasterizer, all through the C-ABI.

Bars (BASELINE.json): tile-key lists, sorted indices and per-tile ranges BIT-EXACT; images within 1e-4 max-abs (fp32);
gradients within 1e-3 relative L2.  The oracle is oracle/raster_oracle.c (parity unpinned by the reference: its
rasterizer source is absent, see the oracle's header)."""
import numpy as np
import pytest
import torch

import gpu_util
import parity_util as util
from d3ga_b200 import synth
from test_oracle_raster import kat_camera, kat_scene

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-4
GRAD_TOL = 1e-3


def cases():
    rng = np.random.default_rng(0)
    out = {}
    out["kat"] = (kat_scene(), kat_camera(), 64, 64)
    out["random"] = (util.random_gaussians(800, rng, sigma=0.06), synth.make_camera(96, 64, 10, 5), 96, 64)
    out["ragged_big"] = (util.random_gaussians(120, rng, sigma=0.3), synth.make_camera(50, 37, -20, 10), 50, 37)
    sc = synth.make_workload("tiny")
    m, c = util.scene_geometry(sc)
    out["tiny_avatar"] = (dict(means=m, cov6=c, opacity=sc.opacity, colors=sc.colors, scales=sc.scales,
                               rot=sc.rotation, shs=np.random.default_rng(1).normal(0, .3, (sc.N, 16, 3)).astype(np.float32)),
                          synth.make_camera(sc.W, sc.H, 25, 10), sc.W, sc.H)
    return out


CASES = cases()
BG = [0.2, 0.5, 0.8]


def assert_image_close(got, ref, what="image"):
    diff = np.abs(got.astype(np.float64) - ref.astype(np.float64))
    bad = int((diff > IMG_TOL).sum())
    # exp() differs by ulps between libm and the GPU's ex2-based __expf; a Gaussian sitting exactly on the 1/255 alpha
    # cut or the 1e-4 transmittance stop can flip for an isolated pixel.  None are tolerated on these seeded scenes.
    assert bad == 0, f"{what}: {bad} pixels beyond {IMG_TOL}, max {diff.max():.3e}"


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("path", ["precomp", "sh"])
def test_geometry_stage_bit_exact(cuda_lib, oracle_libs, name, path):
    g, cam, W, H = CASES[name]
    if path == "sh" and "shs" not in g:
        pytest.skip("no SH inputs")
    _, ref = util.oracle_forward(g, cam, W, H, BG, path)
    got = gpu_util.forward_stages(g, cam, W, H, BG, path, stop_after="geometry")
    assert got["num_rendered"] == ref["num_rendered"] and not got["overflow"]
    assert np.array_equal(got["radii"], ref["radii"])
    assert np.array_equal(got["tiles_touched"], ref["tiles_touched"])
    # emission order is an implementation choice (depth-first binning emits in depth order, upstream in Gaussian
    # order): the duplicate LIST must be the same set of (key, Gaussian) pairs, the prefix sums must end at D
    assert np.array_equal(got["depths"].view(np.uint32), ref["depths"].view(np.uint32))
    assert np.array_equal(got["xy"].view(np.uint32), ref["xy"].view(np.uint32))
    assert np.array_equal(got["conic_opacity"].view(np.uint32), ref["conic_opacity"].view(np.uint32))
    if not got["direct"]:
        # (the direct tile scatter only counts in stage 1: its duplicates are first written, sorted, by stage 2 --
        # test_sort_and_ranges_bit_exact compares that list)
        assert int(got["offsets"].max(initial=0)) == ref["num_rendered"]
        o_got = np.lexsort((got["vals_unsorted"], got["keys_unsorted"]))
        o_ref = np.lexsort((ref["vals_unsorted"], ref["keys_unsorted"]))
        assert np.array_equal(got["keys_unsorted"][o_got], ref["keys_unsorted"][o_ref])
        assert np.array_equal(got["vals_unsorted"][o_got], ref["vals_unsorted"][o_ref])
    if path == "sh":
        assert np.array_equal(got["cov3D"].view(np.uint32), ref["cov3D"].view(np.uint32))
        assert np.array_equal(got["clamped"], ref["clamped"])
        assert np.abs(got["rgb"] - ref["rgb"]).max() < 1e-5


@pytest.mark.parametrize("name", list(CASES))
def test_sort_and_ranges_bit_exact(cuda_lib, oracle_libs, name):
    g, cam, W, H = CASES[name]
    _, ref = util.oracle_forward(g, cam, W, H, BG)
    got = gpu_util.forward_stages(g, cam, W, H, BG)
    assert np.array_equal(got["keys"], ref["keys"])
    assert np.array_equal(got["point_list"], ref["point_list"])
    assert np.array_equal(got["ranges"], ref["ranges"])
    assert int(got["header"][5]) == ref["sort_bits"]


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("path,aa", [("precomp", False), ("sh", False), ("precomp", True)])
def test_image_within_tolerance(cuda_lib, oracle_libs, name, path, aa):
    g, cam, W, H = CASES[name]
    if path == "sh" and "shs" not in g:
        pytest.skip("no SH inputs")
    _, ref = util.oracle_forward(g, cam, W, H, BG, path, aa)
    got = gpu_util.forward_stages(g, cam, W, H, BG, path, aa)
    assert_image_close(got["color"], ref["color"])
    assert_image_close(got["invdepth"], ref["invdepth"], "invdepth")
    assert np.abs(got["final_T"] - ref["final_T"]).max() < IMG_TOL
    assert np.mean(got["n_contrib"] == ref["n_contrib"]) == 1.0


@pytest.mark.parametrize("name", ["random", "ragged_big", "tiny_avatar"])
@pytest.mark.parametrize("path,aa,inv", [("precomp", False, False), ("precomp", False, True), ("sh", False, False),
                                         ("sh", True, True)])
def test_backward_within_tolerance(cuda_lib, oracle_libs, name, path, aa, inv):
    g, cam, W, H = CASES[name]
    orc, ref = util.oracle_forward(g, cam, W, H, BG, path, aa)
    rng = np.random.default_rng(3)
    w = rng.uniform(-1, 1, size=(3, H, W)).astype(np.float32)
    wd = rng.uniform(-1, 1, size=(H, W)).astype(np.float32) if inv else None
    rg = orc.backward(ref, w, wd)
    got = gpu_util.forward_stages(g, cam, W, H, BG, path, aa)
    gg = gpu_util.backward(got, W, H, cam, w, wd, aa)
    names = ["means3D", "means2D", "opacity"] + (["colors", "cov3D"] if path == "precomp" else ["sh", "scales", "rotations"])
    for k in names:
        r = rg["mean2D" if k == "means2D" else k]
        assert np.isfinite(gg[k]).all(), k
        assert util.rel_l2(gg[k], r) < GRAD_TOL, (k, util.rel_l2(gg[k], r))


def test_all_culled_and_empty(cuda_lib, oracle_libs):
    cam = kat_camera()
    g = kat_scene()
    g["means"][:, 2] = -1.0   # everything behind the camera
    got = gpu_util.forward_stages(g, cam, 64, 64, BG)
    assert got["num_rendered"] == 0 and np.all(got["radii"] == 0) and np.all(got["ranges"] == 0)
    assert np.allclose(got["color"], np.asarray(BG, np.float32)[:, None, None])
    gg = gpu_util.backward(got, 64, 64, cam, np.ones((3, 64, 64), np.float32))
    assert all(np.all(v == 0) for v in gg.values() if v is not None)
    empty = {k: v[:0] for k, v in g.items()}
    got = gpu_util.forward_stages(empty, cam, 64, 64, BG)
    assert got["num_rendered"] == 0
    assert np.allclose(got["color"], np.asarray(BG, np.float32)[:, None, None])


def test_capacity_overflow_is_reported_and_retried(cuda_lib, oracle_libs):
    g, cam, W, H = CASES["random"]
    _, ref = util.oracle_forward(g, cam, W, H, BG)
    got = gpu_util.forward_stages(g, cam, W, H, BG, capacity=ref["num_rendered"] // 2)
    assert got["overflow"] == 1 and got["num_rendered"] == ref["num_rendered"]
    # the Python drop-in grows the workspace and re-runs both stages
    from d3ga_b200 import diff_gaussian_rasterization as dgr

    dgr._capacity_hint[(torch.device("cuda", 0), g["means"].shape[0], W, H)] = 64
    img = render_api(g, cam, W, H)[0]
    assert_image_close(img.detach().cpu().numpy(), ref["color"])


def render_api(g, cam, W, H, path="precomp", grads=False, device="cuda"):
    from d3ga_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

    d = lambda a: gpu_util.dev(a).to(device)  # noqa: E731
    rs = GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=d(np.asarray(BG, np.float32)),
        scale_modifier=1.0, viewmatrix=d(cam.viewmatrix), projmatrix=d(cam.projmatrix), sh_degree=3,
        campos=d(cam.campos), prefiltered=False, debug=False, antialiasing=False)
    t = {k: d(v).requires_grad_(grads) for k, v in g.items() if k in ("means", "opacity", "colors", "cov6", "shs", "scales", "rot")}
    means2D = (torch.zeros_like(t["means"], requires_grad=True) + 0)   # renderer.py:122-128
    if grads:
        means2D.retain_grad()
    kw = dict(colors_precomp=t["colors"], cov3D_precomp=t["cov6"]) if path == "precomp" else dict(
        shs=t["shs"], scales=t["scales"], rotations=t["rot"])
    out = GaussianRasterizer(raster_settings=rs)(means3D=t["means"], means2D=means2D, opacities=t["opacity"], **kw)
    return out[0], out, t, means2D


@pytest.mark.parametrize("path", ["precomp", "sh"])
def test_python_dropin_autograd(cuda_lib, oracle_libs, path):
    """The call renderer.py:130-141 makes, element [0] used as the image, gradients through autograd."""
    g, cam, W, H = CASES["tiny_avatar"]
    orc, ref = util.oracle_forward(g, cam, W, H, BG, path)
    img, out, t, means2D = render_api(g, cam, W, H, path, grads=True)
    assert len(out) == 3 and out[1].dtype == torch.int32 and out[2].shape == (1, H, W)
    assert_image_close(img.detach().cpu().numpy(), ref["color"])
    assert np.array_equal(out[1].cpu().numpy(), ref["radii"])
    w = torch.from_numpy(np.random.default_rng(4).uniform(-1, 1, (3, H, W)).astype(np.float32)).cuda()
    (img * w).sum().backward()
    rg = orc.backward(ref, w.cpu().numpy())
    pairs = [("means", "means3D"), ("opacity", "opacity")] + (
        [("colors", "colors"), ("cov6", "cov3D")] if path == "precomp" else [("shs", "sh"), ("scales", "scales"), ("rot", "rotations")])
    for a, b in pairs:
        assert util.rel_l2(t[a].grad.cpu().numpy(), rg[b]) < GRAD_TOL, a
    assert util.rel_l2(means2D.grad.cpu().numpy(), rg["mean2D"]) < GRAD_TOL


def report_parity(record):
    """Appends one JSON line per measured parity figure to $D3GA_PARITY_LOG (DESIGN.md's tables are made from it)."""
    import json
    import os

    path = os.environ.get("D3GA_PARITY_LOG")
    if path:
        with open(path, "a") as f:
            f.write(json.dumps(record) + "\n")
    print("[parity]", json.dumps(record))


@pytest.mark.parametrize("workload", ["c1", "c2", "c3", "metric"])
def test_full_size_properties(cuda_lib, oracle_libs, workload):
    """BASELINE.json sizes: oracle comparison where it finishes in seconds plus size-independent properties."""
    sc = synth.make_workload(workload)
    means, cov6 = util.scene_geometry(sc)
    g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
    cam = synth.make_camera(sc.W, sc.H, 20, 10)
    W, H = sc.W, sc.H
    got = gpu_util.forward_stages(g, cam, W, H, [1, 1, 1])
    D = got["num_rendered"]
    keys, pl, ranges = got["keys"], got["point_list"], got["ranges"].astype(np.int64)
    bits = int(got["header"][5])
    masked = keys & np.uint64((1 << bits) - 1)
    assert np.all(masked[1:] >= masked[:-1])                                   # sortedness
    assert np.array_equal(np.bincount(pl, minlength=sc.N), got["tiles_touched"])   # every duplicate survived the sort
    lens = ranges[:, 1] - ranges[:, 0]
    assert lens.sum() == D and np.all(lens >= 0)                               # ranges partition the list
    nz = np.nonzero(lens)[0]
    assert np.array_equal((keys[ranges[nz, 0]] >> np.uint64(32)).astype(np.int64), nz)
    # stability: within equal keys the Gaussian index increases
    same = keys[1:] == keys[:-1]
    assert np.all(pl[1:][same] > pl[:-1][same])
    # idempotence
    again = gpu_util.forward_stages(g, cam, W, H, [1, 1, 1])
    assert np.array_equal(again["color"], got["color"]) and np.array_equal(again["keys"], keys)
    # oracle comparison at full size (bit-exact lists, image tolerance with the documented flip allowance)
    orc, ref = util.oracle_forward(g, cam, W, H, [1, 1, 1])
    assert np.array_equal(keys, ref["keys"]) and np.array_equal(pl, ref["point_list"])
    assert np.array_equal(got["ranges"], ref["ranges"])
    # The image bar is 1e-4 max-abs.  A (pixel, Gaussian) pair whose alpha sits on the 1/255 cut is decided in upstream's
    # own arithmetic (raster_blend.cu: exact_cut), so the contributor sets agree; what is left is the transmittance stop
    # T (1 - alpha) < 1e-4, whose T carries the rounding of every alpha before it and cannot be decided "exactly" without
    # making every alpha bit-identical.  Measured and asserted: the number of values beyond the bar and their maximum;
    # every one of them must be a pixel whose contributor count differs from the oracle's (a flipped stop), nothing else.
    diff = np.abs(got["color"].astype(np.float64) - ref["color"].astype(np.float64))
    off = diff > IMG_TOL
    n_off, mx = int(off.sum()), float(diff.max())
    nc_diff = got["n_contrib"] != ref["n_contrib"]
    px_off = off.any(axis=0).reshape(-1)
    report_parity(dict(test="full_size_image", workload=workload, values=int(diff.size), beyond_1e_4=n_off, max_abs=mx,
                       pixels_with_other_contributor_count=int(nc_diff.sum()),
                       offenders_explained_by_contributor_count=int((px_off & nc_diff).sum()), offenders=int(px_off.sum())))
    assert not (px_off & ~nc_diff).any(), "a pixel beyond 1e-4 with the oracle's own contributor set"
    assert px_off.sum() <= 4 and mx <= 1e-2, (n_off, mx)
    # backward: linearity in dL/dimage and oracle agreement
    rng = np.random.default_rng(9)
    w = rng.uniform(-1, 1, size=(3, H, W)).astype(np.float32)
    g1 = gpu_util.backward(got, W, H, cam, w)
    g2 = gpu_util.backward(got, W, H, cam, 2 * w)
    assert util.rel_l2(g2["means3D"], 2 * g1["means3D"]) < 1e-5
    rg = orc.backward(ref, w)
    errs = {k: util.rel_l2(g1[k], rg[k]) for k in ("means3D", "cov3D", "opacity", "colors")}
    report_parity(dict(test="full_size_gradients", workload=workload, **errs))
    for k, e in errs.items():
        assert e < GRAD_TOL, (k, e)
        # Regression guard, far inside the bar: fed the oracle's own inputs the kernels agree with it to ~1e-6 (the lists
        # are bit-identical, so nothing discontinuous separates the two).  A bookkeeping bug in the backward's staging once
        # showed as 3e-3 on one camera of the pipelined test while this comparison still passed the 1e-3 bar.
        assert e < 5e-5, (k, e)


def test_mark_visible(cuda_lib, oracle_libs):
    """GaussianRasterizer.markVisible: z_view > 0.2 with the preprocess's arithmetic -- every Gaussian the forward gives
    a radius is marked, and the mask equals the float64 test away from the threshold."""
    from d3ga_b200 import diff_gaussian_rasterization as dgr

    g, cam, W, H = CASES["random"]
    _, ref = util.oracle_forward(g, cam, W, H, BG)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    rs = dgr.GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=torch.zeros(3, device="cuda"),
        scale_modifier=1.0, viewmatrix=t(cam.viewmatrix), projmatrix=t(cam.projmatrix), sh_degree=0, campos=t(cam.campos),
        prefiltered=False, debug=False, antialiasing=False)
    got = dgr.GaussianRasterizer(raster_settings=rs).markVisible(t(g["means"])).cpu().numpy()
    assert got.dtype == np.bool_ and got.shape == (g["means"].shape[0],)
    z = g["means"].astype(np.float64) @ cam.viewmatrix.astype(np.float64)[:3, 2] + float(cam.viewmatrix[3, 2])
    clear = np.abs(z - 0.2) > 1e-5
    assert np.array_equal(got[clear], (z > 0.2)[clear]) and 0 < got.sum() < got.size
    assert got[ref["radii"] > 0].all()
    assert dgr.GaussianRasterizer(raster_settings=rs).markVisible(torch.zeros((0, 3), device="cuda")).numel() == 0


def test_views_in_flight_hint_changes_nothing_but_the_row_length(cuda_lib, oracle_libs):
    """d3ga_set_views_in_flight(n) only re-sizes the direct tile scatter's rows (a view then takes ~1/n of the GPU): lists,
    ranges and image must stay bit-identical.  Checked on the 300 k-Gaussian 1024^2 scene, where n = 1, 4 and 64 give rows of
    a few hundred, a thousand and 8192 Gaussians (several, one and almost no rows per SM)."""
    sc = synth.make_workload("c2")
    means, cov6 = util.scene_geometry(sc)
    g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
    cam = synth.make_camera(sc.W, sc.H, 20, 10)
    lib = cuda_lib
    outs = []
    try:
        for n in (1, 4, 64):
            lib.d3ga_set_views_in_flight(n)
            outs.append(gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1]))
    finally:
        lib.d3ga_set_views_in_flight(1)
    a = outs[0]
    if not a["direct"]:
        pytest.skip("D3GA_SORT_MODE selects another binning path: the hint only sizes the direct tile scatter")
    for b in outs[1:]:
        assert np.array_equal(a["point_list"], b["point_list"]) and np.array_equal(a["keys"], b["keys"])
        assert np.array_equal(a["ranges"], b["ranges"]) and np.array_equal(a["color"], b["color"])
        assert a["num_rendered"] == b["num_rendered"]


def test_hint_flipped_between_the_stages_changes_nothing(cuda_lib, oracle_libs):
    """Stage 1 files the row plan it chose with the workspace and stage 2 takes it from there: a views-in-flight hint that
    changes BETWEEN the two stages of a view (another thread, a capture in progress) must not touch the lists.  Before,
    both stages derived the plan from the process-wide hint and a flip silently corrupted the sorted list."""
    sc = synth.make_workload("c1")
    means, cov6 = util.scene_geometry(sc)
    g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
    cam = synth.make_camera(sc.W, sc.H, 20, 10)
    lib = cuda_lib
    ref = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1])
    if not ref["direct"]:
        pytest.skip("D3GA_SORT_MODE selects another binning path")
    try:
        for first, second in ((1, 16), (16, 1), (4, 64)):
            lib.d3ga_set_views_in_flight(first)
            got = gpu_util.forward_stages(g, cam, sc.W, sc.H, [1, 1, 1], between_stages=lambda: lib.d3ga_set_views_in_flight(second))
            assert np.array_equal(ref["point_list"], got["point_list"]) and np.array_equal(ref["keys"], got["keys"])
            assert np.array_equal(ref["ranges"], got["ranges"]) and np.array_equal(ref["color"], got["color"])
    finally:
        lib.d3ga_set_views_in_flight(1)
    # stage 2 on a workspace no stage 1 has run on is refused, not guessed
    # (a buffer at an address no geometry workspace can have had: inside a larger allocation, off its 512-byte grid)
    t = ref["tensors"]
    big = torch.zeros(t["geom"].numel() + 1024, dtype=torch.uint8, device="cuda")
    geom2 = big[256:256 + t["geom"].numel()]
    geom2.copy_(t["geom"])
    out = torch.zeros((3, sc.H, sc.W), device="cuda")
    rc = lib.d3ga_raster_forward_render(ref["P"], sc.W, sc.H, t["bg"].data_ptr(), t["col"].data_ptr(), geom2.data_ptr(),
                                        t["binning"].data_ptr(), ref["capacity"], t["image"].data_ptr(), out.data_ptr(), None,
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == -4   # D3GA_ENOTREADY


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_second_device_in_one_process(cuda_lib, oracle_libs):
    """Function attributes (the dynamic shared-memory opt-ins of the blend backward, the sorts and the scatter) are
    per device: a view on cuda:1 after one on cuda:0 in the same process must launch and match the oracle."""
    g, cam, W, H = CASES["tiny_avatar"]
    orc, ref = util.oracle_forward(g, cam, W, H, BG)
    w = np.random.default_rng(4).uniform(-1, 1, (3, H, W)).astype(np.float32)
    rg = orc.backward(ref, w)
    for d in (0, 1, 0):
        with torch.cuda.device(d):
            img, out, t, _ = render_api(g, cam, W, H, grads=True, device=f"cuda:{d}")
            (img * torch.from_numpy(w).to(img.device)).sum().backward()
            torch.cuda.synchronize()
            assert_image_close(img.detach().cpu().numpy(), ref["color"])
            for a, b in (("means", "means3D"), ("opacity", "opacity"), ("colors", "colors"), ("cov6", "cov3D")):
                assert util.rel_l2(t[a].grad.cpu().numpy(), rg[b]) < GRAD_TOL, (d, a)


def test_other_binning_modes_are_bit_exact_too(cuda_lib, oracle_libs):
    """The two other binning strategies (D3GA_SORT_MODE=depthfirst: the fallback for tile grids the scatter cannot take --
    depth sort of the Gaussians, emission in depth order, radix passes on the tile digits; =global: one 64-bit LSD sort of
    the duplicates in Gaussian order, upstream's own sequence, kept as a cross-check) are chosen when the library first
    sorts, so each is exercised in a fresh process; both must give the same lists as the default."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for mode in ("depthfirst", "global"):
        env = dict(os.environ, D3GA_SORT_MODE=mode)
        r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu",
                            os.path.join(root, "tests", "test_gpu_raster.py"), "-k",
                            "test_sort_and_ranges_bit_exact or test_geometry_stage_bit_exact or "
                            "(test_full_size_properties and c2)"],
                           env=env, cwd=root, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, mode + "\n" + r.stdout[-2000:] + r.stderr[-2000:]


def test_c4_stress_billion_duplicates(cuda_lib):
    """BASELINE.json config 4 (4 M Gaussians, scales x4, 3840x2160): ~1.3e9 (tile, Gaussian) duplicates -- beyond 2^30,
    far beyond what the CPU oracle can check in seconds, so only size-independent properties are verified, on the
    device: sortedness, ranges partition the list, every Gaussian appears tiles_touched times, capacity retry path."""
    from d3ga_b200 import diff_gaussian_rasterization as dgr
    from d3ga_b200.tetra_sampler import deform_gaussians

    sc = synth.make_workload("c4")
    cam = synth.make_camera(sc.W, sc.H, 20, 10)
    t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    means, cov6 = deform_gaussians(t(sc.tetpoints), t(sc.tets), t(sc.tetra_id), t(sc.barys), t(sc.canonical_gradient),
                                   t(sc.scales), t(sc.rotation))
    rs = dgr.GaussianRasterizationSettings(
        image_height=sc.H, image_width=sc.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=torch.ones(3, device="cuda"),
        scale_modifier=1.0, viewmatrix=t(cam.viewmatrix), projmatrix=t(cam.projmatrix), sh_degree=0, campos=t(cam.campos),
        prefiltered=False, debug=False, antialiasing=False)
    color, radii, _, st = dgr.rasterize_forward(means, None, t(sc.colors), t(sc.opacity), None, None, cov6, rs,
                                                want_invdepth=False)
    torch.cuda.synchronize()
    D = st.num_rendered
    assert D > (1 << 30) and torch.isfinite(color).all()
    lay = dgr._lib.raster_layout(sc.N, sc.W, sc.H, st.capacity)
    hdr = st.geom[:256].cpu().numpy().view(np.uint32)
    assert int(hdr[0]) == D and int(hdr[1]) == 0
    sel = int(hdr[3])
    voff = lay["bin_vals1" if sel else "bin_vals0"]
    keys = torch.zeros(D, dtype=torch.int64, device="cuda")
    dgr._lib.check(dgr._lib.load().d3ga_raster_sorted_keys(
        sc.N, sc.W, sc.H, st.geom.data_ptr(), st.binning.data_ptr(), st.capacity, st.image.data_ptr(), keys.data_ptr(),
        torch.cuda.current_stream().cuda_stream), "sorted_keys")
    tiles_of_key = keys >> 32
    assert bool((tiles_of_key[1:] >= tiles_of_key[:-1]).all())
    depth_of_key = keys & 0xFFFFFFFF
    same_tile = tiles_of_key[1:] == tiles_of_key[:-1]
    assert bool((depth_of_key[1:][same_tile] >= depth_of_key[:-1][same_tile]).all())
    del same_tile, depth_of_key
    vals = st.binning[voff:voff + 4 * D].view(torch.int32)
    tiles = st.geom[lay["geom_tiles"]:lay["geom_tiles"] + 4 * sc.N].view(torch.int32)
    assert int(tiles.sum(dtype=torch.int64)) == D
    assert bool((torch.bincount(vals, minlength=sc.N) == tiles).all())
    ntiles = ((sc.W + 15) // 16) * ((sc.H + 15) // 16)
    ranges = st.image[lay["img_ranges"]:lay["img_ranges"] + 8 * ntiles].view(torch.int32).view(ntiles, 2).long() & 0xFFFFFFFF
    lens = ranges[:, 1] - ranges[:, 0]
    assert int(lens.sum()) == D and bool((lens >= 0).all())
    nz = torch.nonzero(lens).flatten()
    assert bool((tiles_of_key[ranges[nz, 0]] == nz).all())


def edge_cases():
    rng = np.random.default_rng(21)
    out = {}
    # opacity 1 at pixel centres: alpha hits the 0.99 clamp (gradient flows through it like upstream)
    g = util.random_gaussians(64, rng, sigma=0.08, with_behind=False)
    g["opacity"][:] = 1.0
    out["alpha_clamp"] = (g, synth.make_camera(48, 48, 0, 0), 48, 48)
    # one Gaussian far larger than the image, one sub-pixel, one exactly on a tile corner
    g = util.random_gaussians(3, rng, with_behind=False)
    g["means"][:] = [[0, 0, 0], [0.05, 0.02, 0], [0, 0, 0]]
    g["cov6"][:] = 0
    g["cov6"][0, [0, 3, 5]] = 4.0
    g["cov6"][1, [0, 3, 5]] = 1e-10
    g["cov6"][2, [0, 3, 5]] = 1e-3
    g["scales"][:] = [[2, 2, 2], [1e-5, 1e-5, 1e-5], [0.03, 0.03, 0.03]]
    out["huge_and_tiny"] = (g, synth.make_camera(64, 64, 0, 0), 64, 64)
    # indefinite / zero covariance inputs, a point far off screen and one behind the camera
    g = util.random_gaussians(16, rng, with_behind=True)
    g["cov6"][0] = 0
    g["cov6"][1] = [1e-3, 2e-3, 0, 1e-3, 0, 1e-3]   # not positive definite
    g["means"][2] = [50, 0, 0]
    out["degenerate"] = (g, synth.make_camera(40, 24, 5, 5), 40, 24)
    for (w, h) in [(1, 1), (16, 16), (17, 33)]:
        out[f"size_{w}x{h}"] = (util.random_gaussians(257, rng, sigma=0.2, with_behind=False), synth.make_camera(w, h, 10, -10), w, h)
    out["single"] = (util.random_gaussians(1, rng, sigma=0.1, with_behind=False), synth.make_camera(32, 32), 32, 32)
    return out


EDGE = edge_cases()


@pytest.mark.parametrize("name", list(EDGE))
@pytest.mark.parametrize("path", ["precomp", "sh"])
def test_edge_cases_match_oracle(cuda_lib, oracle_libs, name, path):
    g, cam, W, H = EDGE[name]
    orc, ref = util.oracle_forward(g, cam, W, H, BG, path)
    got = gpu_util.forward_stages(g, cam, W, H, BG, path)
    assert np.array_equal(got["radii"], ref["radii"])
    assert np.array_equal(got["keys"], ref["keys"]) and np.array_equal(got["point_list"], ref["point_list"])
    assert np.array_equal(got["ranges"], ref["ranges"])
    assert_image_close(got["color"], ref["color"])
    assert np.array_equal(got["n_contrib"], ref["n_contrib"])
    w = np.random.default_rng(5).uniform(-1, 1, size=(3, H, W)).astype(np.float32)
    rg = orc.backward(ref, w)
    gg = gpu_util.backward(got, W, H, cam, w)
    names = ["means3D", "means2D", "opacity"] + (["colors", "cov3D"] if path == "precomp" else ["sh", "scales", "rotations"])
    for k in names:
        r = rg["mean2D" if k == "means2D" else k]
        assert np.isfinite(gg[k]).all(), k
        if np.linalg.norm(r) > 0:
            assert util.rel_l2(gg[k], r) < GRAD_TOL, (k, util.rel_l2(gg[k], r))
        else:
            assert np.abs(gg[k]).max() == 0, k


@pytest.mark.parametrize("deg,M", [(0, 1), (1, 4), (2, 9), (2, 16), (3, 16)])
def test_sh_degrees(cuda_lib, oracle_libs, deg, M):
    """Active SH degree below the stored coefficient count (D3GA raises sh_degree during training)."""
    from d3ga_b200 import _lib

    rng = np.random.default_rng(deg * 10 + M)
    g = util.random_gaussians(300, rng, sigma=0.06, with_behind=False)
    g["shs"] = np.ascontiguousarray(g["shs"][:, :M])
    cam = synth.make_camera(64, 48, 15, 5)
    orc, ref = util.oracle_forward(g, cam, 64, 48, BG, "sh", sh_degree=deg)
    lib = _lib.load()
    d = gpu_util.dev
    t = dict(means=d(g["means"]), op=d(g["opacity"]), sh=d(g["shs"]), sc=d(g["scales"]), rot=d(g["rot"]))
    P, W, H = 300, 64, 48
    cap = 1 << 16
    gb, bb, ib = _lib.workspace_bytes(P, W, H, cap)
    geom, binning, image = (torch.zeros(n, dtype=torch.uint8, device="cuda") for n in (gb, bb, ib))
    radii = torch.zeros(P, dtype=torch.int32, device="cuda")
    color = torch.zeros(3, H, W, device="cuda")
    view, proj, campos, bg = d(cam.viewmatrix), d(cam.projmatrix), d(cam.campos), d(np.asarray(BG, np.float32))
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.d3ga_raster_forward_geometry(P, deg, M, t["means"].data_ptr(), t["sc"].data_ptr(), 1.0, t["rot"].data_ptr(),
                                                t["op"].data_ptr(), t["sh"].data_ptr(), None, None, view.data_ptr(),
                                                proj.data_ptr(), campos.data_ptr(), W, H, cam.tanfovx, cam.tanfovy, 0, 0,
                                                geom.data_ptr(), binning.data_ptr(), cap, radii.data_ptr(), None, s), "geometry")
    _lib.check(lib.d3ga_raster_forward_render(P, W, H, bg.data_ptr(), None, geom.data_ptr(), binning.data_ptr(), cap,
                                              image.data_ptr(), color.data_ptr(), None, s), "render")
    torch.cuda.synchronize()
    assert_image_close(color.cpu().numpy(), ref["color"])
    w = rng.uniform(-1, 1, size=(3, H, W)).astype(np.float32)
    rg = orc.backward(ref, w)
    scratch = torch.empty(int(lib.d3ga_raster_backward_scratch_bytes(P)), dtype=torch.uint8, device="cuda")
    z = lambda *sh: torch.full(sh, float("nan"), device="cuda")  # noqa: E731
    gm2, gop, gm3, gsh, gsc, grot = z(P, 3), z(P), z(P, 3), z(P, M, 3), z(P, 3), z(P, 4)
    _lib.check(lib.d3ga_raster_backward(P, deg, M, W, H, bg.data_ptr(), t["means"].data_ptr(), None, t["op"].data_ptr(),
                                        t["sc"].data_ptr(), 1.0, t["rot"].data_ptr(), None, t["sh"].data_ptr(),
                                        view.data_ptr(), proj.data_ptr(), campos.data_ptr(), cam.tanfovx, cam.tanfovy, 0,
                                        radii.data_ptr(), geom.data_ptr(), binning.data_ptr(), cap, image.data_ptr(),
                                        d(w).data_ptr(), None, scratch.data_ptr(), gm2.data_ptr(), gop.data_ptr(), None,
                                        gm3.data_ptr(), None, gsh.data_ptr(), gsc.data_ptr(), grot.data_ptr(), s), "backward")
    torch.cuda.synchronize()
    assert util.rel_l2(gsh.cpu().numpy(), rg["sh"]) < GRAD_TOL
    assert util.rel_l2(gm3.cpu().numpy(), rg["means3D"]) < GRAD_TOL
    assert np.all(gsh.cpu().numpy()[:, (deg + 1) ** 2:] == 0)


def test_shared_geometry_second_render(cuda_lib, oracle_libs):
    """models/trainer.py:102-110: the RGB render and the silhouette render of a frame pass the SAME means / covariance /
    opacity tensors and an equal camera; the second call must re-use the first call's binning and still give exactly
    the image and gradients of an independent render."""
    from d3ga_b200 import diff_gaussian_rasterization as dgr
    from d3ga_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

    g, cam, W, H = CASES["tiny_avatar"]
    d = gpu_util.dev

    def settings(bg):   # a fresh Camera per render() call, like renderer.py:70
        return GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=d(np.asarray(bg, np.float32)),
            scale_modifier=1.0, viewmatrix=d(cam.viewmatrix), projmatrix=d(cam.projmatrix), sh_degree=0,
            campos=d(cam.campos), prefiltered=False, debug=False, antialiasing=False)

    means = d(g["means"]).requires_grad_(True)
    cov = d(g["cov6"]).requires_grad_(True)
    op = d(g["opacity"]).requires_grad_(True)
    rgb = d(g["colors"]).requires_grad_(True)
    sil = torch.ones_like(rgb).requires_grad_(True)
    dgr._geom_cache.clear()
    h0, m0 = dgr.share_stats["hits"], dgr.share_stats["misses"]
    img1 = GaussianRasterizer(settings([0.3, 0.6, 0.9]))(means3D=means, means2D=torch.zeros_like(means), opacities=op,
                                                        colors_precomp=rgb, cov3D_precomp=cov)[0]
    img2 = GaussianRasterizer(settings([0, 0, 0]))(means3D=means, means2D=torch.zeros_like(means), opacities=op,
                                                  colors_precomp=sil, cov3D_precomp=cov)[0]
    assert dgr.share_stats["hits"] == h0 + 1 and dgr.share_stats["misses"] == m0 + 1
    rng = np.random.default_rng(8)
    w1, w2 = (d(rng.uniform(-1, 1, (3, H, W)).astype(np.float32)) for _ in range(2))
    ((img1 * w1).sum() + (img2 * w2).sum()).backward()
    # oracle: two independent renders
    g2 = dict(g, colors=np.ones_like(g["colors"]))
    o1, r1 = util.oracle_forward(g, cam, W, H, [0.3, 0.6, 0.9])
    o2, r2 = util.oracle_forward(g2, cam, W, H, [0, 0, 0])
    assert_image_close(img1.detach().cpu().numpy(), r1["color"])
    assert_image_close(img2.detach().cpu().numpy(), r2["color"])
    ga, gb = o1.backward(r1, w1.cpu().numpy()), o2.backward(r2, w2.cpu().numpy())
    assert util.rel_l2(means.grad.cpu().numpy(), ga["means3D"] + gb["means3D"]) < GRAD_TOL
    assert util.rel_l2(cov.grad.cpu().numpy(), ga["cov3D"] + gb["cov3D"]) < GRAD_TOL
    assert util.rel_l2(op.grad.cpu().numpy().ravel(), ga["opacity"] + gb["opacity"]) < GRAD_TOL
    assert util.rel_l2(rgb.grad.cpu().numpy(), ga["colors"]) < GRAD_TOL
    assert util.rel_l2(sil.grad.cpu().numpy(), gb["colors"]) < GRAD_TOL
    # an in-place update of the geometry (optimizer step) or another camera must miss
    with torch.no_grad():
        means.add_(0.0)
    h1 = dgr.share_stats["hits"]
    GaussianRasterizer(settings([0, 0, 0]))(means3D=means, means2D=torch.zeros_like(means), opacities=op,
                                           colors_precomp=sil, cov3D_precomp=cov)
    assert dgr.share_stats["hits"] == h1
    cam2 = synth.make_camera(W, H, 80, 0)
    rs2 = settings([0, 0, 0])._replace(viewmatrix=d(cam2.viewmatrix), projmatrix=d(cam2.projmatrix), campos=d(cam2.campos))
    img3 = GaussianRasterizer(rs2)(means3D=means, means2D=torch.zeros_like(means), opacities=op, colors_precomp=sil,
                                   cov3D_precomp=cov)[0]
    assert dgr.share_stats["hits"] == h1
    _, r3 = util.oracle_forward(g2, cam2, W, H, [0, 0, 0])
    assert_image_close(img3.detach().cpu().numpy(), r3["color"])


@pytest.mark.parametrize("case,layout", [("tiny_avatar", "cov"), ("tiny_avatar", "scale_rot"), ("c1", "cov")])
def test_pair_render_is_the_two_renders_in_one_pass(cuda_lib, oracle_libs, case, layout):
    """SURVEY.md 8(f) rank 1, the fused form: GaussianRasterizer.forward_pair gives the RGB image and the silhouette image of
    models/trainer.py:102-110 from ONE binning, ONE blend forward and ONE blend backward.  Both images must be bit-identical
    to two separate renders (same alphas, same transmittance, same contributor set), the geometry / opacity gradients must be
    the SUM of the two renders' (checked against the oracle's two independent views), the first image's colours get their
    own gradient and the second image's colours none."""
    from d3ga_b200 import diff_gaussian_rasterization as dgr
    from d3ga_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

    if case == "c1":
        sc = synth.make_workload("c1")
        means_np, cov_np = util.scene_geometry(sc)
        g = dict(means=means_np, cov6=cov_np, opacity=sc.opacity, colors=sc.colors)
        cam, W, H = synth.make_camera(sc.W, sc.H, 20, 10), sc.W, sc.H
    else:
        g, cam, W, H = CASES["tiny_avatar"]
    if layout == "scale_rot" and "scales" not in g:
        pytest.skip("case has no scale / rotation form")
    d = gpu_util.dev
    rng = np.random.default_rng(12)
    P = g["means"].shape[0]
    # piecewise-constant second colours, like the per-garment mask colours of models/cage_net.py:243
    sil_np = np.where((np.arange(P) % 3 == 0)[:, None], [1.0, 0.0, 0.5], [0.2, 1.0, 0.0]).astype(np.float32)
    bg_a, bg_b = [0.3, 0.6, 0.9], [0.0, 0.1, 0.0]

    def settings(bg):
        return GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=d(np.asarray(bg, np.float32)),
            scale_modifier=1.0, viewmatrix=d(cam.viewmatrix), projmatrix=d(cam.projmatrix), sh_degree=0,
            campos=d(cam.campos), prefiltered=False, debug=False, antialiasing=False)

    def leaves():
        out = dict(means=d(g["means"]).requires_grad_(True), op=d(g["opacity"]).requires_grad_(True),
                   rgb=d(g["colors"]).requires_grad_(True))
        if layout == "cov":
            out["cov"] = d(g["cov6"]).requires_grad_(True)
            out["geo"] = dict(cov3D_precomp=out["cov"])
        else:
            out["sc"], out["rot"] = d(g["scales"]).requires_grad_(True), d(g["rot"]).requires_grad_(True)
            out["geo"] = dict(scales=out["sc"], rotations=out["rot"])
        return out

    w1, w2 = (d(rng.uniform(-1, 1, (3, H, W)).astype(np.float32)) for _ in range(2))
    sil = d(sil_np)
    # two separate renders
    a = leaves()
    dgr._geom_cache.clear()
    i1 = GaussianRasterizer(settings(bg_a))(means3D=a["means"], means2D=torch.zeros_like(a["means"]), opacities=a["op"],
                                            colors_precomp=a["rgb"], **a["geo"])[0]
    i2 = GaussianRasterizer(settings(bg_b))(means3D=a["means"], means2D=torch.zeros_like(a["means"]), opacities=a["op"],
                                            colors_precomp=sil, **a["geo"])[0]
    ((i1 * w1).sum() + (i2 * w2).sum()).backward()
    # one pair render
    b = leaves()
    j1, j2, radii = GaussianRasterizer(settings(bg_a)).forward_pair(
        means3D=b["means"], means2D=torch.zeros_like(b["means"]), opacities=b["op"], colors_b=sil,
        bg_b=d(np.asarray(bg_b, np.float32)), colors_precomp=b["rgb"], **b["geo"])
    ((j1 * w1).sum() + (j2 * w2).sum()).backward()
    assert torch.equal(j1, i1) and torch.equal(j2, i2)
    for k in a:
        if k == "geo":
            continue
        ga_, gb_ = a[k].grad, b[k].grad
        assert gb_ is not None and float((ga_ - gb_).norm() / ga_.norm()) < 2e-5, k     # float atomics reorder
    # and against the oracle's two independent views
    if layout == "cov":
        g2 = dict(g, colors=sil_np)
        o1, r1 = util.oracle_forward(g, cam, W, H, bg_a)
        o2, r2 = util.oracle_forward(g2, cam, W, H, bg_b)
        assert_image_close(j1.detach().cpu().numpy(), r1["color"])
        assert_image_close(j2.detach().cpu().numpy(), r2["color"])
        ga, gb = o1.backward(r1, w1.cpu().numpy()), o2.backward(r2, w2.cpu().numpy())
        assert util.rel_l2(b["means"].grad.cpu().numpy(), ga["means3D"] + gb["means3D"]) < GRAD_TOL
        assert util.rel_l2(b["cov"].grad.cpu().numpy(), ga["cov3D"] + gb["cov3D"]) < GRAD_TOL
        assert util.rel_l2(b["op"].grad.cpu().numpy().ravel(), ga["opacity"] + gb["opacity"]) < GRAD_TOL
        assert util.rel_l2(b["rgb"].grad.cpu().numpy(), ga["colors"]) < GRAD_TOL
    # the second image's colours cannot ask for a gradient
    with pytest.raises(RuntimeError):
        GaussianRasterizer(settings(bg_a)).forward_pair(
            means3D=b["means"], means2D=torch.zeros_like(b["means"]), opacities=b["op"],
            colors_b=sil.clone().requires_grad_(True), bg_b=d(np.asarray(bg_b, np.float32)), colors_precomp=b["rgb"],
            **b["geo"])
