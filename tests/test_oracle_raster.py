"""CPU tests of the rasterizer oracle (oracle/raster_oracle.c): hand-computed known answers, f32-vs-f64 agreement and
a finite-difference check of the analytic backward.  The reference holds no golden vectors for this path (SURVEY.md
section 4) -- these are the vectors this repo pins its oracle with."""
import numpy as np
import pytest

from d3ga_b200 import synth
from oracle.raster_oracle import RasterOracle
import parity_util as util


def kat_camera(W=64, H=64):
    """Identity view; projection with x' = x, y' = y, w' = z; tan(fov/2) = 1 -> focal = W/2."""
    view = np.eye(4, dtype=np.float32)
    proj = np.zeros((4, 4), np.float32)
    proj[0, 0] = proj[1, 1] = 1.0
    proj[2, 3] = 1.0
    proj[2, 2] = 1.0
    return synth.Camera(W, H, 1.0, 1.0, view, proj, np.zeros(3, np.float32))


def kat_scene():
    means = np.array([[0, 0, 1], [0.5, 0, 2], [0, 0, 0.1], [5, 0, 1]], np.float32)
    sig = np.array([0.05, 0.1, 0.05, 0.05], np.float32)
    cov6 = np.zeros((4, 6), np.float32)
    cov6[:, 0] = cov6[:, 3] = cov6[:, 5] = sig ** 2
    return dict(means=means, cov6=cov6, opacity=np.full((4, 1), 0.5, np.float32),
                colors=np.array([[1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 1]], np.float32))


def test_key_list_known_answer(oracle_libs):
    """Three-Gaussian scene whose radii, rects and (tile|depth) keys are derived by hand in the comments."""
    cam, g = kat_camera(), kat_scene()
    _, st = util.oracle_forward(g, cam, 64, 64, [0, 0, 0])
    # G0: z=1 -> pixel 31.5; cov2D = 32^2 * .0025 + .3 = 2.86 -> radius ceil(3*sqrt(2.86)) = 6 -> tiles x,y in {1,2}
    # G1: z=2, x=.5 -> pixel (39.5, 31.5); cxx = (16^2 + 4^2) * .01 + .3 = 3.02 -> radius 6 -> tile x = 2, y in {1,2}
    # G2: z = 0.1 <= 0.2 culled.  G3: pixel x = 191.5 -> rect clamps to zero area -> culled.
    assert st["radii"].tolist() == [6, 6, 0, 0]
    assert st["tiles_touched"].tolist() == [4, 2, 0, 0]
    assert st["offsets"].tolist() == [4, 6, 6, 6]
    assert np.allclose(st["xy"][:2], [[31.5, 31.5], [39.5, 31.5]])
    d0, d1 = 0x3F800000, 0x40000000
    assert st["keys_unsorted"].tolist() == [(5 << 32) | d0, (6 << 32) | d0, (9 << 32) | d0, (10 << 32) | d0,
                                            (6 << 32) | d1, (10 << 32) | d1]
    assert st["vals_unsorted"].tolist() == [0, 0, 0, 0, 1, 1]
    assert st["sort_bits"] == 32 + 5   # 16 tiles -> getHigherMsb(16) = 5
    assert st["keys"].tolist() == [(5 << 32) | d0, (6 << 32) | d0, (6 << 32) | d1, (9 << 32) | d0,
                                   (10 << 32) | d0, (10 << 32) | d1]
    assert st["point_list"].tolist() == [0, 0, 1, 0, 0, 1]
    r = st["ranges"]
    assert r[5].tolist() == [0, 1] and r[6].tolist() == [1, 3] and r[9].tolist() == [3, 4] and r[10].tolist() == [4, 6]
    assert r[0].tolist() == [0, 0] and r[15].tolist() == [0, 0]


def test_single_gaussian_alpha_known_answer(oracle_libs):
    cam, g = kat_camera(), kat_scene()
    _, st = util.oracle_forward(g, cam, 64, 64, [0.25, 0.5, 0.75])
    # pixel (31,31): d = (.5,.5), conic = 1/2.86 -> alpha0 = .5 * exp(-.5 * .5 / 2.86)
    a0 = 0.5 * np.exp(-0.25 / 2.86)
    # G1 at (39.5,31.5): d = (8.5,.5), conic (1/3.02, 0, 1/2.86)
    a1 = 0.5 * np.exp(-0.5 * (8.5 ** 2 / 3.02 + 0.25 / 2.86))
    assert a1 < 1 / 255   # skipped by the alpha cut
    T = 1 - a0
    px = st["color"][:, 31, 31]
    assert np.allclose(px, [a0 + T * 0.25, T * 0.5, T * 0.75], atol=2e-6)
    assert st["n_contrib"].reshape(64, 64)[31, 31] == 1
    assert np.isclose(st["final_T"].reshape(64, 64)[31, 31], T, atol=1e-6)
    assert np.isclose(st["invdepth"][0, 31, 31], a0 * 1.0, atol=1e-6)
    # a pixel no Gaussian reaches shows the background
    assert np.allclose(st["color"][:, 0, 0], [0.25, 0.5, 0.75])


def test_cull_boundaries(oracle_libs):
    cam = kat_camera()
    means = np.array([[0, 0, 0.2], [0, 0, np.nextafter(np.float32(0.2), np.float32(1))]], np.float32)
    cov6 = np.zeros((2, 6), np.float32)
    cov6[:, [0, 3, 5]] = 1e-4
    g = dict(means=means, cov6=cov6, opacity=np.full((2, 1), 0.5, np.float32), colors=np.ones((2, 3), np.float32))
    _, st = util.oracle_forward(g, cam, 64, 64, [0, 0, 0])
    assert st["radii"][0] == 0 and st["radii"][1] > 0   # z_view <= 0.2 is culled, the next float is kept


def test_empty_and_ragged(oracle_libs):
    cam = synth.make_camera(50, 37)
    rng = np.random.default_rng(0)
    g = util.random_gaussians(0, rng)
    _, st = util.oracle_forward(g, cam, 50, 37, [1, 1, 1])
    assert st["num_rendered"] == 0 and np.allclose(st["color"], 1.0)
    g = util.random_gaussians(300, rng)
    _, st = util.oracle_forward(g, cam, 50, 37, [1, 1, 1])   # 50x37 is not a multiple of 16
    assert st["ranges"].shape[0] == 4 * 3
    assert np.all(st["keys"][1:] >= st["keys"][:-1])
    lens = st["ranges"][:, 1].astype(np.int64) - st["ranges"][:, 0]
    assert lens.sum() == st["num_rendered"]


@pytest.mark.parametrize("path", ["precomp", "sh"])
def test_f32_matches_f64(oracle_libs, path):
    rng = np.random.default_rng(2)
    cam = synth.make_camera(96, 64, azim_deg=10, elev_deg=5)
    g = util.random_gaussians(400, rng, with_behind=False)
    _, a = util.oracle_forward(g, cam, 96, 64, [0.1, 0.2, 0.3], path)
    _, b = util.oracle_forward(g, cam, 96, 64, [0.1, 0.2, 0.3], path, dtype=np.float64)
    assert np.mean(a["radii"] == b["radii"]) > 0.99
    assert np.abs(a["color"] - b["color"]).max() < 5e-3     # rare threshold flips allowed
    assert np.mean(np.abs(a["color"] - b["color"]) < 1e-4) > 0.995


@pytest.mark.parametrize("path,aa", [("precomp", False), ("sh", False), ("sh", True)])
def test_backward_matches_finite_differences(oracle_libs, path, aa):
    """Central differences in float64 validate the analytic backward the CUDA kernels are compared with."""
    rng = np.random.default_rng(5)
    W, H = 48, 32
    cam = synth.make_camera(W, H, azim_deg=15, elev_deg=5)
    g = util.random_gaussians(40, rng, spread=0.2, with_behind=False)
    g = {k: v.astype(np.float64) for k, v in g.items()}
    g["rot"] = g["rot"] * rng.uniform(0.8, 1.2, size=(40, 1))
    w = rng.uniform(-1, 1, size=(3, H, W))
    wd = rng.uniform(-1, 1, size=(H, W))
    bg = [0.3, 0.6, 0.9]

    def run(**over):
        gg = dict(g)
        gg.update(over)
        return util.oracle_forward(gg, cam, W, H, bg, path, aa, dtype=np.float64)

    orc, st = run()
    grads = orc.backward(st, w, wd)
    names = dict(means3D="means", opacity="opacity")
    names.update(dict(colors="colors", cov3D="cov6") if path == "precomp" else dict(sh="shs", scales="scales", rotations="rot"))
    for gname, key in names.items():
        base = g[key]
        for _ in range(6):
            idx = tuple(int(rng.integers(0, s)) for s in base.shape)
            eps = 1e-6 * max(1e-3, abs(base[idx])) if gname == "cov3D" else 1e-6
            hi, lo = base.copy(), base.copy()
            hi[idx] += eps
            lo[idx] -= eps
            f = lambda s: (s["color"] * w).sum() + (s["invdepth"][0] * wd).sum()  # noqa: E731
            fd = (f(run(**{key: hi})[1]) - f(run(**{key: lo})[1])) / (2 * eps)
            an = grads[gname].reshape(base.shape)[idx]
            assert abs(fd - an) <= 1e-4 * (abs(fd) + abs(an)) + 1e-9, (gname, idx, fd, an)


@pytest.mark.parametrize("chunk,rows_per_seg", [(32, 3), (96, 1), (4096, 4)])
def test_counting_identity_of_the_direct_tile_scatter(oracle_libs, chunk, rows_per_seg):
    """The arithmetic csrc/raster_direct.cu rests on, replayed in numpy on the oracle's own duplicate list: with the
    Gaussians in (depth bits, index) order and cut into rows of `chunk`, the slot of duplicate (g, tile) in the sorted list
    is  start[tile] + (duplicates of the tile in earlier segments) + (in earlier rows of g's segment) + (before g in
    its row)  -- counts only, no sort of the duplicates.  Must reproduce the oracle's sorted ids, keys and ranges."""
    rng = np.random.default_rng(chunk)
    g = util.random_gaussians(700, rng, sigma=0.1, with_behind=True)
    cam, W, H = synth.make_camera(112, 80, 15, 5), 112, 80
    _, st = util.oracle_forward(g, cam, W, H, [0, 0, 0])
    tiles = (st["keys_unsorted"] >> np.uint64(32)).astype(np.int64)          # one entry per duplicate, Gaussian order
    gid = st["vals_unsorted"].astype(np.int64)
    D, Tl = len(gid), ((W + 15) // 16) * ((H + 15) // 16)
    assert D == st["num_rendered"] and D > 2000
    # depth order of the Gaussians: culled ones carry the key 0xffffffff and sort to the end (preprocess_kernel)
    key = np.where(st["tiles_touched"] > 0, st["depths"].view(np.uint32), np.uint32(0xFFFFFFFF))
    order = np.argsort(key, kind="stable")
    slot_of = np.empty_like(order)
    slot_of[order] = np.arange(len(order))
    slot = slot_of[gid]                                                      # the duplicate's Gaussian in depth order
    row, seg = slot // chunk, slot // chunk // rows_per_seg
    nrows, nseg = int(row.max()) + 1, int(seg.max()) + 1
    M = np.zeros((nrows, Tl), np.int64)                                      # direct_count_kernel
    np.add.at(M, (row, tiles), 1)
    assert M.max() <= min(chunk, 65535)
    segsum = np.add.reduceat(M, np.arange(0, nrows, rows_per_seg), axis=0)   # direct_segsum_kernel
    assert segsum.shape[0] == nseg
    tot = segsum.sum(0)
    segx = np.cumsum(segsum, 0) - segsum
    start = np.cumsum(tot) - tot                                             # direct_plan_kernel
    ranges = np.where(tot[:, None] > 0, np.stack([start, start + tot], 1), 0)
    before = np.zeros_like(M)                                                # direct_base_kernel
    for s in range(nseg):
        blk = M[s * rows_per_seg:(s + 1) * rows_per_seg]
        before[s * rows_per_seg:(s + 1) * rows_per_seg] = np.cumsum(blk, 0) - blk
    assert before.max() <= 65535
    # direct_scatter_kernel: rank inside the row = earlier Gaussians of the row on the same tile
    in_row = np.zeros(D, np.int64)
    o = np.lexsort((slot, tiles, row))
    r_s, t_s = row[o], tiles[o]
    new_group = np.r_[True, (r_s[1:] != r_s[:-1]) | (t_s[1:] != t_s[:-1])]
    idx = np.arange(D)
    in_row[o] = idx - np.maximum.accumulate(np.where(new_group, idx, 0))
    dst = start[tiles] + segx[seg, tiles] + before[row, tiles] + in_row
    assert len(np.unique(dst)) == D and dst.max() == D - 1
    point_list = np.empty(D, np.uint32)
    point_list[dst] = gid
    keys = np.empty(D, np.uint64)
    keys[dst] = (tiles.astype(np.uint64) << np.uint64(32)) | st["depths"].view(np.uint32)[gid].astype(np.uint64)
    assert np.array_equal(point_list, st["point_list"])
    assert np.array_equal(keys, st["keys"])
    assert np.array_equal(ranges.astype(np.uint32), st["ranges"])


def test_analytic_backward_matches_torch_autograd(oracle_libs):
    """The C oracle's ANALYTIC backward (upstream's derivation, restated) against an independently derived one:
    oracle/raster_torch.py writes the forward in differentiable float64 PyTorch and autograd produces the gradient.
    Same image first (so both differentiate the same function), then dL/d{means3D, cov3D, opacity, colours}."""
    import torch as th

    import parity_util as util
    from d3ga_b200 import synth
    from oracle import raster_torch
    from oracle.raster_oracle import RasterOracle

    rng = np.random.default_rng(21)
    for (W, H, n, sigma) in ((48, 40, 70, 0.06), (64, 64, 120, 0.03)):
        g = util.random_gaussians(n, rng, sigma=sigma)
        cam = synth.make_camera(W, H, 15, -10)
        bg = [0.3, 0.6, 0.1]
        orc = RasterOracle(np.float64)
        f64 = lambda a: np.asarray(a, np.float64)  # noqa: E731
        st = orc.forward(f64(g["means"]), f64(g["opacity"]), f64(cam.viewmatrix), f64(cam.projmatrix), f64(cam.campos), W, H,
                         cam.tanfovx, cam.tanfovy, bg, colors_precomp=f64(g["colors"]), cov3D_precomp=f64(g["cov6"]))
        assert st["num_rendered"] > 0
        w = rng.uniform(-1, 1, (3, H, W))
        ga = orc.backward(st, w)
        leaves = [th.tensor(f64(g[k]).reshape(s), requires_grad=True)
                  for k, s in (("means", (n, 3)), ("cov6", (n, 6)), ("opacity", (n,)), ("colors", (n, 3)))]
        img = raster_torch.render(*leaves, cam.viewmatrix, cam.projmatrix, W, H, cam.tanfovx, cam.tanfovy, bg,
                                  st["ranges"], st["point_list"])
        assert np.abs(img.detach().numpy() - st["color"]).max() < 1e-12
        (img * th.from_numpy(w)).sum().backward()
        # (1e-6, not round-off: upstream regularises the conic's derivative, denom2inv = 1 / (denom^2 + 1e-7), which the
        #  oracle restates and autograd of the exact forward does not have -- measured 6e-8 on dL/dcov3D)
        for leaf, name in zip(leaves, ("means3D", "cov3D", "opacity", "colors")):
            assert util.rel_l2(leaf.grad.numpy(), np.asarray(ga[name]).reshape(leaf.shape)) < 1e-6, name
