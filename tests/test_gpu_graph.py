"""-m gpu tests of training.GraphedViewStep: one view captured into a CUDA graph must give the eager path's loss and
gradients, for device-resident and host-fed inputs, and must notice (and recover from) a binning overflow that the
eager path would have handled in the middle of the view."""
import numpy as np
import pytest
import torch

from d3ga_b200 import diff_gaussian_rasterization as dgr
from d3ga_b200 import synth, training

pytestmark = pytest.mark.gpu


def setup(name="tiny", n_cams=8):
    dev = torch.device("cuda", 0)
    scene = synth.make_workload(name)
    cams = synth.camera_ring(scene.W, scene.H, n_cams)
    avatar = training.Avatar(scene, dev)
    bank = training.CameraBank(cams, dev)
    bg = torch.ones(3, device=dev)
    wh = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (3, scene.H, scene.W)).astype(np.float32)).pin_memory()
    return dev, scene, cams, avatar, bank, bg, wh


def rel(a, b):
    return float((a - b).norm() / b.norm())


def report_parity(record):
    import json
    import os

    path = os.environ.get("D3GA_PARITY_LOG")
    if path:
        with open(path, "a") as f:
            f.write(json.dumps(record) + "\n")
    print("[parity]", json.dumps(record))


def test_graphed_step_matches_eager(cuda_lib):
    dev, scene, cams, avatar, bank, bg, wh = setup()
    w = wh.to(dev)
    ids = [0, 3, 5, 6]
    tot_e = training.multi_view_step(avatar, bank, ids, w, bg).clone()
    g_e = avatar.grads.flat.clone()
    gv = training.GraphedViewStep(avatar, cams[0], bg)
    tot_g = training.multi_view_step_graphed(avatar, bank, ids, gv, weights=w).clone()
    assert gv.captures == 1 and gv.replays == len(ids) and gv.launches_per_replay >= 10
    assert abs(float(tot_g) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
    assert rel(avatar.grads.flat, g_e) < 1e-5          # float atomics reorder, nothing else differs
    # a second step re-uses the graph; host-fed inputs (camera, cage vertices and image uploaded a view ahead)
    feed = training.HostFeed(wh, dev, cam_host=bank.host, tetpoints_host=avatar.tetpoints_host)
    loss_h = torch.zeros((), dtype=torch.float32).pin_memory()
    training.multi_view_step_graphed(avatar, bank, ids, gv, from_host=True, feed=feed, loss_host=loss_h)
    torch.cuda.synchronize()
    assert gv.captures == 1
    assert abs(float(loss_h) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
    assert rel(avatar.grads.flat, g_e) < 1e-5
    loss_e = torch.zeros((), dtype=torch.float32).pin_memory()
    training.multi_view_step(avatar, bank, ids, w, bg, from_host=True, feed=feed, loss_host=loss_e)
    assert abs(float(loss_e) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
    assert rel(avatar.grads.flat, g_e) < 1e-5


def test_pipelined_views_match_eager(cuda_lib):
    """Two graph-replayed views in flight on two streams (shadow avatars, own gradient buffers) give the eager step's
    loss and gradients, device-resident and host-fed (an odd number of views: the lanes end unevenly)."""
    dev, scene, cams, avatar, bank, bg, wh = setup()
    w = wh.to(dev)
    ids = [0, 3, 5, 6, 2]
    tot_e = training.multi_view_step(avatar, bank, ids, w, bg).clone()
    g_e = avatar.grads.flat.clone()
    pv = training.PipelinedViews(avatar, cams[0], bg, lanes=2)
    tot_p = training.multi_view_step_pipelined(avatar, bank, ids, pv, weights=w).clone()
    assert pv.captures == 2
    assert abs(float(tot_p) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
    assert rel(avatar.grads.flat, g_e) < 1e-5
    feed = training.HostFeed(wh, dev, cam_host=bank.host, tetpoints_host=avatar.tetpoints_host, slots=4)
    loss_h = torch.zeros((), dtype=torch.float32).pin_memory()
    for _ in range(2):
        training.multi_view_step_pipelined(avatar, bank, ids, pv, from_host=True, feed=feed, loss_host=loss_h)
        assert abs(float(loss_h) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
        assert rel(avatar.grads.flat, g_e) < 1e-5
    assert pv.captures == 2


def test_graphed_step_recovers_from_overflow(cuda_lib):
    """Capture with a binning workspace that is too small for the other cameras: the replay must report the overflow,
    the step must be redone on a re-captured graph and still equal the eager result."""
    dev, scene, cams, avatar, bank, bg, wh = setup()
    w = wh.to(dev)
    ids = [0, 1, 2, 3]
    tot_e = training.multi_view_step(avatar, bank, ids, w, bg).clone()
    g_e = avatar.grads.flat.clone()
    gv = training.GraphedViewStep(avatar, cams[0], bg)
    gv.capture(bank.dev[0])
    key = (dev, scene.N, scene.W, scene.H)
    assert key in dgr._capacity_hint
    # shrink the hint and capture again without the eager warm-up correcting it
    small = 64
    dgr._capacity_hint[key] = small
    gv.capture(bank.dev[0], warm=False)
    before = gv.captures
    tot_g = training.multi_view_step_graphed(avatar, bank, ids, gv, weights=w).clone()
    assert gv.captures == before + 1, "the overflow must force a re-capture"
    assert dgr._capacity_hint[key] > small
    assert abs(float(tot_g) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
    assert rel(avatar.grads.flat, g_e) < 1e-5


def test_overflow_of_a_middle_view_is_seen(cuda_lib):
    """Only the MIDDLE view of a lane overflows its binning workspace (the views around it fit): the step must still
    notice, re-capture and equal the eager result.  The counter of a lane is created before its capture; when it was
    first touched inside the capture every replay wiped it and only a step's last view was ever checked -- an
    overflowed view in the middle rendered background only, silently."""
    dev, scene, cams, avatar, bank, bg, wh = setup(n_cams=16)
    w = wh.to(dev)
    Ds = []
    for i in range(len(cams)):
        training.multi_view_step(avatar, bank, [i], w, bg)
        Ds.append(int(dgr.last_stats["num_rendered"]))
    small, big = int(np.argmin(Ds)), int(np.argmax(Ds))
    assert Ds[big] > Ds[small] + 8
    for lanes in (1, 2):
        ids = [small, big, small] if lanes == 1 else [small, small, big, big, small, small]
        tot_e = training.multi_view_step(avatar, bank, ids, w, bg).clone()
        g_e = avatar.grads.flat.clone()
        key = (dev, scene.N, scene.W, scene.H)
        pv = training.PipelinedViews(avatar, cams[small], bg, lanes=lanes)
        pv.capture(bank.dev[small])
        dgr._capacity_hint[key] = Ds[big] - 4          # fits every view but the big one
        for g in pv.gvs:
            g.capture(bank.dev[small], warm=False)
        before = pv.captures
        tot = training.multi_view_step_pipelined(avatar, bank, ids, pv, weights=w).clone()
        assert pv.captures == before + lanes, "the overflow of the middle view must force a re-capture"
        assert dgr._capacity_hint[key] >= Ds[big]
        assert abs(float(tot) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
        assert rel(avatar.grads.flat, g_e) < 1e-5


def test_pipelined_metric_step_matches_oracle(cuda_lib, oracle_libs):
    """The path bench.py times -- 4 graph-replayed views in flight (PipelinedViews: shadow avatars, rows sized for the
    views in flight) on the METRIC workload (1 M Gaussians, 1024^2) -- against the oracle: the step's loss and all six
    parameter gradients summed over 8 cameras of the ring, device-resident and host-fed.  Bar: 1e-3 relative L2.

    Two references (parity_util.oracle_view): (a) the rasterizer oracle evaluated at the means / covariances the product's
    own deformation kernel produced (asserted at the bar; the deformation kernel itself is held to 2e-6 of the
    deformation oracle right here), and (b) the oracle end to end, whose rasterizer sees the torch-CPU deformation's
    last-bit-different inputs -- reported, and held to 5e-3, because at that point the comparison measures the
    discontinuity noise of float32 rasterization (f32 vs f64 oracle: 3-6e-4 per view), not the kernels."""
    import parity_util as util
    from d3ga_b200.tetra_sampler import deform_gaussians

    dev, scene, cams, avatar, bank, bg, wh = setup("metric", n_cams=160)
    w = wh.to(dev)
    ids = [0, 20, 45, 70, 95, 110, 135, 159]           # 4 elevations x front / side / back
    with torch.no_grad():
        m_gpu, c_gpu = deform_gaussians(avatar.tetpoints, avatar.tets, avatar.tetra_id, avatar.barys, avatar.cg,
                                        avatar.scaling, avatar.rotation, log_scales=True)
    m_gpu, c_gpu = m_gpu.cpu().numpy(), c_gpu.cpu().numpy()
    m_ref, c_ref = util.scene_geometry(scene)
    assert util.rel_l2(m_gpu, m_ref) < 2e-6 and util.rel_l2(c_gpu, c_ref) < 2e-6
    refs = {}
    for tag, inputs in (("oracle_at_product_deformation", (m_gpu, c_gpu)), ("oracle_end_to_end", None)):
        loss_ref, grads_ref = 0.0, None
        for i in ids:
            l, g = util.oracle_view(scene, cams[i], wh.numpy(), raster_inputs=inputs)
            loss_ref += l
            grads_ref = g if grads_ref is None else {k: grads_ref[k] + g[k] for k in g}
        refs[tag] = (loss_ref, grads_ref)
    training.multi_view_step(avatar, bank, [ids[0]], w, bg)      # settles the capacity hint
    pv = training.PipelinedViews(avatar, cams[0], bg, lanes=4)
    feed = training.HostFeed(wh, dev, cam_host=bank.host, tetpoints_host=avatar.tetpoints_host, slots=8)
    loss_h = torch.zeros((), dtype=torch.float32).pin_memory()
    names = ("tetpoints", "barys", "scaling", "rotation", "opacity", "colors")
    for from_host in (False, True):
        if from_host:
            training.multi_view_step_pipelined(avatar, bank, ids, pv, from_host=True, feed=feed, loss_host=loss_h)
            torch.cuda.synchronize()
            loss = float(loss_h)
        else:
            loss = float(training.multi_view_step_pipelined(avatar, bank, ids, pv, weights=w))
        got = {k: getattr(avatar, k).grad.cpu().numpy() for k in names}
        for tag, bar in (("oracle_at_product_deformation", 1e-3), ("oracle_end_to_end", 5e-3)):
            loss_ref, grads_ref = refs[tag]
            rec = {"test": "pipelined_metric_step", "reference": tag, "from_host": from_host, "lanes": 4, "cameras": len(ids),
                   "loss_rel": abs(loss - loss_ref) / abs(loss_ref)}
            rec.update({k: util.rel_l2(got[k], grads_ref[k]) for k in names})
            report_parity(rec)
            assert rec["loss_rel"] < 1e-4, rec
            for k in names:
                assert rec[k] < bar, rec
    assert pv.captures == 4


def test_tet_sorted_storage_is_the_same_avatar(cuda_lib):
    """synth.sort_scene_by_tet only permutes the Gaussians: same loss, gradients permuted the same way."""
    dev, scene, cams, avatar, bank, bg, wh = setup()
    w = wh.to(dev)
    ids = [0, 5]
    tot = training.multi_view_step(avatar, bank, ids, w, bg).clone()
    order = torch.from_numpy(np.argsort(scene.tetra_id, kind="stable")).to(dev)
    sorted_scene = synth.sort_scene_by_tet(scene)
    assert (np.diff(sorted_scene.tetra_id) >= 0).all() and sorted_scene.N == scene.N
    av2 = training.Avatar(sorted_scene, dev)
    tot2 = training.multi_view_step(av2, bank, ids, w, bg).clone()
    assert abs(float(tot2) - float(tot)) <= 2e-5 * abs(float(tot))
    assert rel(av2.tetpoints.grad, avatar.tetpoints.grad) < 1e-4
    for name in ("barys", "scaling", "rotation", "opacity", "colors"):
        assert rel(getattr(av2, name).grad, getattr(avatar, name).grad[order]) < 1e-4, name


def test_graphed_frame_pair_matches_eager(cuda_lib):
    """training.frame_pair_step (deformation with the cage energy + the fused RGB / silhouette render + backward) captured
    into graphs and replayed with two frames in flight gives the eager frame loop's loss and gradients."""
    dev, scene, cams, avatar, bank, bg, wh = setup()
    w = wh.to(dev)
    ids = [0, 3, 5, 6]
    avatar.grads.zero_()
    tot_e = None
    for i in ids:
        rs = bank.settings(i, avatar.W, avatar.H, bg)
        loss = training.frame_pair_step(avatar, rs, w)
        tot_e = loss.detach().clone() if tot_e is None else tot_e + loss.detach()
    g_e = avatar.grads.flat.clone()
    # the eager two-call frame of the reference's trainer gives the same gradients
    avatar.grads.zero_()
    for i in ids:
        training.frame_step(avatar, bank, i, w, bg)
    assert rel(avatar.grads.flat, g_e) < 1e-5
    pv = training.PipelinedViews(avatar, cams[0], bg, lanes=2, step=training.frame_pair_step)
    tot_p = training.multi_view_step_pipelined(avatar, bank, ids, pv, weights=w)
    assert abs(float(tot_p) - float(tot_e)) <= 1e-5 * abs(float(tot_e))
    assert rel(avatar.grads.flat, g_e) < 1e-5
