"""-m gpu parity tests of rows A0/A1/A2: the fused deformation and the barycentric embedding, through the C-ABI."""
import os

import numpy as np
import pytest
import torch

import parity_util as util
from d3ga_b200 import synth
from d3ga_b200.tetra_sampler import Tetra, compute_bary, deform_gaussians
from oracle import deform_oracle as do
from oracle import tet_oracle
from test_oracle_tet import lattice_case

pytestmark = pytest.mark.gpu
t = torch.from_numpy


def test_deform_matches_reference_golden(cuda_lib):
    G = np.load(os.path.join(util.GOLDEN, "deform_golden.npz"))
    c = lambda k: t(G[k]).cuda()  # noqa: E731
    means, cov6 = deform_gaussians(c("tetpoints"), c("tets"), c("tetra_id"), c("barys"), c("canonical_gradient"),
                                   c("scales"), c("rotation"))
    assert np.allclose(means.cpu().numpy(), G["means3D"], atol=1e-6)
    assert util.rel_l2(cov6.cpu().numpy(), G["cov6"]) < 1e-5


def test_fused_block_matches_reference_golden(cuda_lib):
    """The fused launch -- deltas, exp, normalisation (models/cage_net.py:213-215), the deformation (:218-230) and the cage
    energy (:225, lib/cage.py:348-361) -- against the outputs of the reference's own source lines (deform_golden.npz)."""
    G = np.load(os.path.join(util.GOLDEN, "deform_golden.npz"))
    c = lambda k: t(G[k]).cuda()  # noqa: E731
    means, cov6, energy = deform_gaussians(c("tetpoints"), c("tets"), c("tetra_id"), c("barys"), c("canonical_gradient"),
                                           c("scaling"), c("rotation"), delta_bary=c("delta_bary"),
                                           delta_scale=c("delta_scale"), delta_rot=c("delta_rot"), log_scales=True,
                                           dn_inv=c("Dn_inv"))
    assert np.allclose(means.cpu().numpy(), G["fused_means3D"], atol=1e-6)
    assert util.rel_l2(cov6.cpu().numpy(), G["fused_cov6"]) < 1e-5
    assert np.allclose(energy.cpu().numpy(), G["fem_energy"], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("workload", ["tiny", "c2", "metric"])
def test_fused_block_forward_backward_vs_oracle(cuda_lib, workload):
    """Forward and every gradient of the fused launch (base tensors, deltas, cage vertices incl. the energy's share)
    against torch autograd over the oracle's restatement, up to the 1 M-Gaussian metric workload."""
    sc = synth.make_workload(workload)
    rng = np.random.default_rng(1)
    N = sc.N
    base = dict(tetpoints=sc.tetpoints, barys=sc.barys, scaling=np.log(sc.scales), rotation=sc.rotation * 1.3,
                delta_bary=(0.02 * rng.normal(size=(N, 4))).astype(np.float32),
                delta_scale=(0.2 * rng.normal(size=(N, 3))).astype(np.float32),
                delta_rot=(0.2 * rng.normal(size=(N, 4))).astype(np.float32))
    cpu = {k: t(np.ascontiguousarray(v, dtype=np.float32)).clone().requires_grad_(True) for k, v in base.items()}
    gpu = {k: v.detach().cuda().requires_grad_(True) for k, v in cpu.items()}
    tets, tid, cg = t(sc.tets), t(sc.tetra_id), t(sc.canonical_gradient)
    dn_inv = torch.linalg.inv(do.tet_gradient(t(sc.canon_points)[tets]))
    m_ref, c_ref = do.deform_block(cpu["tetpoints"], tets, tid, cpu["barys"], cpu["delta_bary"], cg, cpu["scaling"],
                                   cpu["delta_scale"], cpu["rotation"], cpu["delta_rot"])
    e_ref = do.fem_energy(cpu["tetpoints"], tets, dn_inv)
    m, c, e = deform_gaussians(gpu["tetpoints"], tets.cuda(), tid.cuda(), gpu["barys"], cg.cuda(), gpu["scaling"],
                               gpu["rotation"], delta_bary=gpu["delta_bary"], delta_scale=gpu["delta_scale"],
                               delta_rot=gpu["delta_rot"], log_scales=True, dn_inv=dn_inv.cuda())
    assert np.allclose(m.detach().cpu().numpy(), m_ref.detach().numpy(), atol=2e-6)
    assert util.rel_l2(c.detach().cpu().numpy(), c_ref.detach().numpy()) < 1e-5
    assert np.allclose(e.detach().cpu().numpy(), e_ref.detach().numpy(), rtol=1e-4, atol=1e-5)
    wm = t(rng.normal(size=m_ref.shape).astype(np.float32))
    wc = t((rng.normal(size=c_ref.shape) * 1e3).astype(np.float32))
    we = t(rng.uniform(0.5, 1.5, size=e_ref.shape).astype(np.float32))
    ((m_ref * wm).sum() + (c_ref * wc).sum() + (e_ref * we).sum()).backward()
    ((m * wm.cuda()).sum() + (c * wc.cuda()).sum() + (e * we.cuda()).sum()).backward()
    for k in base:
        assert util.rel_l2(gpu[k].grad.cpu().numpy(), cpu[k].grad.numpy()) < 1e-3, k
    # a base tensor and its delta have the same gradient (the kernel writes it once)
    assert torch.equal(gpu["barys"].grad, gpu["delta_bary"].grad) and torch.equal(gpu["scaling"].grad, gpu["delta_scale"].grad)


@pytest.mark.parametrize("sparse", [False, True])
@pytest.mark.parametrize("layout", ["as_built", "tet_sorted"])
@pytest.mark.parametrize("mode,pitch4", [("auto", False), ("ordered", False), ("ordered", True), ("padded", False),
                                         ("plain", False)])
def test_backward_accumulation_modes_agree_with_oracle(cuda_lib, monkeypatch, layout, mode, pitch4, sparse):
    """Every way the backward sums the cage-vertex gradients (warp-summed through the sort permutation, warp-summed in
    natural order on tet-sorted rows, padded vector reductions, plain atomics), on rows in both layouts and with a
    Gaussian count that is not a multiple of the warp (the row streams go through shared memory 32 rows at a time)."""
    import d3ga_b200.tetra_sampler as ts
    monkeypatch.setattr(ts, "DEFORM_BACKWARD", mode)
    monkeypatch.setattr(ts, "DEFORM_PITCH4", pitch4)
    sc = synth.make_workload("c2")
    if layout == "tet_sorted":
        sc = synth.sort_scene_by_tet(sc)
    N = sc.N - 13
    assert N % 32 != 0
    rng = np.random.default_rng(5)
    base = dict(tetpoints=sc.tetpoints, barys=sc.barys[:N], scaling=np.log(sc.scales[:N]), rotation=sc.rotation[:N])
    cpu = {k: t(np.ascontiguousarray(v, dtype=np.float32)).clone().requires_grad_(True) for k, v in base.items()}
    gpu = {k: v.detach().cuda().requires_grad_(True) for k, v in cpu.items()}
    tets, tid, cg = t(sc.tets), t(np.ascontiguousarray(sc.tetra_id[:N])), t(np.ascontiguousarray(sc.canonical_gradient[:N]))
    assert bool((tid[1:] >= tid[:-1]).all()) == (layout == "tet_sorted")
    zeros4, zeros3 = torch.zeros(N, 4), torch.zeros(N, 3)
    m_ref, c_ref = do.deform_block(cpu["tetpoints"], tets, tid, cpu["barys"], zeros4, cg, cpu["scaling"], zeros3,
                                   cpu["rotation"], zeros4)
    m, c = deform_gaussians(gpu["tetpoints"], tets.cuda(), tid.cuda(), gpu["barys"], cg.cuda(), gpu["scaling"],
                            gpu["rotation"], log_scales=True)
    assert np.allclose(m.detach().cpu().numpy(), m_ref.detach().numpy(), atol=2e-6)
    assert util.rel_l2(c.detach().cpu().numpy(), c_ref.detach().numpy()) < 1e-5
    wm = t(rng.normal(size=m_ref.shape).astype(np.float32))
    wc = t((rng.normal(size=c_ref.shape) * 1e3).astype(np.float32))
    if sparse:
        # what a rendered view hands back: most Gaussians were never blended and get an exactly-zero incoming gradient
        # (the kernel writes their zeros and compacts the rest over the CTA); some CTAs stay dense, one is all zero
        keep = torch.from_numpy(rng.random(N) < 0.12)
        keep[1024:1536] = True
        keep[2048:2304] = False
        wm[~keep] = 0
        wc[~keep] = 0
    ((m_ref * wm).sum() + (c_ref * wc).sum()).backward()
    ((m * wm.cuda()).sum() + (c * wc.cuda()).sum()).backward()
    if sparse:
        for k in ("barys", "scaling", "rotation"):
            assert not gpu[k].grad[~keep.cuda()].any(), k
    for k in base:
        assert gpu[k].grad.shape == cpu[k].grad.shape
        assert util.rel_l2(gpu[k].grad.cpu().numpy(), cpu[k].grad.numpy()) < 1e-3, k


def test_zero_quaternion_is_nan_like_the_reference(cuda_lib):
    """A zero quaternion: F.normalize (eps 1e-12, cage_net.py:215) leaves it zero, build_rotation then divides 0 by its
    zero norm (utils/general_utils.py:59-61) -> NaN covariance in the reference.  The kernel does the same (and only
    for that Gaussian); a tiny but non-zero quaternion is a valid rotation in both."""
    sc = synth.make_workload("tiny")
    rot = sc.rotation.copy()
    rot[3] = 0.0
    rot[5] *= 1e-15
    args = [t(sc.tetpoints), t(sc.tets), t(sc.tetra_id), t(sc.barys), t(sc.canonical_gradient), t(sc.scales), t(rot)]
    m_ref, c_ref = do.deform_gaussians(*args[:6], torch.nn.functional.normalize(args[6]))
    m, c = deform_gaussians(*[a.cuda() for a in args])
    c, c_ref = c.cpu().numpy(), c_ref.numpy()
    assert np.isnan(c_ref[3]).all() and np.isnan(c[3]).all()
    ok = np.ones(sc.N, bool)
    ok[3] = False
    assert np.isfinite(c[ok]).all() and util.rel_l2(c[ok], c_ref[ok]) < 1e-5
    assert np.allclose(m.cpu().numpy(), m_ref.numpy(), atol=2e-6)


@pytest.mark.parametrize("workload", ["tiny", "c1", "c3"])
def test_deform_forward_backward_vs_oracle(cuda_lib, workload):
    sc = synth.make_workload(workload)
    names = ["tetpoints", "barys", "scales", "rotation"]
    cpu = {k: t(getattr(sc, k)).clone().requires_grad_(True) for k in names}
    cpu["rotation"] = (t(sc.rotation) * 1.3).clone().requires_grad_(True)
    gpu = {k: v.detach().cuda().requires_grad_(True) for k, v in cpu.items()}
    fixed = dict(tets=t(sc.tets), tetra_id=t(sc.tetra_id), cg=t(sc.canonical_gradient))
    m_ref, c_ref = do.deform_gaussians(cpu["tetpoints"], fixed["tets"], fixed["tetra_id"], cpu["barys"], fixed["cg"],
                                       cpu["scales"], cpu["rotation"])
    m, c = deform_gaussians(gpu["tetpoints"], fixed["tets"].cuda(), fixed["tetra_id"].cuda(), gpu["barys"],
                            fixed["cg"].cuda(), gpu["scales"], gpu["rotation"])
    assert np.allclose(m.detach().cpu().numpy(), m_ref.detach().numpy(), atol=2e-6)
    assert util.rel_l2(c.detach().cpu().numpy(), c_ref.detach().numpy()) < 1e-5
    rng = np.random.default_rng(0)
    wm = t(rng.normal(size=m_ref.shape).astype(np.float32))
    wc = t((rng.normal(size=c_ref.shape) * 1e3).astype(np.float32))
    ((m_ref * wm).sum() + (c_ref * wc).sum()).backward()
    ((m * wm.cuda()).sum() + (c * wc.cuda()).sum()).backward()
    for k in names:
        assert util.rel_l2(gpu[k].grad.cpu().numpy(), cpu[k].grad.numpy()) < 1e-3, k


def test_deform_negative_tetra_id_wraps(cuda_lib):
    sc = synth.make_workload("tiny")
    tid = t(sc.tetra_id).clone()
    tid[::10] = -1     # compute_bary's miss value; torch indexing wraps it to the last tet
    args = [t(sc.tetpoints), t(sc.tets), tid, t(sc.barys), t(sc.canonical_gradient), t(sc.scales), t(sc.rotation)]
    m_ref, c_ref = do.deform_gaussians(*args)
    m, c = deform_gaussians(*[a.cuda() for a in args])
    assert np.allclose(m.cpu().numpy(), m_ref.numpy(), atol=2e-6)
    assert util.rel_l2(c.cpu().numpy(), c_ref.numpy()) < 1e-5


def test_compute_bary_lattice(cuda_lib, oracle_libs):
    pts, tets, faces, topo, tid, b, p = lattice_case(2000, seed=2)
    cage = Tetra.from_tensors(t(pts), t(tets), device="cuda")
    assert np.array_equal(cage.tetra_faces.cpu().numpy(), faces)
    assert np.array_equal(cage.triangle_to_tetra.cpu().numpy(), topo)
    barys, tetra_id, active = compute_bary(t(p).cuda(), cage.get_positions(), cage.get_triangles(),
                                           cage.triangle_to_tetra, cage)
    assert tetra_id.dtype == torch.int64 and active.dtype == torch.int32 and barys.shape == (len(p), 4)
    ob, otid, oact, _ = tet_oracle.compute_bary(p, pts[tets], pts[faces], topo)
    assert np.array_equal(tetra_id.cpu().numpy(), otid)
    assert np.array_equal(active.cpu().numpy(), oact)
    assert np.allclose(barys.cpu().numpy(), ob, atol=1e-5)
    outside = np.zeros(len(p), bool)
    outside[::50] = True
    assert np.array_equal(otid[~outside], tid[~outside]) and np.all(otid[outside] == -1)


def test_compute_bary_at_c1_size(cuda_lib, oracle_libs):
    """SURVEY.md section 8 row A0 at the size it quotes: 100 k points in the 12 x 12 x 12 lattice cage (10 368 tets,
    ~21.7 k unique faces) against the tet oracle -- same tets, same miss flags, barycentrics to 1e-5."""
    pts, tets = synth.make_cage(12, 12, 12)
    from d3ga_b200.tetra_sampler import unique_faces_and_topology

    faces, topo = unique_faces_and_topology(tets)
    rng = np.random.default_rng(7)
    n = 100_000
    tid = rng.integers(0, tets.shape[0], n)
    b = rng.dirichlet(np.ones(4), n).astype(np.float32)
    while (b.min(1) < 0.02).any():
        bad = b.min(1) < 0.02
        b[bad] = rng.dirichlet(np.ones(4), int(bad.sum())).astype(np.float32)
    p = np.einsum("nk,nkc->nc", b, pts[tets[tid]]).astype(np.float32)
    p[::97] += 5.0                                            # outside the cage: tetra_id -1, zero barycentrics
    cage = Tetra.from_tensors(t(pts), t(tets), device="cuda")
    barys, tetra_id, active = compute_bary(t(p).cuda(), cage.get_positions(), cage.get_triangles(),
                                           cage.triangle_to_tetra, cage)
    ob, otid, oact, _ = tet_oracle.compute_bary(p, pts[tets], pts[faces], topo)
    assert np.array_equal(tetra_id.cpu().numpy(), otid) and np.array_equal(active.cpu().numpy(), oact)
    assert np.allclose(barys.cpu().numpy(), ob, atol=1e-5)
    inside = np.ones(n, bool)
    inside[::97] = False
    assert np.array_equal(otid[inside], tid[inside]) and np.all(otid[~inside] == -1)


def test_compute_bary_backward_raises(cuda_lib):
    pts, tets, faces, topo, tid, b, p = lattice_case(16)
    cage = Tetra.from_tensors(t(pts), t(tets), device="cuda")
    x = t(p).cuda().requires_grad_(True)
    barys, _, _ = compute_bary(x, cage.get_positions(), cage.get_triangles(), cage.triangle_to_tetra, cage)
    with pytest.raises(NotImplementedError):
        barys.sum().backward()


def test_fem_energy_forward_backward_vs_oracle(cuda_lib):
    """the fused neo-Hookean cage energy (lib/cage.py:349-361) against its torch restatement."""
    from d3ga_b200.tetra_sampler import fem_energy

    sc = synth.make_workload("c1")
    rest = t(sc.canon_points)
    tets = t(sc.tets)
    dn_inv = torch.linalg.inv(do.tet_gradient(rest[tets]))
    x_cpu = t(sc.tetpoints).clone().requires_grad_(True)
    e_ref = do.fem_energy(x_cpu, tets, dn_inv)
    x_gpu = t(sc.tetpoints).cuda().requires_grad_(True)
    e = fem_energy(x_gpu, tets.cuda(), dn_inv.cuda())
    assert e.shape == e_ref.shape
    assert np.allclose(e.detach().cpu().numpy(), e_ref.detach().numpy(), rtol=1e-4, atol=1e-5)
    w = t(np.random.default_rng(1).uniform(0.5, 1.5, size=e_ref.shape).astype(np.float32))
    (e_ref * w).mean().backward()
    (e * w.cuda()).mean().backward()
    assert util.rel_l2(x_gpu.grad.cpu().numpy(), x_cpu.grad.numpy()) < 1e-4
    # rest pose: zero energy (its gradient is NOT zero there: dE/dF = F at F = I, it only cancels at interior vertices)
    e0 = fem_energy(rest.cuda(), tets.cuda(), dn_inv.cuda())
    assert float(e0.abs().max()) < 1e-4
