"""CPU tests pinning oracle/deform_oracle.py against the reference's OWN functions (tests/golden/deform_golden.npz was
produced by tests/golden/make_deform_golden.py executing utils/general_utils.py and tetra_sampler/tetra.py from the
reference tree) and against analytic known answers."""
import math
import os

import numpy as np
import torch as th

import parity_util as util
from d3ga_b200 import synth
from oracle import deform_oracle as do

G = np.load(os.path.join(util.GOLDEN, "deform_golden.npz"))
t = th.from_numpy


def test_pieces_match_reference_functions():
    rot, scales = t(G["rotation"]), t(G["scales"])
    assert np.allclose(do.build_rotation(rot).numpy(), G["R"], atol=1e-6)
    assert np.allclose(do.build_scaling_rotation(scales, rot).numpy(), G["L"], atol=1e-7)
    X = t(G["tetpoints"])[t(G["tets"])][t(G["tetra_id"])]
    assert np.allclose((do.tet_gradient(X) @ t(G["canonical_gradient"])).numpy(), G["J"], atol=1e-5)
    cg = do.canonical_gradient(t(G["canon"]), t(G["tets"]), t(G["tetra_id"]))
    assert np.allclose(cg.numpy(), G["canonical_gradient"], rtol=1e-4, atol=1e-3)


def test_deform_matches_reference_composition():
    means, cov6 = do.deform_gaussians(t(G["tetpoints"]), t(G["tets"]), t(G["tetra_id"]), t(G["barys"]),
                                      t(G["canonical_gradient"]), t(G["scales"]), t(G["rotation"]))
    assert np.allclose(means.numpy(), G["means3D"], atol=1e-6)
    assert util.rel_l2(cov6.numpy(), G["cov6"]) < 1e-6


def _scene():
    return synth.make_scene("kat", 500, 96, 64, 64, seed=1)


def test_identity_deformation():
    """tetpoints == canonical  =>  J = I, cov = R S^2 R^T, means = sum b_k v_k of the canonical tet."""
    sc = _scene()
    means, cov6 = do.deform_gaussians(t(sc.canon_points), t(sc.tets), t(sc.tetra_id), t(sc.barys),
                                      t(sc.canonical_gradient), t(sc.scales), t(sc.rotation))
    L = do.build_scaling_rotation(t(sc.scales), t(sc.rotation))
    expect = do.strip_symmetric(L @ L.transpose(1, 2))
    assert util.rel_l2(cov6.numpy(), expect.numpy()) < 1e-3   # fp32 inverse of the canonical gradient
    canon_pts = np.einsum("nkc,nk->nc", sc.canon_points[sc.tets[sc.tetra_id]], sc.barys)
    assert np.allclose(means.numpy(), canon_pts, atol=1e-6)


def test_rigid_motion():
    """tetpoints = Q x + t  =>  cov = Q Sigma Q^T."""
    sc = _scene()
    a = 0.7
    Q = np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]], np.float32)
    moved = (sc.canon_points @ Q.T + np.array([0.1, -0.2, 0.3], np.float32)).astype(np.float32)
    _, cov6 = do.deform_gaussians(t(moved), t(sc.tets), t(sc.tetra_id), t(sc.barys), t(sc.canonical_gradient),
                                  t(sc.scales), t(sc.rotation))
    L = do.build_scaling_rotation(t(sc.scales), t(sc.rotation))
    S = (L @ L.transpose(1, 2)).numpy()
    expect = np.einsum("ij,njk,lk->nil", Q, S, Q)
    got = np.zeros_like(expect)
    iu = [(0, 0), (0, 1), (0, 2), (1, 1), (1, 2), (2, 2)]
    for k, (i, j) in enumerate(iu):
        got[:, i, j] = got[:, j, i] = cov6[:, k].numpy()
    assert util.rel_l2(got, expect) < 1e-3


def test_fem_energy_zero_at_rest():
    sc = _scene()
    X = t(sc.canon_points)
    Dn_inv = th.linalg.inv(do.tet_gradient(X[t(sc.tets)]))
    e = do.fem_energy(X, t(sc.tets), Dn_inv)
    assert float(e.abs().max()) < 1e-4


def test_fused_block_oracle_matches_reference_golden():
    """oracle.deform_block (cage_net.py:213-215 + :218-230) and oracle.fem_energy (lib/cage.py:348-361) against what the
    reference's own source lines produced (tests/golden/make_deform_golden.py executes them as they stand)."""
    import os

    import numpy as np
    import torch as th

    from oracle import deform_oracle as do

    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "deform_golden.npz"))
    t = lambda k: th.from_numpy(G[k])  # noqa: E731
    m, c = do.deform_block(t("tetpoints"), t("tets"), t("tetra_id"), t("barys"), t("delta_bary"), t("canonical_gradient"),
                           t("scaling"), t("delta_scale"), t("rotation"), t("delta_rot"))
    assert np.allclose(m.numpy(), G["fused_means3D"], atol=1e-6)
    assert np.abs(c.numpy() - G["fused_cov6"]).max() <= 1e-5 * np.abs(G["fused_cov6"]).max()
    e = do.fem_energy(t("tetpoints"), t("tets"), t("Dn_inv"))
    assert np.allclose(e.numpy(), G["fem_energy"], rtol=1e-5, atol=1e-6)
