"""CPU tests pinning oracle/tet_oracle.c: against vectors produced by the REFERENCE's own header code
(tests/golden/tet_golden.npz, generated through oracle/_ref by tests/golden/make_tet_golden.py), against oracle/_ref
live when it is present, and on a lattice known-answer test (sample (tet, bary) -> point -> same tet and bary)."""
import os

import numpy as np
import pytest

import parity_util as util
from d3ga_b200 import synth
from d3ga_b200.tetra_sampler import unique_faces_and_topology
from oracle import tet_oracle

G = np.load(os.path.join(util.GOLDEN, "tet_golden.npz"))


def test_unit_tet_probe(oracle_libs):
    # unit tet, p = (.1,.2,.3) -> bary (.4,.1,.2,.3)
    b = tet_oracle.barycentric(G["points"][:1], G["tets"][:1])
    assert np.allclose(b, [[0.4, 0.1, 0.2, 0.3]], atol=1e-6)


def test_matches_reference_golden(oracle_libs):
    b = tet_oracle.barycentric(G["points"], G["tets"])
    assert np.array_equal(tet_oracle.point_in_tet(G["points"], G["tets"]), G["inside"])
    assert np.allclose(b, G["barys"], rtol=1e-5, atol=1e-5)


@pytest.mark.skipif(tet_oracle.ref() is None, reason="oracle/_ref not built (needs /root/reference)")
def test_matches_reference_live(oracle_libs):
    rng = np.random.default_rng(11)
    tets = rng.normal(size=(300, 4, 3)).astype(np.float32)
    pts = rng.normal(size=(300, 3)).astype(np.float32)
    assert np.array_equal(tet_oracle.point_in_tet(pts, tets), tet_oracle.ref_point_in_tet(pts, tets))
    assert np.allclose(tet_oracle.barycentric(pts, tets), tet_oracle.ref_barycentric(pts, tets), rtol=1e-5, atol=1e-5)


def lattice_case(n_points=400, seed=0):
    pts, tets = synth.make_cage(3, 4, 2)
    faces, topo = unique_faces_and_topology(tets)
    rng = np.random.default_rng(seed)
    tid = rng.integers(0, tets.shape[0], size=n_points)
    b = rng.dirichlet(np.ones(4), size=n_points)
    b = np.where(b.min(1, keepdims=True) < 0.03, 0.25, b).astype(np.float32)
    b /= b.sum(1, keepdims=True)
    X = pts[tets[tid]]
    p = np.einsum("nk,nkc->nc", b, X).astype(np.float32)
    p[::50] += 10.0   # a few points far outside the cage
    return pts, tets, faces, topo, tid, b, p


def test_lattice_embedding(oracle_libs):
    pts, tets, faces, topo, tid, b, p = lattice_case()
    barys, got_tid, active, face = tet_oracle.compute_bary(p, pts[tets], pts[faces], topo)
    outside = np.zeros(len(p), bool)
    outside[::50] = True
    assert np.all(got_tid[outside] == -1) and np.all(active[outside] == -1) and np.all(barys[outside] == 0)
    assert np.array_equal(got_tid[~outside], tid[~outside])
    assert np.allclose(barys[~outside], b[~outside], atol=1e-4)
    # the nearest face is a face of the containing tet
    assert np.all((topo[face[~outside]] == tid[~outside, None]).any(1))


def test_topology_matches_brute_force():
    _, tets = synth.make_cage(2, 2, 2)
    faces, topo = unique_faces_and_topology(tets)
    assert len(np.unique(faces, axis=0)) == len(faces)
    for f, (a, b) in zip(faces, topo):
        owners = [i for i, tt in enumerate(tets) if set(f) <= set(tt)]
        assert owners == [x for x in (a, b) if x != -1]   # increasing tet id, -1 padded


def test_tetra_matches_the_reference_constructor():
    """Tetra.from_tensors against tests/golden/tetra_golden.npz: unique faces, the face -> (tet, tet | -1) map, the
    surface split and gradient() as the reference's own Tetra.__init__ / build() produced them
    (tetra_sampler/tetra.py:12-130, executed by tests/golden/make_tetra_golden.py) -- tets in arbitrary order with
    arbitrary vertex order."""
    import os

    import torch

    from d3ga_b200.tetra_sampler import Tetra

    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "tetra_golden.npz"))
    for k in range(int(G["n_cases"])):
        pts, tets = torch.from_numpy(G[f"{k}_points"]), torch.from_numpy(G[f"{k}_tetras"])
        t = Tetra.from_tensors(pts, tets, device="cpu")
        assert np.array_equal(t.tetra_faces.numpy(), G[f"{k}_tetra_faces"])
        assert np.array_equal(t.triangle_to_tetra.numpy(), G[f"{k}_triangle_to_tetra"])
        assert t.triangle_to_tetra.dtype == torch.int32
        assert np.array_equal(t.surface_triangle.numpy(), G[f"{k}_surface_triangle"])
        assert np.array_equal(t.surface_tetras.numpy(), G[f"{k}_surface_tetras"])
        assert np.array_equal(t.gradient(pts[tets]).numpy(), G[f"{k}_gradient"])
        assert [t.n(), t.n_points()] == list(G[f"{k}_n"])
