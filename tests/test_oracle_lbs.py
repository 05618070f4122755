"""CPU tests pinning oracle/lbs_oracle.py against vectors produced by the REFERENCE's own lbsmodel/body_model.py,
lib/blueman.py (Blueman.get) and lib/smplman.py (Smplman.deform) -- tests/golden/lbs_golden.npz, see
tests/golden/make_lbs_golden.py.  The reference computes in float32; the oracle is run in float64 (tolerance = float32
round-off through a ~12-joint chain on millimetre-scale coordinates) and in float32."""
import os

import numpy as np
import pytest

import parity_util as util
from d3ga_b200 import synth
from oracle import lbs_oracle as lo

G = np.load(os.path.join(util.GOLDEN, "lbs_golden.npz"))


def skeleton(dtype):
    model_json, cfg, _, _ = synth.make_skeleton(int(G["V"]), seed=int(G["seed"]))
    return lo.Skeleton(model_json, cfg, dtype=dtype)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_skeleton_states_and_matrices(dtype):
    sk = skeleton(dtype)
    assert np.array_equal(sk.parents, G["joint_parents"].reshape(-1))
    assert np.array_equal(sk.skin_indices, G["skin_indices"])
    assert np.array_equal(sk.skin_weights.astype(np.float32), G["skin_weights"])
    params = np.concatenate([G["motion"], G["scales"]], 1)
    jp = lo.param_transform(sk.transform, sk.transform_offsets, params.astype(dtype))
    assert np.abs(jp - G["joint_params"]).max() < 1e-5
    assert np.abs(sk.bind_state - G["bind_state"]).max() < 2e-3          # millimetres; |t| ~ 2000
    st = sk.states(params)
    assert np.abs(st[..., :3] - G["states"][..., :3]).max() < 2e-3
    assert np.abs(st[..., 3:] - G["states"][..., 3:]).max() < 1e-5
    mats = sk.matrices(params)
    assert np.abs(mats[..., :3] - G["mats"][..., :3]).max() < 1e-5
    assert np.abs(mats[..., 3] - G["mats"][..., 3]).max() < 5e-3


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_skinning_unposing_root(dtype):
    sk = skeleton(dtype)
    m, s = G["motion"].astype(dtype), G["scales"].astype(dtype)
    assert np.abs(sk.forward(m, s) - G["posed"]).max() < 5e-3
    assert np.abs(sk.forward(m, s, G["verts_in"]) - G["posed_in"]).max() < 5e-3
    assert np.abs(sk.unpose(m, s, G["posed_in"]) - G["unposed"]).max() < 1e-2
    assert np.abs(sk.unpose(m, s, G["posed_in"]) - G["verts_in"]).max() < 2e-2    # round trip
    t_root, R_root = sk.root_rigid_transform(m)
    assert np.abs(t_root - G["t_root"]).max() < 2e-3
    assert np.abs(R_root - G["R_root"]).max() < 1e-5


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_cage_pose_matches_blueman_get(dtype):
    sk = skeleton(dtype)
    assert np.array_equal(lo.rot180(), G["blue_rot180"][0])
    g0, RT0, cm = lo.cage_pose(sk, G["motion"][:1], G["scales"][:1], G["global_scaling"], G["template"][None])
    assert np.abs(RT0 - G["blue_RT0"]).max() < 1e-5
    assert np.abs(cm - G["blue_center_mass"]).max() < 1e-5               # metres
    assert np.abs(g0 - G["blue_geom0"]).max() < 1e-5
    g1, _, _ = lo.cage_pose(sk, G["motion"][1:2], G["scales"][:1], G["global_scaling"], G["template"][None],
                            delta=G["blue_delta"], center_mass=G["blue_center_mass"])
    assert np.abs(g1 - G["blue_geom1"]).max() < 1e-5
    if dtype == np.float64:                                             # gradient wrt delta by central differences
        rng = np.random.default_rng(0)
        d = rng.normal(size=G["blue_delta"].shape)
        h = 1e-4
        f = lambda x: (lo.cage_pose(sk, G["motion"][1:2], G["scales"][:1], G["global_scaling"], G["template"][None],
                                    delta=x, center_mass=G["blue_center_mass"])[0] * G["blue_weight"]).sum()
        fd = (f(G["blue_delta"] + h * d) - f(G["blue_delta"] - h * d)) / (2 * h)
        assert np.isclose(fd, (G["blue_grad_delta"] * d).sum(), rtol=1e-4)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_smpl_deform_matches_smplman(dtype):
    c = lambda k: G[k].astype(dtype)
    assert np.abs(lo.batch_rodrigues(c("smpl_Rh")) - G["rodrigues"]).max() < 1e-6
    g = lo.smpl_deform(c("smpl_weights"), G["smpl_nn_ids"], c("smpl_vertices"), c("smpl_A"), c("smpl_Bs"), c("smpl_Rh"),
                       c("smpl_Th"), c("smpl_delta"))
    assert np.abs(g - G["smpl_geom"]).max() < 1e-5


def test_parameter_file_parser_matches_reference():
    """d3ga_b200.lbsmodel.load_momentum_cfg (host logic, no GPU) against the reference's parser on the same text."""
    import io
    import json

    from d3ga_b200.lbsmodel import load_momentum_cfg

    model_json, cfg, _, _ = synth.make_skeleton(int(G["V"]), seed=int(G["seed"]))
    out = load_momentum_cfg(model_json, io.StringIO(str(G["cfg_text"])))
    assert np.array_equal(out["transform"], G["cfg_transform"])
    assert out["model_param_names"] == list(G["cfg_param_names"])
    assert out["limits"] == json.loads(str(G["cfg_limits"]))
    assert out["nr_scaling_params"] == int(G["cfg_nr_scaling"]) and out["nr_position_params"] == int(G["cfg_nr_position"])
    assert out["transform_offsets"].shape == (1, out["transform"].shape[0])


def test_skin_tables_match_reference():
    """LinearBlendSkinning.__init__'s ragged -> [V, 8] tables (host logic) against the reference's buffers."""
    from d3ga_b200.lbsmodel import LinearBlendSkinning

    model_json, cfg, _, _ = synth.make_skeleton(int(G["V"]), seed=int(G["seed"]))
    fn = LinearBlendSkinning(model_json, cfg)
    assert np.array_equal(fn.skin_indices.numpy(), G["skin_indices"])
    assert np.array_equal(fn.skin_weights.numpy(), G["skin_weights"])
    assert np.array_equal(fn.joint_parents.numpy(), G["joint_parents"])
    dense = fn.param_transform.transform.numpy()
    pt = fn.param_transform
    rebuilt = np.zeros_like(dense)
    rp = pt.csr_row_ptr.numpy()
    for r in range(dense.shape[0]):
        rebuilt[r, pt.csr_col_idx.numpy()[rp[r]:rp[r + 1]]] = pt.csr_vals.numpy()[rp[r]:rp[r + 1]]
    assert np.array_equal(rebuilt, dense)


def test_tetra_sampler_import_surface_of_smplman():
    """lib/smplman.py:9,16 imports tetra_sampler.lbs.batch_rodrigues and tetra_sampler.body_model.SMPLlayer: both resolve
    when this package stands in for tetra_sampler, and batch_rodrigues reproduces the reference's own outputs."""
    import importlib
    import sys

    import torch

    import d3ga_b200.tetra_sampler as ts
    saved = {k: sys.modules.get(k) for k in ("tetra_sampler", "tetra_sampler.lbs", "tetra_sampler.body_model")}
    try:
        sys.modules["tetra_sampler"] = ts
        sys.modules.pop("tetra_sampler.lbs", None)
        sys.modules.pop("tetra_sampler.body_model", None)
        lbs = importlib.import_module("tetra_sampler.lbs")
        bm = importlib.import_module("tetra_sampler.body_model")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "lbs_golden.npz"))
    R = lbs.batch_rodrigues(torch.from_numpy(G["smpl_Rh"]))
    assert R.dtype == torch.float32 and np.abs(R.numpy() - G["rodrigues"]).max() < 1e-6
    assert np.allclose(lbs.batch_rodrigues(torch.zeros(2, 3)).numpy(), np.eye(3)[None].repeat(2, 0), atol=1e-6)
    assert hasattr(bm, "SMPLlayer")
    if not os.environ.get("D3GA_REFERENCE_TETRA_SAMPLER"):
        with pytest.raises(ImportError, match="D3GA_REFERENCE_TETRA_SAMPLER"):
            bm.SMPLlayer("models")
