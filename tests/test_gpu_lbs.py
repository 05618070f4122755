"""-m gpu parity tests of the cage-vertex skinning kernels through the C-ABI: against the vectors
the reference's own lbsmodel/body_model.py, Blueman.get and Smplman.deform produced (tests/golden/lbs_golden.npz) and
against the float64 oracle at the cage size of the benchmark workload.

Tolerances (floating point; coordinates are millimetres with |x| up to ~2000, so one float32 ulp is ~1.2e-4 and a
12-joint chain accumulates a few of them): 5e-3 mm on skeleton translations and skinned millimetre positions,
1e-5 on rotations / quaternions / metre-scale outputs, 1e-4 relative L2 on gradients."""
import os

import numpy as np
import pytest
import torch

import parity_util as util
from d3ga_b200 import lbsmodel, synth
from oracle import lbs_oracle as lo

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(util.GOLDEN, "lbs_golden.npz"))


def cu(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a)).to("cuda", dtype)


def golden_module():
    model_json, cfg, lbs_scale, _ = synth.make_skeleton(int(G["V"]), seed=int(G["seed"]))
    return lbsmodel.LBSModule(model_json, cfg, G["template"], lbs_scale, G["global_scaling"]).cuda()


def test_skeleton_states_and_matrices_match_reference(cuda_lib):
    mod = golden_module()
    fn = mod.lbs_fn
    assert np.abs(fn.bind_state.cpu().numpy() - G["bind_state"]).max() < 2e-3
    motion, scales = cu(G["motion"]), cu(G["scales"])
    states = fn.compute_rigid_transforms(motion[:, :6], motion[:, 6:], scales).cpu().numpy()
    assert np.abs(states[..., :3] - G["states"][..., :3]).max() < 5e-3
    assert np.abs(states[..., 3:] - G["states"][..., 3:]).max() < 1e-5
    mats = fn.compute_rigid_transforms_matrix(motion[:, :6], motion[:, 6:], scales).cpu().numpy()
    assert np.abs(mats[..., :3] - G["mats"][..., :3]).max() < 1e-5
    assert np.abs(mats[..., 3] - G["mats"][..., 3]).max() < 5e-3
    # the module-level functions of the reference API: transformed parameters -> states -> matrices
    st2 = lbsmodel.solve_skeleton_state(cu(G["joint_params"]), fn.joint_offset, fn.joint_rotation, fn.joint_parents)
    assert np.abs(st2.cpu().numpy()[..., :3] - G["states"][..., :3]).max() < 5e-3
    assert np.abs(st2.cpu().numpy()[..., 3:] - G["states"][..., 3:]).max() < 1e-5
    m2 = lbsmodel.states_to_matrix(cu(G["bind_state"]), cu(G["states"])).cpu().numpy()
    assert np.abs(m2[..., :3] - G["mats"][..., :3]).max() < 1e-5
    assert np.abs(m2[..., 3] - G["mats"][..., 3]).max() < 5e-3
    t_root, R_root = fn.compute_root_rigid_transform(motion)
    assert np.abs(t_root.cpu().numpy() - G["t_root"]).max() < 5e-3
    assert np.abs(R_root.cpu().numpy() - G["R_root"]).max() < 1e-5


def test_skinning_and_unposing_match_reference(cuda_lib):
    mod = golden_module()
    fn = mod.lbs_fn
    motion, scales = cu(G["motion"]), cu(G["scales"])
    assert np.abs(fn(motion, scales).cpu().numpy() - G["posed"]).max() < 5e-3
    posed_in = fn(motion, scales, cu(G["verts_in"]))
    assert np.abs(posed_in.cpu().numpy() - G["posed_in"]).max() < 5e-3
    unposed = fn.unpose(motion, scales, cu(G["posed_in"])).cpu().numpy()
    assert np.abs(unposed - G["unposed"]).max() < 1e-2
    assert np.abs(unposed - G["verts_in"]).max() < 2e-2
    # .skinning / .unskinning with explicit states
    sk = fn.skinning(cu(G["bind_state"]), cu(G["verts_in"]), cu(G["states"])).cpu().numpy()
    assert np.abs(sk - G["posed_in"]).max() < 5e-3
    un = fn.unskinning(cu(G["bind_state"]), cu(G["states"]), cu(G["posed_in"])).cpu().numpy()
    assert np.abs(un - G["unposed"]).max() < 1e-2
    # LBSModule.pose = skinning * global_scaling ; unpose inverts it
    posed = mod.pose(motion[:1])
    ref = lo.Skeleton(*synth.make_skeleton(int(G["V"]), seed=int(G["seed"]))[:2]).forward(
        G["motion"][:1].astype(np.float64), G["scales"][:1].astype(np.float64), G["template"][None]) * G["global_scaling"]
    assert np.abs(posed.cpu().numpy() - ref).max() < 5e-3
    back = mod.unpose(posed, motion[:1]).cpu().numpy()
    assert np.abs(back).max() < 2e-2           # unpose returns the offset from the template (body_model.py:680-681)


def test_cage_pose_matches_blueman_get(cuda_lib):
    mod = golden_module()
    rot180 = cu(G["blue_rot180"])
    g0, RT0 = mod.cage_pose(cu(G["motion"][:1]), pre_mul=rot180, return_rt=True)
    assert np.abs(RT0.cpu().numpy() - G["blue_RT0"]).max() < 1e-5
    cm = -1 * g0.mean(dim=1, keepdim=True)                                  # blueman.py:154-155
    assert np.abs(cm.cpu().numpy() - G["blue_center_mass"]).max() < 1e-5
    assert np.abs((g0 + cm).cpu().numpy() - G["blue_geom0"]).max() < 1e-5
    delta = cu(G["blue_delta"]).requires_grad_(True)
    g1 = mod.cage_pose(cu(G["motion"][1:2]), delta=delta, pre_mul=rot180, center_mass=cu(G["blue_center_mass"]))
    assert np.abs(g1.detach().cpu().numpy() - G["blue_geom1"]).max() < 1e-5
    (g1 * cu(G["blue_weight"])).sum().backward()
    assert util.rel_l2(delta.grad.cpu().numpy(), G["blue_grad_delta"]) < 1e-4


def test_cage_pose_batch_and_template_gradient(cuda_lib):
    """Batch of 2 poses sharing one delta: gradients of both poses add up; gradient to explicit vertices."""
    mod = golden_module()
    rot180 = cu(G["blue_rot180"])
    delta = cu(G["blue_delta"]).requires_grad_(True)
    g = mod.cage_pose(cu(G["motion"]), delta=delta, pre_mul=rot180)
    w = cu(np.random.default_rng(2).normal(size=(2, int(G["V"]), 3)))
    (g * w).sum().backward()
    parts = []
    for b in range(2):
        d = cu(G["blue_delta"]).requires_grad_(True)
        (mod.cage_pose(cu(G["motion"][b:b + 1]), delta=d, pre_mul=rot180) * w[b:b + 1]).sum().backward()
        parts.append(d.grad)
    assert util.rel_l2(delta.grad.cpu().numpy(), (parts[0] + parts[1]).cpu().numpy()) < 1e-5
    # finite differences in float64 through the oracle
    sk = lo.Skeleton(*synth.make_skeleton(int(G["V"]), seed=int(G["seed"]))[:2])
    f = lambda x: (lo.cage_pose(sk, G["motion"], G["scales"][:1], G["global_scaling"], G["template"][None], delta=x,
                                center_mass=np.zeros((1, 1, 3)))[0] * w.cpu().numpy()).sum()
    dirn = np.random.default_rng(3).normal(size=G["blue_delta"].shape)
    fd = (f(G["blue_delta"] + 1e-4 * dirn) - f(G["blue_delta"] - 1e-4 * dirn)) / 2e-4
    assert np.isclose(fd, float((delta.grad.cpu().numpy() * dirn).sum()), rtol=1e-3)
    verts = cu(G["verts_in"]).requires_grad_(True)
    out = mod.lbs_fn(cu(G["motion"]), cu(G["scales"]), verts)
    (out * w).sum().backward()
    mats = G["mats"].astype(np.float64)
    expect = np.einsum("bvkij,vk,bvi->bvj", mats[:, G["skin_indices"]][..., :3], G["skin_weights"].astype(np.float64),
                       w.cpu().numpy().astype(np.float64))
    assert util.rel_l2(verts.grad.cpu().numpy(), expect) < 1e-5


def test_smpl_deform_matches_smplman(cuda_lib):
    delta = cu(G["smpl_delta"]).requires_grad_(True)
    out = lbsmodel.smpl_deform(cu(G["smpl_weights"]), cu(G["smpl_nn_ids"], torch.int64), cu(G["smpl_vertices"]),
                               cu(G["smpl_A"]), cu(G["smpl_Bs"]), cu(G["smpl_Rh"]), cu(G["smpl_Th"]), delta)
    assert np.abs(out.detach().cpu().numpy() - G["smpl_geom"]).max() < 1e-5
    (out * cu(G["smpl_weight"])).sum().backward()
    assert util.rel_l2(delta.grad.cpu().numpy(), G["smpl_grad_delta"]) < 1e-5


def test_full_size_cage_vs_oracle_and_edge_cases(cuda_lib):
    """C3-sized cage (V ~ 20 k vertices), 16 poses; deep chain (a 300-joint line); errors raised loudly."""
    V = 20000
    model_json, cfg, lbs_scale, motion = synth.make_skeleton(V, seed=1)
    template = np.asarray(model_json["SkinnedModel"]["RestPositions"], np.float32)
    mod = lbsmodel.LBSModule(model_json, cfg, template, lbs_scale, np.float32(1.0)).cuda()
    rng = np.random.default_rng(4)
    motions = (motion + rng.normal(size=(16, motion.shape[1])).astype(np.float32) * 0.2).astype(np.float32)
    got = mod.cage_pose(cu(motions), pre_mul=cu(lo.rot180())).cpu().numpy()
    sk = lo.Skeleton(model_json, cfg)
    ref, _, _ = lo.cage_pose(sk, motions, lbs_scale, np.float64(1.0), template[None], center_mass=np.zeros((1, 1, 3)))
    assert np.abs(got - ref).max() < 2e-5                      # metres
    # a 300-joint chain: 300 dependent levels in one CTA
    nj = 300
    bones = [{"Name": f"j{i}", "Parent": (10 ** 6 if i == 0 else i - 1), "PreRotation": [0.0, 0.0, 0.0, 1.0],
              "TranslationOffset": [1.0, 0.5, 0.0]} for i in range(nj)]
    tr = np.zeros((7 * nj, nj), np.float32)
    for i in range(nj):
        tr[7 * i + 5, i] = 1.0
    chain = {"Skeleton": {"Bones": bones},
             "SkinnedModel": {"RestPositions": [[0.0, 0.0, 0.0]], "RestVertexNormals": [[0.0, 0.0, 1.0]],
                              "SkinningWeights": [[nj - 1, 1.0]], "SkinningOffsets": [0, 1],
                              "Faces": {"Indices": [0, 0, 0], "TextureIndices": [0, 0, 0]},
                              "TextureCoordinates": [0.0, 0.0]}}
    ccfg = {"channel_names": cfg["channel_names"], "transform": tr, "transform_offsets": np.zeros((1, 7 * nj), np.float32),
            "limits": [], "nr_scaling_params": 0, "nr_position_params": nj}
    fn = lbsmodel.LinearBlendSkinning(chain, ccfg).cuda()
    pose = (rng.normal(size=(1, nj)) * 0.02).astype(np.float32)
    st = fn.compute_rigid_transforms(cu(pose[:, :6]), cu(pose[:, 6:]), cu(np.zeros((1, 0), np.float32))).cpu().numpy()
    ref_st = lo.Skeleton(chain, ccfg).states(pose.astype(np.float64))
    assert np.abs(st - ref_st).max() < 2e-3
    # loud failures: CPU tensors, pose gradients, too many joints
    with pytest.raises(RuntimeError):
        mod.cage_pose(torch.from_numpy(motions))
    with pytest.raises(NotImplementedError):
        mod.cage_pose(cu(motions).requires_grad_(True))
    big = [{"Name": f"j{i}", "Parent": (10 ** 6 if i == 0 else 0), "PreRotation": [0.0, 0.0, 0.0, 1.0],
            "TranslationOffset": [0.0, 0.0, 0.0]} for i in range(1025)]
    bcfg = dict(ccfg, transform=np.zeros((7 * 1025, 1), np.float32), transform_offsets=np.zeros((1, 7 * 1025), np.float32),
                nr_position_params=1)
    bfn = lbsmodel.LinearBlendSkinning(dict(chain, Skeleton={"Bones": big}), bcfg).cuda()
    with pytest.raises(RuntimeError, match="fixed capacity"):
        _ = bfn.bind_state
    # exactly the maximum (1024 joints: 68 KB of shared memory, the opt-in path)
    mcfg = dict(ccfg, transform=np.zeros((7 * 1024, 1), np.float32), transform_offsets=np.zeros((1, 7 * 1024), np.float32),
                nr_position_params=1)
    mfn = lbsmodel.LinearBlendSkinning(dict(chain, Skeleton={"Bones": big[:1024]}), mcfg).cuda()
    bs = mfn.bind_state.cpu().numpy()
    assert bs.shape == (1, 1024, 8) and np.array_equal(bs[0, :, 6], np.ones(1024, np.float32))
