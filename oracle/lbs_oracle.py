"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference's cage-vertex linear blend skinning
(SURVEY.md section 8(f) rank 4).  Nothing in the product path may import this file; only tests/, smoke() and
bench.py's cpu_baseline leg do.

Pinned: tests/golden/lbs_golden.npz holds outputs of the reference's OWN lbsmodel/body_model.py +
lbsmodel/quaternion.py executed on CPU on a synthetic skeleton (tests/golden/make_lbs_golden.py);
tests/test_oracle_lbs.py checks every function below against them.

Each function cites the reference lines it follows.  dtype follows the inputs (float64 for the tight oracle,
float32 to imitate the reference's arithmetic).  Quaternions are (x, y, z, w).
"""
from __future__ import annotations

import numpy as np


# ---- lbsmodel/quaternion.py -------------------------------------------------------------------------------------------
def quat_mul(q, r):
    """Quaternion.batchMul (quaternion.py:169-236): signed products against permuted r."""
    x = q[..., 0] * r[..., 3] + q[..., 1] * r[..., 2] - q[..., 2] * r[..., 1] + q[..., 3] * r[..., 0]
    y = -q[..., 0] * r[..., 2] + q[..., 1] * r[..., 3] + q[..., 2] * r[..., 0] + q[..., 3] * r[..., 1]
    z = q[..., 0] * r[..., 1] - q[..., 1] * r[..., 0] + q[..., 2] * r[..., 3] + q[..., 3] * r[..., 2]
    w = -q[..., 0] * r[..., 0] - q[..., 1] * r[..., 1] - q[..., 2] * r[..., 2] + q[..., 3] * r[..., 3]
    return np.stack([x, y, z, w], -1)


def quat_rot(q, v):
    """Quaternion.batchRot (quaternion.py:238-252): v + 2 (w (a x v) + a x (a x v)), a = q.xyz."""
    a = q[..., :3]
    av = np.cross(a, v)
    aav = np.cross(a, av)
    return v + 2 * (av * q[..., 3:4] + aav)


def quat_invert(q):
    """Quaternion.batchInvert (quaternion.py:254-265): conjugate / |q|^2."""
    sign = np.array([-1.0, -1.0, -1.0, 1.0], q.dtype)
    return q * sign * (1.0 / np.sum(q * q, -1, keepdims=True))


def quat_from_xyz(r):
    """Quaternion.batchFromXYZ (quaternion.py:267-304): XYZ-Euler, half angles (-x/2, y/2, z/2)."""
    rm = r * np.array([-0.5, 0.5, 0.5], r.dtype)
    c, s = np.cos(rm), np.sin(rm)
    return np.stack([
        -s[..., 0] * (c[..., 1] * c[..., 2]) - c[..., 0] * (s[..., 1] * s[..., 2]),
        c[..., 0] * (s[..., 1] * c[..., 2]) - s[..., 0] * (c[..., 1] * s[..., 2]),
        c[..., 0] * (c[..., 1] * s[..., 2]) + s[..., 0] * (s[..., 1] * c[..., 2]),
        c[..., 0] * (c[..., 1] * c[..., 2]) - s[..., 0] * (s[..., 1] * s[..., 2]),
    ], -1)


# ---- lbsmodel/body_model.py -------------------------------------------------------------------------------------------
def param_transform(transform, offsets, pose):
    """ParameterTransform.forward (body_model.py:39-47): transform @ pose^T + offsets -> [B, 7 NJ]."""
    return (transform @ pose.T).T + offsets.reshape(1, -1)


def solve_skeleton_state(param, joint_offset, joint_rotation, joint_parents):
    """solve_skeleton_state (body_model.py:311-347): local (t, q, 2^s) per joint composed down the parent chain.
    param [B, 7 NJ]; returns [B, NJ, 8] = (translation 3, quaternion 4, scale 1)."""
    B = param.shape[0]
    jp = param.reshape(B, -1, 7)
    lt = jp[:, :, 0:3] + joint_offset[None]
    lr = quat_mul(np.broadcast_to(joint_rotation[None], jp[:, :, :4].shape), quat_from_xyz(jp[:, :, 3:6]))
    ls = np.power(np.array(2.0, param.dtype), jp[:, :, 6:7])
    state = np.zeros((B, jp.shape[1], 8), param.dtype)
    for j, p in enumerate(np.asarray(joint_parents).reshape(-1)):
        p = int(p)
        if p != -1:
            gr = quat_mul(state[:, p, 3:7], lr[:, j])
            gt = quat_rot(state[:, p, 3:7], lt[:, j] * state[:, p, 7:8]) + state[:, p, 0:3]
            gs = state[:, p, 7:8] * ls[:, j]
            state[:, j] = np.concatenate([gt, gr, gs], -1)
        else:
            state[:, j] = np.concatenate([lt[:, j], lr[:, j], ls[:, j]], -1)
    return state


def states_to_matrix(bind_state, target_states):
    """states_to_matrix (body_model.py:350-387): target o bind^-1 as [B, NJ, 3, 4] (rotation*scale | translation)."""
    br = quat_invert(bind_state[:, :, 3:7])
    bs = 1.0 / bind_state[:, :, 7:8]
    bt = quat_rot(br, -bind_state[:, :, 0:3]) * bs
    tr = quat_mul(target_states[:, :, 3:7], np.broadcast_to(br, target_states[:, :, 3:7].shape))
    ts = target_states[:, :, 7:8] * bs
    tt = quat_rot(target_states[:, :, 3:7], bt * target_states[:, :, 7:8]) + target_states[:, :, 0:3]
    x, y, z, w = (tr[..., k] for k in range(4))
    twx, twy, twz = 2.0 * x * w, 2.0 * y * w, 2.0 * z * w
    txx, txy, txz = 2.0 * x * x, 2.0 * y * x, 2.0 * z * x
    tyy, tyz, tzz = 2.0 * y * y, 2.0 * z * y, 2.0 * z * z
    col0 = np.stack([1.0 - (tyy + tzz), txy + twz, txz - twy], -1) * ts
    col1 = np.stack([txy - twz, 1.0 - (txx + tzz), tyz + twx], -1) * ts
    col2 = np.stack([txz + twy, tyz - twx, 1.0 - (txx + tyy)], -1) * ts
    return np.stack([col0, col1, col2, tt], -1)


def skinning(mat, skin_indices, skin_weights, vertices):
    """LinearBlendSkinning.skinning (body_model.py:208-234): sum_k w_k (M_k [v;1]).  mat [B,NJ,3,4];
    vertices [B or 1, V, 3]; returns [B, V, 3]."""
    B = mat.shape[0]
    v = np.broadcast_to(vertices, (B,) + vertices.shape[1:])
    vh = np.concatenate([v, np.ones_like(v[..., :1])], -1)
    vs = np.einsum("bvkij,bvj->bvki", mat[:, skin_indices], vh)
    return (vs * skin_weights[None, :, :, None]).sum(2)


def unskinning(mat, skin_indices, skin_weights, verts):
    """LinearBlendSkinning.unskinning (body_model.py:248-283): invert each vertex's blended 4x4 and apply it."""
    sm = (mat[:, skin_indices] * skin_weights[None, :, :, None, None]).sum(2)
    B, V = sm.shape[:2]
    m4 = np.zeros((B, V, 4, 4), sm.dtype)
    m4[:, :, :3] = sm
    m4[:, :, 3, 3] = 1.0
    vh = np.concatenate([verts, np.ones_like(verts[..., :1])], -1)
    return np.einsum("bvij,bvj->bvi", np.linalg.inv(m4), vh)[..., :3]


class Skeleton:
    """The buffers LinearBlendSkinning.__init__ derives from the model dictionary (body_model.py:50-125)."""

    def __init__(self, model_json, cfg, num_max_skin_joints=8, dtype=np.float64):
        bones = model_json["Skeleton"]["Bones"]
        nj = len(bones)
        self.dtype = dtype
        self.parents = np.array([-1 if b["Parent"] > nj else b["Parent"] for b in bones], np.int64)
        # the reference stores these as float32 tensors: round through float32 whatever dtype we compute in
        self.joint_rotation = np.array([b["PreRotation"] for b in bones], np.float32).astype(dtype)
        self.joint_offset = np.array([b["TranslationOffset"] for b in bones], np.float32).astype(dtype)
        sm = model_json["SkinnedModel"]
        self.mesh_vertices = np.array(sm["RestPositions"], np.float32).astype(dtype)
        w = np.array([e[1] for e in sm["SkinningWeights"]], np.float32).astype(dtype)
        idx = np.array([e[0] for e in sm["SkinningWeights"]], np.int64)
        off = np.array(sm["SkinningOffsets"], np.int64)
        V = len(off) - 1
        self.skin_weights = np.zeros((V, num_max_skin_joints), dtype)
        self.skin_indices = np.zeros((V, num_max_skin_joints), np.int64)
        for k in range(num_max_skin_joints):                         # body_model.py:91-99
            left = off[:-1] + k
            ok = left < off[1:]
            self.skin_weights[ok, k] = w[left[ok]]
            self.skin_indices[ok, k] = idx[left[ok]]
        self.transform = np.asarray(cfg["transform"], np.float32).astype(dtype)
        self.transform_offsets = np.asarray(cfg["transform_offsets"], np.float32).astype(dtype)
        self.nr_total_params = cfg["nr_scaling_params"] + cfg["nr_position_params"]
        zero = np.zeros((1, self.nr_total_params), dtype)
        self.bind_state = self.states(zero)                          # body_model.py:106-110

    def states(self, params):
        return solve_skeleton_state(param_transform(self.transform, self.transform_offsets, params.astype(self.dtype)),
                                    self.joint_offset, self.joint_rotation, self.parents)

    def matrices(self, params):
        return states_to_matrix(self.bind_state, self.states(params))

    def forward(self, poses, scales, verts_unposed=None):
        """LinearBlendSkinning.forward (body_model.py:285-308)."""
        mat = self.matrices(np.concatenate([poses, scales], 1))
        v = self.mesh_vertices[None] if verts_unposed is None else verts_unposed.astype(self.dtype)
        return skinning(mat, self.skin_indices, self.skin_weights, v)

    def unpose(self, poses, scales, verts):
        """LinearBlendSkinning.unpose (body_model.py:236-246)."""
        mat = self.matrices(np.concatenate([poses, scales], 1))
        return unskinning(mat, self.skin_indices, self.skin_weights, verts.astype(self.dtype))

    def root_rigid_transform(self, poses):
        """LinearBlendSkinning.compute_root_rigid_transform (body_model.py:176-190): scales are ZERO here; returns
        (t_root [B,3], R_root [B,3,3]) = joint 1's matrix."""
        zeros = np.zeros((poses.shape[0], self.nr_total_params - poses.shape[1]), self.dtype)
        mat = self.matrices(np.concatenate([poses.astype(self.dtype), zeros], 1))
        return mat[:, 1, :, 3], mat[:, 1, :, :3]


def rot180():
    """Blueman.initialization (lib/blueman.py:41-47): z(180) @ y(180) = diag(-1, 1, -1) up to rounding; the reference
    builds it with scipy in float64 and casts to float32."""
    c, s = np.cos(np.pi), np.sin(np.pi)
    y = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    z = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
    m = np.eye(4)
    m[:3, :3] = z @ y
    return m.astype(np.float32)


def cage_pose(skel: Skeleton, motion, lbs_scale, global_scaling, template, delta=None, center_mass=None):
    """Blueman.get (lib/blueman.py:101-116,139-167): skin the cage template (mm), undo the root joint's rigid motion,
    convert to metres, shift by the centre of mass.  template [1,V,3] in the skeleton's units; delta [1,V,3] or None
    (:105-106: (template/100 + delta)*100).  Returns (geom [B,V,3], RT [B,4,4], center_mass [1,1,3])."""
    dt = skel.dtype
    B = motion.shape[0]
    tmpl = np.broadcast_to(template.astype(dt), (B,) + template.shape[1:])
    if delta is not None:
        tmpl = (tmpl / 100.0 + delta.astype(dt)) * 100.0
    scale = np.broadcast_to(lbs_scale.astype(dt), (B, lbs_scale.shape[-1]))
    geom = skel.forward(motion.astype(dt), scale, tmpl) * np.asarray(global_scaling, dt)     # body_model.py:669-675
    t_root, R_root = skel.root_rigid_transform(motion)
    RT = np.tile(np.eye(4, dtype=dt)[None], (B, 1, 1))
    RT[:, :3, :3] = R_root
    RT[:, :3, 3] = t_root / 1000.0
    RT = np.linalg.inv(RT @ rot180().astype(dt)[None])
    geom = geom / 1000
    gh = np.concatenate([geom, np.ones_like(geom[..., :1])], -1)
    geom = np.einsum("bji,bki->bkj", RT, gh)[:, :, :3]                                       # lib/cage.py:54-58
    if center_mass is None:
        center_mass = -1 * geom.mean(1, keepdims=True)                                       # blueman.py:154-155
    return geom + center_mass, RT, center_mass


# ---- SMPL-side variant (lib/smplman.py) -------------------------------------------------------------------------------
def batch_rodrigues(rot_vecs):
    """tetra_sampler/lbs.py:272-302: angle = |r + 1e-8|, R = I + sin K + (1 - cos) K K."""
    angle = np.linalg.norm(rot_vecs + 1e-8, axis=1, keepdims=True)
    d = rot_vecs / angle
    c, s = np.cos(angle)[:, None], np.sin(angle)[:, None]
    z = np.zeros_like(d[:, :1])
    K = np.concatenate([z, -d[:, 2:3], d[:, 1:2], d[:, 2:3], z, -d[:, 0:1], -d[:, 1:2], d[:, 0:1], z], 1).reshape(-1, 3, 3)
    return np.eye(3, dtype=rot_vecs.dtype)[None] + s * K + (1 - c) * (K @ K)


def smpl_deform(skin_weights, nn_ids, vertices, A, Bs, Rh, Th, delta=None):
    """Smplman.deform (lib/smplman.py:155-171): T = W @ A per dense vertex, out = Rh (T[nn] [v + delta + Bs[nn]; 1]) + Th.
    skin_weights [Vd,J]; nn_ids [V]; vertices [1,V,3]; A [B,J,4,4]; Bs [B,Vd,3]; Rh, Th [B,3]."""
    B, J = A.shape[:2]
    T = (skin_weights[None] @ A.reshape(B, J, 16)).reshape(B, -1, 4, 4)
    v = vertices if delta is None else vertices + delta
    g = v + Bs[:, nn_ids]
    gh = np.concatenate([g, np.ones_like(g[..., :1])], -1)
    g = np.einsum("bvij,bvj->bvi", T[:, nn_ids], gh)[..., :3]
    rot = batch_rodrigues(Rh)
    return g @ rot.transpose(0, 2, 1) + Th[:, None]
