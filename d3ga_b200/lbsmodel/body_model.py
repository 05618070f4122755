"""Host-side mirror of the reference's `lbsmodel.body_model` for the cage-vertex skinning path
(SURVEY.md section 8(f) rank 4), backed by d3ga_b200/csrc/lbs.cu through the C-ABI.

Same names, arguments and return shapes as the reference (lbsmodel/body_model.py): `ParameterTransform`,
`LinearBlendSkinning` (.forward, .skinning, .unpose, .unskinning, .compute_rigid_transforms[_matrix],
.compute_root_rigid_transform, .compute_joints_weights), `solve_skeleton_state`, `states_to_matrix`, `LBSModule`
(.pose, .unpose, .template_pose) -- plus `LBSModule.cage_pose`, the whole of Blueman.get (lib/blueman.py:139-167) in two
kernel launches.  Where the reference walks the joints in a Python loop of ~10 torch launches per joint
(body_model.py:327-345), one CTA solves the skeleton level by level.

Gradients flow to the vertices (`verts_unposed` / `template` / `delta`) only; the pose parameters are dataset inputs in
the reference (models/cage_net.py:186-189) and a pose tensor that requires grad raises NotImplementedError.
There is no CPU path: tensors must be CUDA tensors and the library must be built (d3ga_b200/_lib.py raises otherwise).
"""
from __future__ import annotations

from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch
import torch as th
import torch.nn as nn

from .. import _lib


def _cuda_f32(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")        # CHECK_INPUT, include/common.h:3-7
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _no_pose_grad(*tensors):
    for t in tensors:
        if isinstance(t, torch.Tensor) and t.requires_grad and torch.is_grad_enabled():
            raise NotImplementedError("d3ga_b200.lbsmodel: gradients with respect to pose / scale parameters are not "
                                      "implemented (the reference feeds dataset motion, models/cage_net.py:186-189)")


class ParameterTransform(nn.Module):
    """body_model.py:23-47.  `forward` is kept for API parity (a plain matrix product on whatever device holds the
    buffers); the skinning path uses the CSR form of the same matrix inside the skeleton kernel."""

    def __init__(self, lbs_cfg_dict: Dict[str, Any]):
        super().__init__()
        self.channel_names = list(lbs_cfg_dict["channel_names"])
        self.limits = lbs_cfg_dict["limits"]
        self.nr_scaling_params = lbs_cfg_dict["nr_scaling_params"]
        self.nr_position_params = lbs_cfg_dict["nr_position_params"]
        self.nr_total_params = self.nr_scaling_params + self.nr_position_params
        transform = torch.as_tensor(np.asarray(lbs_cfg_dict["transform"]), dtype=torch.float32)
        offsets = torch.as_tensor(np.asarray(lbs_cfg_dict["transform_offsets"]), dtype=torch.float32)
        self.register_buffer("transform_offsets", offsets)
        self.register_buffer("transform", transform)
        # CSR of `transform` for the kernel (rows = 7 * joints); zeros contribute nothing to the reference's product
        dense = transform.numpy()
        rows, cols = np.nonzero(dense)
        row_ptr = np.zeros(dense.shape[0] + 1, np.int32)
        np.cumsum(np.bincount(rows, minlength=dense.shape[0]), out=row_ptr[1:])
        self.register_buffer("csr_row_ptr", torch.from_numpy(row_ptr), persistent=False)
        self.register_buffer("csr_col_idx", torch.from_numpy(cols.astype(np.int32)), persistent=False)
        self.register_buffer("csr_vals", torch.from_numpy(dense[rows, cols].astype(np.float32)), persistent=False)

    def forward(self, pose: th.Tensor) -> th.Tensor:
        return self.transform.mm(pose.t()).t() + self.transform_offsets


def _identity_csr(n_rows: int, device):
    ar = torch.arange(n_rows + 1, dtype=torch.int32, device=device)
    return ar, ar[:-1].contiguous(), torch.ones(n_rows, dtype=torch.float32, device=device)


def _solve(pose, scales, scales_stride, with_zero_copy, row_ptr, col_idx, vals, offsets, joint_offset, joint_rotation,
           parents_i32, bind_state, want_states, want_mats):
    lib = _lib.load()
    B, NJ, dev = int(pose.shape[0]), int(joint_offset.shape[0]), pose.device
    n_pose = int(pose.shape[1])
    n_scale = 0 if scales is None else int(scales.shape[-1])
    nb = 2 * B if with_zero_copy else B
    with _lib.device_guard(dev):
        states = torch.empty((nb, NJ, 8), dtype=torch.float32, device=dev) if want_states else None
        mats = torch.empty((nb, NJ, 3, 4), dtype=torch.float32, device=dev) if want_mats else None
        _lib.check(lib.d3ga_lbs_skeleton_forward(
            B, NJ, n_pose, n_scale, _lib.ptr(pose), _lib.ptr(scales), scales_stride, int(with_zero_copy),
            row_ptr.data_ptr(), _lib.ptr(col_idx), _lib.ptr(vals), offsets.data_ptr(), joint_offset.data_ptr(),
            joint_rotation.data_ptr(), parents_i32.data_ptr(), _lib.ptr(bind_state), _lib.ptr(states), _lib.ptr(mats),
            _lib.raw_stream(dev)), "lbs_skeleton_forward")
    return states, mats


def solve_skeleton_state(param: th.Tensor, joint_offset: th.Tensor, joint_rotation: th.Tensor, joint_parents: th.Tensor):
    """body_model.py:311-347: param [B, 7*NJ] (ParameterTransform outputs) -> skeleton states [B, NJ, 8]."""
    _no_pose_grad(param)
    param = _cuda_f32(param, "param")
    dev = param.device
    n_rows = int(param.shape[1])
    row_ptr, col_idx, vals = _identity_csr(n_rows, dev)
    offsets = torch.zeros(n_rows, dtype=torch.float32, device=dev)
    parents = joint_parents.reshape(-1).to(device=dev, dtype=torch.int32).contiguous()
    states, _ = _solve(param, None, 0, False, row_ptr, col_idx, vals, offsets, _cuda_f32(joint_offset, "joint_offset"),
                       _cuda_f32(joint_rotation, "joint_rotation"), parents, None, True, False)
    return states


def states_to_matrix(bind_state: th.Tensor, target_states: th.Tensor, return_transform: bool = False):
    """body_model.py:350-387: [B, NJ, 3, 4] matrices target o bind^-1."""
    if return_transform:
        raise NotImplementedError("states_to_matrix(return_transform=True) is not on the skinning path")
    lib = _lib.load()
    bind_state = _cuda_f32(bind_state, "bind_state")
    target_states = _cuda_f32(target_states.detach(), "target_states")
    B, NJ, dev = int(target_states.shape[0]), int(target_states.shape[1]), target_states.device
    with _lib.device_guard(dev):
        mats = torch.empty((B, NJ, 3, 4), dtype=torch.float32, device=dev)
        _lib.check(lib.d3ga_lbs_states_to_matrix(B, NJ, bind_state.data_ptr(), target_states.data_ptr(), mats.data_ptr(),
                                                 _lib.raw_stream(dev)), "lbs_states_to_matrix")
    return mats


class _Skin(torch.autograd.Function):
    """out = post(global_scaling * sum_k w_k M_k [v;1]); differentiable in `verts` and `delta`."""

    @staticmethod
    def forward(ctx, verts, delta, delta_unit, skin_indices, skin_weights, mats, global_scaling, root_mats, root_joint,
                pre_mul, unit_div, shift, want_rt):
        lib = _lib.load()
        B, NJ = int(mats.shape[0]), int(mats.shape[1])
        V, K, dev = int(skin_weights.shape[0]), int(skin_weights.shape[1]), mats.device
        vb = int(verts.shape[0])
        if vb not in (1, B) or verts.shape[1] != V:
            raise RuntimeError(f"vertices must be [1|{B}, {V}, 3], got {tuple(verts.shape)}")
        db = 0 if delta is None else int(delta.shape[0])
        if delta is not None and (db not in (1, B) or delta.shape[1] != V):
            raise RuntimeError(f"delta must be [1|{B}, {V}, 3], got {tuple(delta.shape)}")
        with _lib.device_guard(dev):
            out = torch.empty((B, V, 3), dtype=torch.float32, device=dev)
            rt = torch.empty((B, 4, 4), dtype=torch.float32, device=dev) if want_rt else None
            _lib.check(lib.d3ga_lbs_skin_forward(
                B, V, K, NJ, verts.data_ptr(), 0 if vb == 1 else 3 * V, _lib.ptr(delta), 0 if db <= 1 else 3 * V,
                float(delta_unit), skin_indices.data_ptr(), skin_weights.data_ptr(), mats.data_ptr(),
                _lib.ptr(global_scaling), _lib.ptr(root_mats), int(root_joint), _lib.ptr(pre_mul), float(unit_div),
                _lib.ptr(shift), out.data_ptr(), _lib.ptr(rt), _lib.raw_stream(dev)), "lbs_skin_forward")
        ctx.save_for_backward(skin_indices, skin_weights, mats, global_scaling, root_mats, pre_mul)
        ctx.meta = (B, V, K, NJ, vb, db, float(delta_unit), int(root_joint), float(unit_div))
        if want_rt:
            ctx.mark_non_differentiable(rt)
            return out, rt
        return out

    @staticmethod
    def backward(ctx, g, *unused):
        lib = _lib.load()
        skin_indices, skin_weights, mats, global_scaling, root_mats, pre_mul = ctx.saved_tensors
        B, V, K, NJ, vb, db, delta_unit, root_joint, unit_div = ctx.meta
        dev = mats.device
        g = g.contiguous()
        with _lib.device_guard(dev):
            dv = torch.empty((B, V, 3), dtype=torch.float32, device=dev)
            _lib.check(lib.d3ga_lbs_skin_backward(
                B, V, K, NJ, skin_indices.data_ptr(), skin_weights.data_ptr(), mats.data_ptr(), _lib.ptr(global_scaling),
                _lib.ptr(root_mats), root_joint, _lib.ptr(pre_mul), unit_div, g.data_ptr(), 1.0, dv.data_ptr(),
                _lib.raw_stream(dev)), "lbs_skin_backward")
        g_verts = g_delta = None
        if ctx.needs_input_grad[0]:
            g_verts = dv if vb == B else dv.sum(0, keepdim=True)
        if db and ctx.needs_input_grad[1]:
            # v = (t / u + delta) * u  ->  dv/ddelta = u ; dv/dt = 1 (up to rounding)
            g_delta = (dv if db == B else dv.sum(0, keepdim=True)) * delta_unit
        return (g_verts, g_delta) + (None,) * 11


class LinearBlendSkinning(nn.Module):
    """body_model.py:49-308.  Buffers have the reference's names and shapes."""

    def __init__(self, model_json: Dict[str, Any], lbs_config_dict: Dict[str, Any], num_max_skin_joints: int = 8,
                 scale_path: str = None):
        super().__init__()
        model = model_json
        self.param_transform = ParameterTransform(lbs_config_dict)
        bones = model["Skeleton"]["Bones"]
        nr_joints = len(bones)
        self.joint_names = [b["Name"] for b in bones]
        parents = np.array([-1 if b["Parent"] > nr_joints else b["Parent"] for b in bones], np.int64)  # body_model.py:70-73
        joint_rotation = torch.tensor([b["PreRotation"] for b in bones], dtype=torch.float32).reshape(nr_joints, 4)
        joint_offset = torch.tensor([b["TranslationOffset"] for b in bones], dtype=torch.float32).reshape(nr_joints, 3)

        skin_model = model["SkinnedModel"]
        mesh_vertices = torch.tensor(skin_model["RestPositions"], dtype=torch.float32)
        mesh_normals = torch.tensor(skin_model["RestVertexNormals"], dtype=torch.float32)
        pairs = np.asarray(skin_model["SkinningWeights"], dtype=np.float64).reshape(-1, 2)
        offsets = np.asarray(skin_model["SkinningOffsets"], dtype=np.int64)
        nr_vertices = len(offsets) - 1
        # ragged (joint, weight) lists -> fixed [V, num_max_skin_joints] tables, zero padded (body_model.py:86-99)
        counts = np.minimum(np.diff(offsets), num_max_skin_joints)
        slot = np.arange(num_max_skin_joints)[None, :]
        valid = slot < counts[:, None]
        src = (offsets[:-1, None] + slot)[valid]
        skin_weights = np.zeros((nr_vertices, num_max_skin_joints), np.float32)
        skin_indices = np.zeros((nr_vertices, num_max_skin_joints), np.int64)
        skin_weights[valid] = pairs[src, 1].astype(np.float32)
        skin_indices[valid] = pairs[src, 0].astype(np.int64)

        self.register_buffer("mesh_vertices", mesh_vertices)
        self.register_buffer("joint_parents", torch.from_numpy(parents).reshape(nr_joints, 1))
        self.register_buffer("joint_parents_i32", torch.from_numpy(parents.astype(np.int32)), persistent=False)
        self.register_buffer("joint_rotation", joint_rotation)
        self.register_buffer("joint_offset", joint_offset)
        self.register_buffer("mesh_normals", mesh_normals)
        self.register_buffer("mesh_faces", torch.tensor(skin_model["Faces"]["Indices"], dtype=torch.int32).view(-1, 3))
        self.register_buffer("mesh_texture_faces",
                             torch.tensor(skin_model["Faces"]["TextureIndices"], dtype=torch.int32).view(-1, 3))
        self.register_buffer("mesh_texture_coords",
                             torch.tensor(skin_model["TextureCoordinates"], dtype=torch.float32).view(-1, 2))
        self.register_buffer("skin_weights", torch.from_numpy(skin_weights))
        self.register_buffer("skin_indices", torch.from_numpy(skin_indices))
        self.register_buffer("rest_vertices", mesh_vertices)
        self.register_buffer("joints_weights", self.compute_joints_weights())
        # the bind state (zero-parameter solve, body_model.py:102-106) is produced by the skeleton kernel the first time
        # the module is used on a GPU; there is no CPU solve in this package.
        self.register_buffer("_bind_state", None, persistent=False)
        if scale_path is not None:
            scale = np.loadtxt(scale_path).astype(np.float32)[np.newaxis]
            scale = scale[:, 0, :] if len(scale.shape) == 3 else scale
            self.register_buffer("scale", torch.tensor(scale))

    # ---- reference properties -------------------------------------------------------------------------------------
    @property
    def num_verts(self):
        return self.mesh_vertices.size(0)

    @property
    def num_joints(self):
        return self.joint_offset.size(0)

    @property
    def num_params(self):
        return self.skin_weights.shape[-1]

    @property
    def bind_state(self) -> th.Tensor:
        """[1, NJ, 8] skeleton state of the all-zero parameter vector."""
        if self._bind_state is None or self._bind_state.device != self.joint_offset.device:
            pt = self.param_transform
            zero = torch.zeros((1, pt.nr_total_params), dtype=torch.float32, device=self.joint_offset.device)
            states, _ = self._solve(zero, None, False, want_states=True, want_mats=False, bind=False)
            self._bind_state = states
        return self._bind_state

    def compute_joints_weights(self, drop_empty=False):
        """body_model.py:162-174: dense [NJ, V] weights from the (index, weight) tables."""
        idxs_verts = torch.arange(self.num_verts)[:, np.newaxis].expand(-1, self.num_params)
        weights_joints = torch.zeros((self.num_joints, self.num_verts), dtype=torch.float32,
                                     device=self.skin_weights.device)
        weights_joints[self.skin_indices, idxs_verts.to(self.skin_indices.device)] = self.skin_weights
        if drop_empty:
            weights_joints = weights_joints[weights_joints.sum(axis=-1).abs() > 0]
        return weights_joints

    # ---- kernels --------------------------------------------------------------------------------------------------
    def _solve(self, pose, scales, with_zero_copy, want_states, want_mats, bind=True):
        """pose [B, n_pose]; scales None, [n_scale], [1, n_scale] or [B, n_scale]."""
        _no_pose_grad(pose, scales)
        pt = self.param_transform
        pose = _cuda_f32(pose.detach(), "poses")
        stride = 0
        if scales is not None:
            scales = _cuda_f32(scales.detach(), "scales")
            if scales.dim() == 2 and scales.shape[0] == pose.shape[0] and pose.shape[0] > 1:
                stride = int(scales.shape[1])
            elif scales.dim() == 2 and scales.shape[0] not in (1, pose.shape[0]):
                raise RuntimeError(f"scales must be [1|{pose.shape[0]}, n], got {tuple(scales.shape)}")
            if pose.shape[1] + scales.shape[-1] != pt.nr_total_params:
                raise RuntimeError(f"poses ({pose.shape[1]}) + scales ({scales.shape[-1]}) must give "
                                   f"{pt.nr_total_params} model parameters")
        elif pose.shape[1] > pt.nr_total_params:
            raise RuntimeError(f"poses has {pose.shape[1]} parameters, the model {pt.nr_total_params}")
        return _solve(pose, scales, stride, with_zero_copy, pt.csr_row_ptr, pt.csr_col_idx, pt.csr_vals,
                      pt.transform_offsets.reshape(-1), self.joint_offset, self.joint_rotation, self.joint_parents_i32,
                      self.bind_state.reshape(-1, 8) if bind else None, want_states, want_mats)

    def _skin(self, verts, mats, **kw):
        verts = _cuda_f32(verts, "vertices")
        if verts.dim() == 2:
            verts = verts[None]
        args = dict(delta=None, delta_unit=1.0, global_scaling=None, root_mats=None, root_joint=0, pre_mul=None,
                    unit_div=1.0, shift=None, want_rt=False)
        args.update(kw)
        return _Skin.apply(verts, args["delta"], args["delta_unit"], self.skin_indices, self.skin_weights, mats,
                           args["global_scaling"], args["root_mats"], args["root_joint"], args["pre_mul"],
                           args["unit_div"], args["shift"], args["want_rt"])

    # ---- reference methods ----------------------------------------------------------------------------------------
    def compute_rigid_transforms(self, global_pose: th.Tensor, local_pose: th.Tensor, scale: th.Tensor):
        """body_model.py:143-149: skeleton states [B, NJ, 8]."""
        states, _ = self._solve(torch.cat([global_pose, local_pose], -1), scale, False, True, False, bind=False)
        return states

    def compute_rigid_transforms_matrix(self, global_pose: th.Tensor, local_pose: th.Tensor, scale: th.Tensor):
        """body_model.py:151-157."""
        _, mats = self._solve(torch.cat([global_pose, local_pose], -1), scale, False, False, True)
        return mats

    def compute_root_rigid_transform(self, poses: th.Tensor) -> Tuple[th.Tensor, th.Tensor]:
        """body_model.py:176-190: (t [B,3], R [B,3,3]) of joint 1, solved with every scale parameter zero."""
        _, mats = self._solve(poses, None, False, False, True)
        return mats[:, 1, :, 3], mats[:, 1, :, :3]

    def skinning(self, bind_state: th.Tensor, vertices: th.Tensor, target_states: th.Tensor):
        """body_model.py:208-234."""
        assert target_states.size()[1:] == bind_state.size()[1:]
        return self._skin(vertices, states_to_matrix(bind_state.reshape(-1, 8), target_states))

    def unskinning(self, bind_state: th.Tensor, target_states: th.Tensor, verts: th.Tensor):
        """body_model.py:248-283."""
        assert target_states.size()[1:] == bind_state.size()[1:]
        return self._unskin(states_to_matrix(bind_state.reshape(-1, 8), target_states), verts)

    def _unskin(self, mats, verts):
        lib = _lib.load()
        verts = _cuda_f32(verts.detach(), "verts")
        B, V, dev = int(mats.shape[0]), int(self.skin_weights.shape[0]), mats.device
        if tuple(verts.shape) != (B, V, 3):
            raise RuntimeError(f"verts must be [{B}, {V}, 3], got {tuple(verts.shape)}")
        with _lib.device_guard(dev):
            out = torch.empty((B, V, 3), dtype=torch.float32, device=dev)
            _lib.check(lib.d3ga_lbs_unskin_forward(B, V, int(self.skin_weights.shape[1]), int(mats.shape[1]),
                                                   verts.data_ptr(), self.skin_indices.data_ptr(),
                                                   self.skin_weights.data_ptr(), mats.data_ptr(), out.data_ptr(),
                                                   _lib.raw_stream(dev)), "lbs_unskin_forward")
        return out

    def unpose(self, poses: th.Tensor, scales: th.Tensor, verts: th.Tensor):
        """body_model.py:236-246."""
        _, mats = self._solve(poses, scales, False, False, True)
        return self._unskin(mats, verts)

    def forward(self, poses: th.Tensor, scales: th.Tensor, verts_unposed: Optional[th.Tensor] = None) -> th.Tensor:
        """body_model.py:285-308: posed vertices [B, V, 3]."""
        _, mats = self._solve(poses, scales, False, False, True)
        return self._skin(self.mesh_vertices[None] if verts_unposed is None else verts_unposed, mats)


class LBSModule(nn.Module):
    """body_model.py:656-688."""

    def __init__(self, lbs_model_json, lbs_config_dict, lbs_template_verts, lbs_scale, global_scaling):
        super().__init__()
        self.lbs_fn = LinearBlendSkinning(lbs_model_json, lbs_config_dict)
        self.register_buffer("lbs_scale", th.as_tensor(lbs_scale, dtype=th.float32))
        self.register_buffer("lbs_template_verts", th.as_tensor(lbs_template_verts, dtype=th.float32))
        self.register_buffer("global_scaling", th.as_tensor(global_scaling))

    def _gs3(self):
        gs = self.global_scaling.to(torch.float32).reshape(-1)
        if gs.numel() == 1:
            gs = gs.expand(3)
        if gs.numel() != 3:
            raise RuntimeError(f"global_scaling must have 1 or 3 elements, got {gs.numel()}")
        return gs.contiguous()

    def pose(self, motion, template: Optional[th.Tensor] = None):
        """body_model.py:669-675: skinned template * global_scaling."""
        if template is None:
            template = self.lbs_template_verts
        _, mats = self.lbs_fn._solve(motion, self.lbs_scale, False, False, True)
        return self.lbs_fn._skin(template, mats, global_scaling=self._gs3())

    def unpose(self, verts, motion):
        """body_model.py:677-682."""
        return self.lbs_fn.unpose(motion, self.lbs_scale, verts / self.global_scaling) - self.lbs_template_verts

    def template_pose(self, motion):
        """body_model.py:684-688."""
        return self.pose(motion)

    def cage_pose(self, motion, delta=None, pre_mul=None, center_mass=None, return_rt=False, delta_unit=100.0,
                  unit_div=1000.0, root_joint=1):
        """The whole of Blueman.get (lib/blueman.py:139-167; .skinning :101-116) in two launches.

        motion [B, n_pose]; delta [1|B, V, 3] or None ((template/100 + delta)*100, :105-106); pre_mul = Blueman.rot180
        [4,4] or [1,4,4]; center_mass [.., 3] or None (None: no shift -- the caller fixes it from the first result like
        blueman.py:154-155).  Returns geom [B, V, 3] in metres (and RT [B,4,4] when return_rt)."""
        fn = self.lbs_fn
        B = int(motion.shape[0])
        _, mats = fn._solve(motion, self.lbs_scale, True, False, True)      # [2B]: with lbs_scale, then zero scales
        if pre_mul is not None:
            pre_mul = _cuda_f32(pre_mul, "pre_mul").reshape(4, 4)
        if center_mass is not None:
            center_mass = _cuda_f32(center_mass.detach(), "center_mass").reshape(3)
        if delta is not None:
            delta = _cuda_f32(delta, "delta")
            if delta.dim() == 2:
                delta = delta[None]
        template = self.lbs_template_verts.detach()
        return fn._skin(template, mats[:B], delta=delta, delta_unit=delta_unit, global_scaling=self._gs3(),
                        root_mats=mats[B:], root_joint=root_joint, pre_mul=pre_mul, unit_div=unit_div, shift=center_mass,
                        want_rt=return_rt)


class _SkinDense(torch.autograd.Function):
    @staticmethod
    def forward(ctx, verts, delta, weights, nn_ids, A, Bs, Rh, Th):
        lib = _lib.load()
        B, J = int(A.shape[0]), int(A.shape[1])
        V, Vd, dev = int(nn_ids.shape[0]), int(weights.shape[0]), A.device
        with _lib.device_guard(dev):
            out = torch.empty((B, V, 3), dtype=torch.float32, device=dev)
            _lib.check(lib.d3ga_lbs_skin_dense_forward(B, V, Vd, J, verts.data_ptr(), _lib.ptr(delta), weights.data_ptr(),
                                                       nn_ids.data_ptr(), A.data_ptr(), Bs.data_ptr(), Rh.data_ptr(),
                                                       Th.data_ptr(), out.data_ptr(), _lib.raw_stream(dev)),
                       "lbs_skin_dense_forward")
        ctx.save_for_backward(weights, nn_ids, A, Rh)
        ctx.has_delta = delta is not None
        return out

    @staticmethod
    def backward(ctx, g):
        lib = _lib.load()
        weights, nn_ids, A, Rh = ctx.saved_tensors
        B, J = int(A.shape[0]), int(A.shape[1])
        V, Vd, dev = int(nn_ids.shape[0]), int(weights.shape[0]), A.device
        g = g.contiguous()
        with _lib.device_guard(dev):
            dv = torch.empty((B, V, 3), dtype=torch.float32, device=dev)
            _lib.check(lib.d3ga_lbs_skin_dense_backward(B, V, Vd, J, weights.data_ptr(), nn_ids.data_ptr(), A.data_ptr(),
                                                        Rh.data_ptr(), g.data_ptr(), dv.data_ptr(), _lib.raw_stream(dev)),
                       "lbs_skin_dense_backward")
        dv = dv.sum(0, keepdim=True)
        return (dv if ctx.needs_input_grad[0] else None, dv if ctx.has_delta and ctx.needs_input_grad[1] else None,
                None, None, None, None, None, None)


def smpl_deform(skin_weights, nn_ids, vertices, A, Bs, Rh, Th, delta=None):
    """Smplman.deform (lib/smplman.py:155-171) in one launch: out = Rodrigues(Rh) (T[nn] [v + delta + Bs[nn]; 1]) + Th with
    T = skin_weights @ A.  skin_weights [Vd, J]; nn_ids [V] int64; vertices, delta [1, V, 3]; A [B, J, 4, 4]; Bs [B, Vd, 3];
    Rh, Th [B, 3].  Gradients flow to vertices and delta."""
    _no_pose_grad(A, Bs, Rh, Th)
    vertices = _cuda_f32(vertices, "vertices").reshape(1, -1, 3)
    if delta is not None:
        delta = _cuda_f32(delta, "delta").reshape(1, -1, 3)
    if nn_ids.dtype != torch.int64 or not nn_ids.is_cuda:
        raise RuntimeError("nn_ids must be a CUDA int64 tensor")
    return _SkinDense.apply(vertices, delta, _cuda_f32(skin_weights, "skin_weights"), nn_ids.contiguous(),
                            _cuda_f32(A.detach(), "A"), _cuda_f32(Bs.detach(), "Bs"), _cuda_f32(Rh.detach(), "Rh"),
                            _cuda_f32(Th.detach(), "Th"))


def compute_root_transform_cuda(lbs_fn, poses, verts=None):
    """Imported by lib/blueman.py:9 but never called there; the reference body (body_model.py:600-622) expects a
    six-output skinning class that is not part of the open-source tree.  Use
    LinearBlendSkinning.compute_root_rigid_transform."""
    raise NotImplementedError("compute_root_transform_cuda: use LinearBlendSkinning.compute_root_rigid_transform")


_CHANNELS = ["tx", "ty", "tz", "rx", "ry", "rz", "sc"]
_FLOAT = r"[-+]?[0-9]*\.?[0-9]+"


def load_momentum_cfg(model, lbs_config_txt_fh, nr_scaling_params=None):
    """Parser of the skeleton parameter file (body_model.py:390-578), producing the dictionary ParameterTransform reads.

    Lines (after stripping `#` comments):
      `<joint>.<channel> = <w> * <param> [+ <w> * <param> ...]`  -> one row of the [7*NJ, NP] transform per channel;
      `limit <name> minmax|minmax_passive [lo, hi] [weight]`       -> an entry of `limits`;
      `parameterset ...`                                           -> ignored.
    Parameters are numbered in order of first appearance; scaling parameters are those whose name starts with "scale".
    """
    import re

    joint_names = [b["Name"] for b in model["Skeleton"]["Bones"]]
    joint_of = {}
    for i, n in enumerate(joint_names):
        joint_of.setdefault(n, i)
    param_names, triplets, limits = [], [], []
    assign_re = re.compile(r"(\w+).(\w+)\s*=\s*(.*)")
    term_re = re.compile(r"\s*([+-]?[0-9]*\.?[0-9]*)\s\*\s(\w+)\s*")
    limit_re = re.compile(r"limit ([\w.]+) (\w+) (.*)")
    range_re = re.compile(r"\[\s*(" + _FLOAT + r")\s*,\s*(" + _FLOAT + r")\s*\](\s*" + _FLOAT + r")?")
    for raw in lbs_config_txt_fh.readlines():
        line = raw[: raw.find("#")]           # like the reference: a line without '#' loses its last character (the newline)
        if "limit" in line:
            m = limit_re.search(line)
            if m is None:
                continue
            fullname, kind, rest = m.groups()
            parts = fullname.split(".")
            joint = joint_of.get(parts[0])
            channel = _CHANNELS.index(parts[1]) if joint is not None and len(parts) == 2 else -1
            if kind in ("minmax_passive", "minmax"):
                r = range_re.search(rest)
                lo, hi = float(r.group(1)), float(r.group(2))
                weight = float(r.group(3)) if r.group(3) is not None else 1.0
                if channel >= 0:
                    limits.append({"type": "LimitMinMaxJointValue", "str": fullname, "valueIndex": joint * 7 + channel,
                                   "limits": [lo, hi], "weight": weight})
                elif fullname in param_names:
                    limits.append({"type": "LimitMinMaxParameter", "str": fullname,
                                   "parameterIndex": param_names.index(fullname), "limits": [lo, hi], "weight": weight})
            continue
        if "parameterset" in line:
            continue
        m = assign_re.search(line)
        if m is None:
            continue
        joint = joint_of.get(m.group(1))
        if joint is None:
            continue
        row = joint * 7 + _CHANNELS.index(m.group(2))
        for term in m.group(3).split("+"):
            t = term_re.search(term.strip())
            if t is None:
                continue
            name = t.group(2)
            if name not in param_names:
                param_names.append(name)
            triplets.append((row, param_names.index(name), float(t.group(1))))
    transform = np.zeros((7 * len(joint_names), len(param_names)), dtype=np.float32)
    for r, c, v in triplets:
        transform[r, c] = v
    out = {"model_param_names": param_names, "joint_names": joint_names, "channel_names": list(_CHANNELS),
           "limits": limits, "transform": transform,
           "transform_offsets": np.zeros((1, 7 * len(joint_names)), dtype=np.float32)}
    if nr_scaling_params is None:
        out["nr_scaling_params"] = len([s for s in param_names if s.startswith("scale")])
        out["nr_position_params"] = len(param_names) - out["nr_scaling_params"]
    return out
