"""Generates tests/golden/lbs_golden.npz by executing the REFERENCE's own cage-LBS code on CPU tensors.

Executed unmodified from /root/reference:
  lbsmodel/quaternion.py + lbsmodel/body_model.py  (ParameterTransform, LinearBlendSkinning, solve_skeleton_state,
      states_to_matrix, LBSModule.pose / unpose)
  lib/blueman.py: Blueman.initialization, .skinning, .get  (and lib/cage.py: CageBase.transform)
  submodules/tetra-sampler/tetra_sampler/lbs.py: batch_rodrigues;  lib/smplman.py: Smplman.deform, .to_homo
Absent third-party imports (loguru, trimesh, networkx, pytorch3d, cager, meshio, ...) are replaced by empty stub
modules; `Tensor.cuda()` is made a no-op while the reference code runs (it hard-codes .cuda()); the Blueman / Smplman
objects are created without running their __init__ (which reads asset files) and given exactly the attributes the
executed methods read.  Inputs: d3ga_b200.synth.make_skeleton (seeded).

    python tests/golden/make_lbs_golden.py      (needs /root/reference)
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.environ.get("D3GA_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
from d3ga_b200 import synth  # noqa: E402


class _Stub(types.ModuleType):
    def __getattr__(self, name):          # any attribute of a stub module is a dummy object
        if name.startswith("__"):
            raise AttributeError(name)
        return type(name, (), {})


for name in ("loguru", "trimesh", "networkx", "pytorch3d", "pytorch3d.transforms", "cager", "cager.ops", "meshio",
             "utils", "utils.mesh_utils", "utils.geometry", "tetra_sampler", "tetra_sampler.body_model", "tetra_sampler.lbs",
             "kornia", "smplx", "pytorch3d.ops", "pytorch3d.structures", "utils.general_utils", "igl", "pymeshlab"):
    sys.modules.setdefault(name, _Stub(name))
sys.modules["loguru"].logger = types.SimpleNamespace(info=lambda *a, **k: None, warning=lambda *a, **k: None)
sys.path.insert(0, REF)
sys.path.insert(0, os.path.join(REF, "submodules", "tetra-sampler"))

body_model = importlib.import_module("lbsmodel.body_model")


def load_file(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


ref_lbs = load_file(os.path.join(REF, "submodules", "tetra-sampler", "tetra_sampler", "lbs.py"), "ref_ts_lbs")
sys.modules["tetra_sampler.lbs"].batch_rodrigues = ref_lbs.batch_rodrigues
blueman = importlib.import_module("lib.blueman")
smplman = importlib.import_module("lib.smplman")

_cuda = torch.Tensor.cuda
torch.Tensor.cuda = lambda self, *a, **k: self
torch.nn.Module.cuda = lambda self, *a, **k: self
out = {}
try:
    V = 300
    model_json, cfg, lbs_scale, motion = synth.make_skeleton(V, seed=5)
    rng = np.random.default_rng(17)
    template = np.asarray(model_json["SkinnedModel"]["RestPositions"], np.float32)
    global_scaling = np.array([1.0, 1.0, 1.0], np.float32) * 1.03
    lbs_module = body_model.LBSModule(model_json, cfg, template, lbs_scale, global_scaling)
    fn = lbs_module.lbs_fn
    t = torch.from_numpy
    motion2 = np.concatenate([motion, motion * 0.5 + 0.1], 0).astype(np.float32)       # batch of 2
    scales2 = np.repeat(lbs_scale, 2, 0)

    params = fn.param_transform(torch.cat([t(motion2), t(scales2)], 1))
    states = body_model.solve_skeleton_state(params, fn.joint_offset, fn.joint_rotation, fn.joint_parents)
    mats = body_model.states_to_matrix(fn.bind_state, states)
    posed = fn(t(motion2), t(scales2))                                   # forward on the rest mesh
    verts_in = t(template[None] + rng.normal(size=(2, V, 3)).astype(np.float32))
    posed_in = fn(t(motion2), t(scales2), verts_in)
    unposed = fn.unpose(t(motion2), t(scales2), posed_in)
    t_root, R_root = fn.compute_root_rigid_transform(t(motion2))
    out.update(V=V, seed=5, motion=motion2, scales=scales2, template=template, global_scaling=global_scaling,
               joint_params=params.numpy(), bind_state=fn.bind_state.numpy(), states=states.numpy(), mats=mats.numpy(),
               skin_indices=fn.skin_indices.numpy(), skin_weights=fn.skin_weights.numpy(),
               joint_parents=fn.joint_parents.numpy(), posed=posed.numpy(), verts_in=verts_in.numpy(),
               posed_in=posed_in.numpy(), unposed=unposed.numpy(), t_root=t_root.numpy(), R_root=R_root.numpy())

    # Blueman.get (lib/blueman.py:139-167) with and without delta; first call fixes center_mass (:154-155)
    man = blueman.Blueman.__new__(blueman.Blueman)
    torch.nn.Module.__init__(man)
    man.lbs_module = lbs_module
    man.center_mass = None
    man.initialization()
    g0, RT0 = man.get(t(motion), return_rt=True)
    delta = torch.from_numpy((rng.normal(size=(1, V, 3)) * 0.01).astype(np.float32)).requires_grad_(True)
    g1 = man.get(t(motion2[1:2]), delta=delta)
    wgt = torch.from_numpy(rng.normal(size=(1, V, 3)).astype(np.float32))
    (g_delta,) = torch.autograd.grad((g1 * wgt).sum(), delta)
    out.update(blue_rot180=man.rot180.numpy(), blue_center_mass=man.center_mass.numpy(), blue_geom0=g0.numpy(),
               blue_RT0=RT0.numpy(), blue_delta=delta.detach().numpy(), blue_geom1=g1.detach().numpy(),
               blue_weight=wgt.numpy(), blue_grad_delta=g_delta.numpy())

    # Smplman.deform (lib/smplman.py:155-171) on synthetic A / Bs / dense weights
    J, Vd, Vc = 24, 500, 200
    sm = smplman.Smplman.__new__(smplman.Smplman)
    torch.nn.Module.__init__(sm)
    w = rng.random((Vd, J)).astype(np.float32) ** 6
    w /= w.sum(1, keepdims=True)
    sm.skin_weights = t(w)
    sm.nn_ids = t(rng.integers(0, Vd, size=Vc))
    sm.body_template_vertices = t(rng.normal(size=(1, Vc, 3)).astype(np.float32))
    sm.lbs_module = types.SimpleNamespace(J_regressor=torch.zeros(J, 1))
    A = np.tile(np.eye(4, dtype=np.float32)[None, None], (1, J, 1, 1))
    A[0, :, :3, :3] = ref_lbs.batch_rodrigues(t(rng.normal(size=(J, 3)).astype(np.float32) * 0.6)).numpy()
    A[0, :, :3, 3] = rng.normal(size=(J, 3)).astype(np.float32) * 0.2
    Bs = (rng.normal(size=(1, Vd, 3)) * 0.01).astype(np.float32)
    Rh = np.array([[0.3, -0.2, 0.5]], np.float32)
    Th = np.array([[0.1, 0.9, -0.4]], np.float32)
    sdelta = torch.from_numpy((rng.normal(size=(1, Vc, 3)) * 0.01).astype(np.float32)).requires_grad_(True)
    sg = sm.deform(t(A), t(Bs), t(Rh), t(Th), sdelta)
    swgt = torch.from_numpy(rng.normal(size=(1, Vc, 3)).astype(np.float32))
    (sg_delta,) = torch.autograd.grad((sg * swgt).sum(), sdelta)
    out.update(smpl_weights=w, smpl_nn_ids=sm.nn_ids.numpy(), smpl_vertices=sm.body_template_vertices.numpy(), smpl_A=A,
               smpl_Bs=Bs, smpl_Rh=Rh, smpl_Th=Th, smpl_delta=sdelta.detach().numpy(), smpl_geom=sg.detach().numpy(),
               smpl_weight=swgt.numpy(), smpl_grad_delta=sg_delta.numpy(),
               rodrigues=ref_lbs.batch_rodrigues(t(Rh)).numpy())

    # load_momentum_cfg (body_model.py:390-578) on a parameter file written from the synthetic transform
    import io
    import json
    names = [b["Name"] for b in model_json["Skeleton"]["Bones"]]
    chans = cfg["channel_names"]
    npos = cfg["nr_position_params"]
    pname = lambda c: (f"p{c}" if c < npos else f"scale_{c - npos}")
    lines = ["# synthetic skeleton parameters", "parameterset all p0 p1"]
    for r in range(cfg["transform"].shape[0]):
        cols = np.nonzero(cfg["transform"][r])[0]
        if len(cols):
            terms = " + ".join(f"{cfg['transform'][r, c]:.3f} * {pname(c)}" for c in cols)
            lines.append(f"{names[r // 7]}.{chans[r % 7]} = {terms}   # row {r}")
    lines += ["limit p7 minmax [-1.5, 0.75] 2.0 #", "limit b_head.rx minmax_passive [ -0.3 , 0.4 ] #",
              "limit nothing minmax [0, 1]", "limit scale_3 minmax [-0.2,0.2] #", ""]
    text = "\n".join(lines)
    parsed = body_model.load_momentum_cfg(model_json, io.StringIO(text))
    out.update(cfg_text=np.array(text), cfg_transform=parsed["transform"], cfg_param_names=np.array(parsed["model_param_names"]),
               cfg_limits=np.array(json.dumps(parsed["limits"])), cfg_nr_scaling=parsed["nr_scaling_params"],
               cfg_nr_position=parsed["nr_position_params"])
finally:
    torch.Tensor.cuda = _cuda
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "lbs_golden.npz"), **out)
print("wrote lbs_golden.npz:", {k: np.asarray(v).shape for k, v in out.items()})
