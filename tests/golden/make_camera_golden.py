"""Generates tests/golden/camera_golden.npz by executing the REFERENCE's own camera and render glue on CPU:

  * lib/cameras.py (Camera, batch_to_camera) and utils/graphics_utils.py (getWorld2View2, getProjectionMatrix),
    unmodified -- the matrices the rasterizer is handed;
  * renderer.py render(), unmodified, with `diff_gaussian_rasterization` resolved to THIS repository's drop-in package
    and its GaussianRasterizer replaced by a recorder: what the reference constructs (the 13 settings fields, by keyword)
    and what it calls the rasterizer with (8 keyword arguments, element [0] of the result) is captured, not imagined.

Modules the reference imports but that are absent here (pytorch3d, kornia, loguru) are empty stubs; `.cuda()` and
`device="cuda"` are neutralised so the code runs on the CPU.  Nothing of the reference is copied: it is imported from
where it lies.      python tests/golden/make_camera_golden.py      (needs /root/reference)
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
import d3ga_b200.diff_gaussian_rasterization as dropin  # noqa: E402

# ---- stubs and CPU neutralisation -------------------------------------------------------------------------------------
for name in ("pytorch3d", "pytorch3d.ops", "pytorch3d.ops.points_normals", "kornia", "kornia.geometry",
             "kornia.geometry.depth", "loguru"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["pytorch3d.ops.points_normals"].estimate_pointcloud_normals = None
sys.modules["kornia.geometry.depth"].spatial_gradient = None
sys.modules["kornia.geometry.depth"].depth_to_normals = None
sys.modules["loguru"].logger = types.SimpleNamespace(info=print, warning=print, error=print)
torch.Tensor.cuda = lambda self, *a, **k: self


def _strip_device(fn):
    def wrapped(*a, **k):
        if k.get("device") is not None and "cuda" in str(k["device"]):
            k["device"] = "cpu"
        return fn(*a, **k)
    return wrapped


torch.tensor = _strip_device(torch.tensor)
torch.zeros_like = _strip_device(torch.zeros_like)
_real_device = torch.device

calls = []


class RecordingRasterizer(torch.nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, **kw):
        rs = self.raster_settings
        calls.append((rs, kw))
        return (torch.zeros(3, rs.image_height, rs.image_width), None, None)


fake = types.ModuleType("diff_gaussian_rasterization")
fake.GaussianRasterizationSettings = dropin.GaussianRasterizationSettings     # the drop-in's own class
fake.GaussianRasterizer = RecordingRasterizer
sys.modules["diff_gaussian_rasterization"] = fake
sys.path.insert(0, REF)
ref_cameras = importlib.import_module("lib.cameras")
ref_renderer = importlib.import_module("renderer")

out = {}
cases = [(512, 512, 0.0, 0.0), (1024, 1024, 20.0, 10.0), (2048, 1334, 135.0, -15.0), (50, 37, -20.0, 30.0)]
for k, (W, H, az, el) in enumerate(cases):
    R, T, fovx, fovy = synth.camera_pose(W, H, az, el)
    batch = dict(camera_id=k, frame_id=7, R=R, T=T, FoVx=fovx, FoVy=fovy, width=W, height=H, crop=[0, 1, 0, 1, W, H])
    cam = ref_cameras.batch_to_camera(batch)
    N = 5
    g = torch.Generator().manual_seed(k)
    pkg = dict(means3D=torch.randn(N, 3, generator=g), cov3D_precomp=torch.rand(N, 6, generator=g),
               opacities=torch.rand(N, 1, generator=g), rgb=torch.rand(N, 3, generator=g), shs=None, sh_degree=0)
    bg = torch.tensor([0.25, 0.5, 0.75])
    res = ref_renderer.render(batch, pkg, bg)
    rs, kw = calls[-1]
    assert res["render"].shape == (3, H, W)
    out[f"{k}_case"] = np.array([W, H, az, el], np.float64)
    out[f"{k}_R"], out[f"{k}_T"], out[f"{k}_fov"] = R, T, np.array([fovx, fovy])
    out[f"{k}_world_view_transform"] = cam.world_view_transform.numpy()
    out[f"{k}_projection_matrix"] = cam.projection_matrix.numpy()
    out[f"{k}_full_proj_transform"] = cam.full_proj_transform.numpy()
    out[f"{k}_camera_center"] = cam.camera_center.numpy()
    # what render() handed the rasterizer
    out[f"{k}_settings_fields"] = np.array(list(rs._fields))
    out[f"{k}_settings_scalars"] = np.array([rs.image_height, rs.image_width, rs.tanfovx, rs.tanfovy, rs.scale_modifier,
                                             rs.sh_degree, rs.prefiltered, rs.debug, rs.antialiasing], np.float64)
    out[f"{k}_settings_viewmatrix"] = rs.viewmatrix.numpy()
    out[f"{k}_settings_projmatrix"] = rs.projmatrix.numpy()
    out[f"{k}_settings_campos"] = rs.campos.numpy()
    out[f"{k}_settings_bg"] = rs.bg.numpy()
    out[f"{k}_call_kwargs"] = np.array(sorted(kw))
    out[f"{k}_call_none"] = np.array(sorted(n for n, v in kw.items() if v is None))
    out[f"{k}_means2D_zero_with_grad"] = np.array([float(kw["means2D"].abs().sum()), float(kw["means2D"].requires_grad)])
# the silhouette render of models/trainer.py:104-110: explicit colours, detach list
calls.clear()
pkg["means3D"].requires_grad_(True)
pkg["cov3D_precomp"].requires_grad_(True)
ref_renderer.render(batch, pkg, torch.zeros(3), colors_precomp=torch.ones(N, 3), detach=["position", "covariance"])
rs, kw = calls[-1]
out["silhouette_detached"] = np.array([float(kw["means3D"].requires_grad), float(kw["cov3D_precomp"].requires_grad),
                                       float(kw["shs"] is None), float(kw["colors_precomp"].sum())])
out["n_cases"] = np.array(len(cases))
np.savez(os.path.join(ROOT, "tests", "golden", "camera_golden.npz"), **out)
print("wrote camera_golden.npz:", len(cases), "cameras;", "rasterizer called with", list(out["0_call_kwargs"]))
