"""Row B of SURVEY.md section 8 pinned to the reference's own code: tests/golden/camera_golden.npz holds what the
reference's Camera (lib/cameras.py:29-75, utils/graphics_utils.py:41-75) computes and what its render()
(renderer.py:69-145) hands the rasterizer, recorded by executing both unmodified (tests/golden/make_camera_golden.py).
CPU only."""
import inspect
import os

import numpy as np

from d3ga_b200 import synth
from d3ga_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "camera_golden.npz"))


def test_make_camera_reproduces_the_reference_camera():
    for k in range(int(G["n_cases"])):
        W, H, az, el = G[f"{k}_case"]
        cam = synth.make_camera(int(W), int(H), az, el)
        R, T, fovx, fovy = synth.camera_pose(int(W), int(H), az, el)
        assert np.array_equal(R, G[f"{k}_R"]) and np.array_equal(T, G[f"{k}_T"])      # same pose went into the reference
        assert np.allclose([fovx, fovy], G[f"{k}_fov"], rtol=0, atol=0)
        assert np.abs(cam.viewmatrix - G[f"{k}_world_view_transform"]).max() < 1e-6
        assert np.abs(cam.projmatrix - G[f"{k}_full_proj_transform"]).max() < 1e-6
        assert np.abs(cam.campos - G[f"{k}_camera_center"]).max() < 1e-6
        # ... and render() passes exactly those on (renderer.py:79-93)
        assert np.array_equal(G[f"{k}_settings_viewmatrix"], G[f"{k}_world_view_transform"])
        assert np.array_equal(G[f"{k}_settings_projmatrix"], G[f"{k}_full_proj_transform"])
        assert np.array_equal(G[f"{k}_settings_campos"], G[f"{k}_camera_center"])
        h, w, tfx, tfy, mod, deg, pre, dbg, aa = G[f"{k}_settings_scalars"]
        assert (h, w) == (H, W) and abs(tfx - cam.tanfovx) < 1e-12 and abs(tfy - cam.tanfovy) < 1e-12
        assert (mod, deg, pre, dbg, aa) == (1.0, 0, 0, 0, 0)


def test_render_constructs_the_dropin_settings_and_calls_the_dropin_signature():
    """The reference's render() built THIS package's GaussianRasterizationSettings by keyword (13 fields) and called the
    rasterizer with the keyword arguments the drop-in's forward() takes; means2D is a zero tensor that requires grad."""
    fields = [str(x) for x in G["0_settings_fields"]]
    assert fields == list(GaussianRasterizationSettings._fields)
    sig = [p for p in inspect.signature(GaussianRasterizer.forward).parameters if p != "self"]
    assert sorted(sig) == [str(x) for x in G["0_call_kwargs"]]
    assert [str(x) for x in G["0_call_none"]] == ["rotations", "scales", "shs"]        # cov3D_precomp + colors_precomp path
    assert list(G["0_means2D_zero_with_grad"]) == [0.0, 1.0]
    # silhouette pass (models/trainer.py:104-110): detached geometry, explicit colours, no SH
    assert list(G["silhouette_detached"][:3]) == [0.0, 0.0, 1.0]
