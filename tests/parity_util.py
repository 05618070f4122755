"""Shared helpers of the parity tests: seeded scenes -> oracle outputs."""
import numpy as np
import torch as th

from d3ga_b200 import synth
from oracle import deform_oracle
from oracle.raster_oracle import RasterOracle

GOLDEN = __import__("os").path.join(__import__("os").path.dirname(__file__), "golden")


def scene_geometry(sc):
    """means3D, cov6 of a synthetic scene through the deformation oracle (CPU torch)."""
    t = th.from_numpy
    means, cov6 = deform_oracle.deform_gaussians(t(sc.tetpoints), t(sc.tets), t(sc.tetra_id), t(sc.barys),
                                                 t(sc.canonical_gradient), t(sc.scales), t(sc.rotation))
    return means.numpy(), cov6.numpy()


def random_gaussians(P, rng, spread=0.3, sigma=0.03, with_behind=True):
    means = rng.uniform(-spread, spread, size=(P, 3)).astype(np.float32)
    if with_behind and P >= 8:
        means[::7, 2] += 10.0   # behind a camera that sits at +z looking down -z
    scales = np.exp(rng.normal(np.log(sigma), 0.5, size=(P, 3))).astype(np.float32)
    q = rng.normal(size=(P, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    A = rng.normal(size=(P, 3, 3)) * sigma
    cov = A @ A.transpose(0, 2, 1) + 1e-6 * np.eye(3)
    cov6 = np.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], 1).astype(np.float32)
    op = rng.uniform(0.05, 0.95, size=(P, 1)).astype(np.float32)
    col = rng.uniform(0, 1, size=(P, 3)).astype(np.float32)
    shs = rng.normal(0, 0.3, size=(P, 16, 3)).astype(np.float32)
    return dict(means=means, scales=scales, rot=q, cov6=cov6, opacity=op, colors=col, shs=shs)


def oracle_forward(g, cam, W, H, bg, path="precomp", aa=False, dtype=np.float32, **kw):
    orc = RasterOracle(dtype)
    args = dict(means3D=g["means"], opacities=g["opacity"], viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix,
                campos=cam.campos, W=W, H=H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg, antialiasing=aa)
    if path == "precomp":
        args.update(colors_precomp=g["colors"], cov3D_precomp=g["cov6"])
    else:
        args.update(shs=g["shs"], sh_degree=3, scales=g["scales"], rotations=g["rot"], scale_modifier=1.0)
    args.update(kw)
    return orc, orc.forward(**args)


def rel_l2(a, b):
    a = np.asarray(a, np.float64).ravel()
    b = np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(1e-30, np.linalg.norm(b)))


def oracle_view(scene, cam, weights, bg=(1.0, 1.0, 1.0), raster_inputs=None):
    """One whole view on the CPU -- deformation oracle -> rasterizer oracle -> loss = sum(image * weights) -> both
    backward passes -- returning (loss, {parameter: gradient}) for the six hot-path parameters.

    raster_inputs = (means3D, cov3D) evaluates the RASTERIZER oracle at those values instead of the deformation oracle's
    own (the chain rule still runs through the deformation oracle).  The rasterizer is a discontinuous function of its
    inputs -- depth order, the 1/255 cut, the transmittance stop -- so two float32 deformations that differ in the last
    bit give images and gradients that differ by ~5e-4 relative L2 at 1 M Gaussians (the float32 and float64 oracles
    differ from each other by that much); feeding the oracle the values the product path actually rasterized removes
    that noise from the comparison, and the deformation itself is compared separately to ~1e-6."""
    t = th.from_numpy
    leaves = {k: t(np.ascontiguousarray(getattr(scene, k))).clone().requires_grad_(True)
              for k in ("tetpoints", "barys", "rotation")}
    leaves["scaling"] = t(np.log(scene.scales)).clone().requires_grad_(True)      # pre-activation, like training.Avatar
    means, cov6 = deform_oracle.deform_gaussians(leaves["tetpoints"], t(scene.tets), t(scene.tetra_id), leaves["barys"],
                                                 t(scene.canonical_gradient), th.exp(leaves["scaling"]), leaves["rotation"])
    m_in, c_in = (means.detach().numpy(), cov6.detach().numpy()) if raster_inputs is None else raster_inputs
    orc = RasterOracle(np.float32)
    st = orc.forward(np.ascontiguousarray(m_in), scene.opacity, cam.viewmatrix, cam.projmatrix, cam.campos, scene.W, scene.H,
                     cam.tanfovx, cam.tanfovy, list(bg), colors_precomp=scene.colors, cov3D_precomp=np.ascontiguousarray(c_in))
    g = orc.backward(st, weights)
    th.autograd.backward([means, cov6], [t(g["means3D"]), t(g["cov3D"])])
    grads = {k: v.grad.numpy().astype(np.float64) for k, v in leaves.items()}
    grads["opacity"] = np.asarray(g["opacity"], np.float64).reshape(scene.opacity.shape)
    grads["colors"] = np.asarray(g["colors"], np.float64)
    loss = float((st["color"].astype(np.float64) * weights.astype(np.float64)).sum())
    return loss, grads
