"""oracle/raster_torch.py -- TEST INFRASTRUCTURE ONLY: a second, independently derived backward for rows R6-R8.

oracle/raster_oracle.c restates upstream's ANALYTIC backward (renderCUDA / computeCov2DCUDA / preprocessCUDA backward of
graphdeco-inria/diff-gaussian-rasterization, absent from the reference tree).  Its only check so far was finite
differences.  This file writes the FORWARD in differentiable PyTorch (float64, CPU) -- projection, EWA covariance, conic,
per-pixel front-to-back blending -- and lets torch.autograd derive the backward; tests/test_oracle_raster.py compares the
two.  Small scenes only (every pixel against every Gaussian: O(P N) memory).

What is taken from the C oracle's forward, as constants: the discrete decisions that carry no gradient -- which Gaussians
are in which tile's list and in which order (radius / rect / depth sort: `ranges`, `point_list`).  Everything with a
gradient is recomputed here.  Conventions that make the autograd gradient the one upstream defines:
  * the 0.99 clamp is straight-through (upstream's backward uses dL/dG = o * dL/dalpha whether or not alpha was clamped);
  * the skip tests (power > 0, alpha < 1/255) and the stop test (T (1 - alpha) < 1e-4) are decisions, not functions;
  * the frustum clamp of t.x / t.z is an ordinary clamp (zero gradient outside: upstream's x_grad_mul / y_grad_mul).
Follows SURVEY.md Appendix B; call site in the reference: renderer.py:130-141.
"""
from __future__ import annotations

import numpy as np
import torch as th


def render(means3D, cov3D, opacity, colors, viewmatrix, projmatrix, W, H, tanfovx, tanfovy, bg, ranges, point_list):
    """means3D [N,3], cov3D [N,6], opacity [N], colors [N,3]: float64 tensors (leaves).  Returns the image [3,H,W]."""
    dt = th.float64
    V = th.as_tensor(np.asarray(viewmatrix, np.float64)).reshape(4, 4)      # row-vector convention: p_view = [p,1] V
    Pm = th.as_tensor(np.asarray(projmatrix, np.float64)).reshape(4, 4)
    N = means3D.shape[0]
    hom = th.cat([means3D, th.ones(N, 1, dtype=dt)], 1)
    t = hom @ V                                                              # view space
    h = hom @ Pm
    pw = 1.0 / (h[:, 3] + 0.0000001)
    ndc = h[:, :2] * pw[:, None]
    px = ((ndc[:, 0] + 1.0) * W - 1.0) * 0.5
    py = ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5
    fx, fy = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    tz = t[:, 2]
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    tx = th.clamp(t[:, 0] / tz, -limx, limx) * tz
    ty = th.clamp(t[:, 1] / tz, -limy, limy) * tz
    zero = th.zeros_like(tz)
    J = th.stack([th.stack([fx / tz, zero, -(fx * tx) / (tz * tz)], 1),
                  th.stack([zero, fy / tz, -(fy * ty) / (tz * tz)], 1)], 1)   # [N,2,3]
    Wrot = V[:3, :3].T                                                       # world -> view rotation (column vectors)
    M = J @ Wrot                                                             # [N,2,3]
    S = th.stack([th.stack([cov3D[:, 0], cov3D[:, 1], cov3D[:, 2]], 1),
                  th.stack([cov3D[:, 1], cov3D[:, 3], cov3D[:, 4]], 1),
                  th.stack([cov3D[:, 2], cov3D[:, 4], cov3D[:, 5]], 1)], 1)
    c2 = M @ S @ M.transpose(1, 2)
    cxx, cxy, cyy = c2[:, 0, 0] + 0.3, c2[:, 0, 1], c2[:, 1, 1] + 0.3
    det = cxx * cyy - cxy * cxy
    con_a, con_b, con_c = cyy / det, -cxy / det, cxx / det

    gx = (W + 15) // 16
    ys, xs = np.mgrid[0:H, 0:W]
    tile_of_pixel = (ys // 16) * gx + xs // 16
    bgt = th.as_tensor(np.asarray(bg, np.float64))
    out = th.zeros(3, H, W, dtype=dt)
    for tile in np.unique(tile_of_pixel):
        r0, r1 = int(ranges[tile, 0]), int(ranges[tile, 1])
        sel = tile_of_pixel == tile
        pyx = np.stack([ys[sel], xs[sel]], 1)
        if r1 <= r0:
            out[:, pyx[:, 0], pyx[:, 1]] = bgt[:, None].expand(3, pyx.shape[0])
            continue
        ids = th.as_tensor(point_list[r0:r1].astype(np.int64))
        pxf = th.as_tensor(pyx[:, 1].astype(np.float64))[:, None]
        pyf = th.as_tensor(pyx[:, 0].astype(np.float64))[:, None]
        dx, dy = px[ids][None, :] - pxf, py[ids][None, :] - pyf               # [pixels, entries]
        power = -0.5 * (con_a[ids] * dx * dx + con_c[ids] * dy * dy) - con_b[ids] * dx * dy
        raw = opacity[ids][None, :] * th.exp(power)
        alpha = raw + (th.clamp(raw, max=0.99) - raw).detach()               # straight-through clamp
        valid = (power <= 0) & (alpha >= 1.0 / 255.0)
        a = th.where(valid, alpha, th.zeros_like(alpha))
        one_minus = 1.0 - a
        T_before = th.cumprod(th.cat([th.ones(a.shape[0], 1, dtype=dt), one_minus[:, :-1]], 1), 1)
        test = (T_before * one_minus).detach()
        stop = valid & (test < 0.0001)
        stopped = th.cumsum(stop.to(th.int64), 1) > 0                        # this entry or an earlier one stopped
        contrib = valid & ~stopped
        w = th.where(contrib, a * T_before, th.zeros_like(a))
        # transmittance left when the pixel is finished: the product over the contributing entries only
        T_final = th.prod(th.where(contrib, one_minus, th.ones_like(one_minus)), 1)
        col = w @ colors[ids]                                                # [pixels, 3]
        res = col + T_final[:, None] * bgt[None, :]
        out[:, pyx[:, 0], pyx[:, 1]] = res.T
    return out
