"""Drop-in for the part of ``tetra_sampler`` that D3GA uses (lib/cage.py:17: ``from tetra_sampler import Tetra,
compute_bary``), plus the fused deformation op that replaces the torch chain of models/cage_net.py:218-230.

* ``Tetra`` -- tet-mesh container with the attributes lib/cage*.py reads (``tetras, points, triangles, tetra_faces,
  triangle_to_tetra, gradient(), get_positions(), get_triangles(), n(), n_points()``), see
  submodules/tetra-sampler/tetra_sampler/tetra.py:12-130.  ``Tetra(path)`` reads a ``cage.mesh`` through meshio like
  the reference when meshio is installed; ``Tetra.from_tensors(points, tetras)`` builds one from tensors.
* ``compute_bary(points, tetras, triangles, tri_to_tetra, tetra_model)`` -> ``(barys[N,4] f32, tetra_id[N] i64,
  active[N] i32)``; backward raises NotImplementedError (tetra_sampler/barys.py:9-33).
* ``deform_gaussians(tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation)`` ->
  ``(means3D[N,3], cov3D_precomp[N,6])``, differentiable in tetpoints, barys, scales and rotation.

The ray/segment samplers and the bundled SMPL layer of the reference package are not on D3GA's render path
 and are not provided.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from d3ga_b200 import _lib


# how the deformation backward sums the cage-vertex gradients: "ordered" = a tet's contributions summed inside the warp
# before the atomics -- Gaussians walked in natural order when tetra_id is already sorted along the rows (coalesced
# reads), else through the sort permutation; "padded" = natural order, one 16-byte vector reduction per (Gaussian,
# vertex) into a padded [V,4] accumulator; "plain" = natural order, 12 scalar atomics per Gaussian.  DEFORM_PITCH4:
# "ordered" also accumulates into the padded buffer (4 vector reductions per tet instead of 12 scalar atomics).
# "auto" (measured at 1 M Gaussians / 100 k tets on B200, tools/deform_sweep.sh): rows in tet order -> "ordered" with the
# padded buffer (61 us; 67 without the padding, 87 "padded", 226 "plain"); rows in any other order -> "padded" (83 us;
# 107-120 "ordered" through the permutation, whose row reads are gathers; 169 "plain").
DEFORM_BACKWARD = os.environ.get("D3GA_DEFORM_BACKWARD", "auto")
DEFORM_PITCH4 = os.environ.get("D3GA_DEFORM_PITCH4", "0") == "1"
# The natural-order forms first select the Gaussians whose incoming gradient is not zero (d3ga_deform_backward_fused_sparse):
# a rendered view hands back exact zeros for everything that was never blended -- 86-93 % of a 1 M-Gaussian avatar per
# camera.  "0" keeps the single dense launch (faster when every Gaussian has a gradient, which no render produces).
DEFORM_SPARSE = os.environ.get("D3GA_DEFORM_SPARSE", "1") == "1"


def _check(t, name, dtype):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")   # CHECK_INPUT of include/common.h:3-7
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def unique_faces_and_topology(tetras: np.ndarray):
    """Unique sorted faces [F,3] and face -> (tet, tet | -1) map [F,2] int32 (tetra.py:45-48,88-130).

    Faces are the rows of ``unique(sort(ABCD))`` in lexicographic order like ``torch.unique(dim=0)``; the two slots of
    a face are filled in increasing tet id, which is the order the reference's dict loop produces.
    """
    t = np.asarray(tetras, dtype=np.int64)
    T = t.shape[0]
    a, b, c, d = t[:, 0], t[:, 1], t[:, 2], t[:, 3]
    faces = np.stack([np.stack([a, b, c], 1), np.stack([a, b, d], 1), np.stack([a, c, d], 1), np.stack([b, c, d], 1)], 1)
    faces = np.sort(faces.reshape(-1, 3), axis=1)
    owner = np.repeat(np.arange(T, dtype=np.int64), 4)
    uniq, inv = np.unique(faces, axis=0, return_inverse=True)
    inv = inv.reshape(-1)
    order = np.lexsort((owner, inv))   # by face, then by tet id
    inv_s, own_s = inv[order], owner[order]
    first = np.ones(inv_s.shape[0], dtype=bool)
    first[1:] = inv_s[1:] != inv_s[:-1]
    topo = -np.ones((uniq.shape[0], 2), dtype=np.int32)
    topo[inv_s[first], 0] = own_s[first]
    second = ~first
    topo[inv_s[second], 1] = own_s[second]
    return uniq.astype(np.int64), topo


class Tetra:
    def __init__(self, path=None):
        if path is None:
            return
        try:
            import meshio   # the reference's reader (tetra.py:5,15); not part of the hot path
        except ImportError as e:   # pragma: no cover
            raise ImportError("Tetra(path) needs meshio to read cage.mesh; use Tetra.from_tensors instead") from e
        mesh = meshio.read(path)
        tris = mesh.cells_dict.get("triangle")
        self._init(torch.from_numpy(np.asarray(mesh.points)).float(), torch.from_numpy(mesh.cells_dict["tetra"]).long(),
                   None if tris is None else torch.from_numpy(tris).long())

    @classmethod
    def from_tensors(cls, points, tetras, triangles=None, device=None):
        self = cls(None)
        self._init(points, tetras, triangles, device)
        return self

    def _init(self, points, tetras, triangles=None, device=None):
        if device is None:
            device = points.device if points.is_cuda else ("cuda" if torch.cuda.is_available() else "cpu")
        self.points = points.float().to(device)
        self.tetras = tetras.long().to(device)
        self.A, self.B, self.C, self.D = (self.tetras[:, k] for k in range(4))
        v0 = torch.stack([self.A, self.B, self.C], dim=1)
        v1 = torch.stack([self.A, self.B, self.D], dim=1)
        v2 = torch.stack([self.A, self.C, self.D], dim=1)
        v3 = torch.stack([self.B, self.C, self.D], dim=1)
        self.ABCD = torch.stack([v0, v1, v2, v3], dim=1)
        faces, topo = unique_faces_and_topology(self.tetras.cpu().numpy())
        self.tetra_faces = torch.from_numpy(faces).to(device)
        self.triangle_to_tetra = torch.from_numpy(topo).int().to(device)
        surf = self.triangle_to_tetra[:, 1] == -1
        self.surface_triangle = self.tetra_faces[surf]
        self.surface_tetras = self.triangle_to_tetra[surf][:, 0].long()
        self.triangles = triangles.long().to(device) if triangles is not None else self.surface_triangle

    def get_positions(self, points=None):
        return (self.points if points is None else points)[self.tetras]

    def get_triangles(self, points=None):
        return (self.points if points is None else points)[self.tetra_faces]

    def gradient(self, tetras):
        v0, v1, v2, v3 = tetras[:, 0], tetras[:, 1], tetras[:, 2], tetras[:, 3]
        return torch.stack([v3 - v0, v2 - v0, v1 - v0], dim=2)

    def n(self):
        return self.tetras.shape[0]

    def n_points(self):
        return self.points.size(0)


class BarycentricCuda(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, tetras, triangles, tri_to_tetra, tetra_model):
        lib = _lib.load()
        points = _check(points, "points", torch.float32)
        tetras = _check(tetras, "tetras", torch.float32)
        triangles = _check(triangles, "triangles", torch.float32)
        tri_to_tetra = _check(tri_to_tetra, "tri_to_tetra", torch.int32)
        n, dev = points.size(0), points.device
        with _lib.device_guard(dev):
            barys = torch.zeros(n, 4, dtype=torch.float32, device=dev)
            tetra_id = torch.zeros(n, dtype=torch.int32, device=dev)
            active_points = torch.zeros(n, dtype=torch.int32, device=dev)
            _lib.check(lib.d3ga_compute_bary(
                n, int(tetras.shape[0]), int(triangles.shape[0]), _lib.ptr(points), tetras.data_ptr(),
                triangles.data_ptr(), tri_to_tetra.data_ptr(), _lib.ptr(barys), _lib.ptr(tetra_id),
                _lib.ptr(active_points), None, _lib.raw_stream(dev)), "compute_bary")
        return barys, tetra_id.long(), active_points

    @staticmethod
    def backward(ctx, grad_output, *args, **kwargs):
        raise NotImplementedError


def compute_bary(points, tetras, triangles, tri_to_tetra, tetra_model):
    return BarycentricCuda.apply(points, tetras, triangles, tri_to_tetra, tetra_model)


# tetra_id is a constant buffer of the model (lib/cage.py:331-337): its sort permutation is computed once and reused by
# every backward (d3ga_deform_backward_ordered sums a tet's vertex gradients in the warp before the atomics).
_order_cache: dict = {}


def _tet_order(tetra_id):
    """The int32 permutation that sorts tetra_id, or True when the rows already lie in tet order (no permutation needed:
    the kernel reads them in natural order and still sums a tet's neighbours in the warp)."""
    key = (tetra_id.data_ptr(), tetra_id._version, int(tetra_id.shape[0]), tetra_id.device)
    order = _order_cache.get(key)
    if order is None:
        if tetra_id.is_cuda and torch.cuda.is_current_stream_capturing():
            return None                                        # a read-back cannot sit in a capture: the caller falls back
        if len(_order_cache) > 16:
            _order_cache.clear()
        if bool((tetra_id[1:] >= tetra_id[:-1]).all()):       # one read-back per constant buffer, before any capture
            order = True
        else:
            order = torch.argsort(tetra_id, stable=True).to(torch.int32).contiguous()
        _order_cache[key] = order
    return order


class _DeformGaussians(torch.autograd.Function):
    """One launch each way for models/cage_net.py:213-230 (csrc/deform.cu).  Inputs 0-6 are the plain op's; 7-9 the
    optional deltas, 10 the log-scales flag, 11 the optional inverse rest-pose edge matrices (then a third output, the
    per-tet cage energy of lib/cage.py:349-361, comes out of the same launch)."""

    @staticmethod
    def forward(ctx, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, delta_bary, delta_scale,
                delta_rot, log_scales, dn_inv):
        lib = _lib.load()
        f32 = torch.float32
        tetpoints = _check(tetpoints, "tetpoints", f32)
        tets = _check(tets, "tets", torch.int64)
        tetra_id = _check(tetra_id, "tetra_id", torch.int64)
        barys = _check(barys, "barys", f32)
        canonical_gradient = _check(canonical_gradient, "canonical_gradient", f32)
        scales = _check(scales, "scales", f32)
        rotation = _check(rotation, "rotation", f32)
        delta_bary = None if delta_bary is None else _check(delta_bary, "delta_bary", f32)
        delta_scale = None if delta_scale is None else _check(delta_scale, "delta_scale", f32)
        delta_rot = None if delta_rot is None else _check(delta_rot, "delta_rot", f32)
        dn_inv = None if dn_inv is None else _check(dn_inv, "dn_inv", f32)
        if delta_scale is not None and not log_scales:
            raise RuntimeError("delta_scale is added before the exp (models/cage_net.py:214): pass log_scales=True")
        N, V, T, dev = int(tetra_id.shape[0]), int(tetpoints.shape[0]), int(tets.shape[0]), tetpoints.device
        with _lib.device_guard(dev):
            means3D = torch.empty((N, 3), dtype=f32, device=dev)
            cov3D = torch.empty((N, 6), dtype=f32, device=dev)
            energy = torch.empty((T,), dtype=f32, device=dev) if dn_inv is not None else None
            _lib.check(lib.d3ga_deform_forward_fused(
                N, V, T, tetpoints.data_ptr(), tets.data_ptr(), _lib.ptr(tetra_id), _lib.ptr(barys), _lib.ptr(delta_bary),
                _lib.ptr(canonical_gradient), _lib.ptr(scales), _lib.ptr(delta_scale), int(bool(log_scales)),
                _lib.ptr(rotation), _lib.ptr(delta_rot), _lib.ptr(dn_inv), _lib.ptr(energy), _lib.ptr(means3D),
                _lib.ptr(cov3D), _lib.raw_stream(dev)), "deform_forward")
        ctx.save_for_backward(tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, delta_bary,
                              delta_scale, delta_rot, dn_inv)
        ctx.log_scales = bool(log_scales)
        ctx.set_materialize_grads(False)
        if energy is None:
            return means3D, cov3D
        return means3D, cov3D, energy

    @staticmethod
    def backward(ctx, g_means, g_cov, g_energy=None):
        lib = _lib.load()
        (tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, delta_bary, delta_scale, delta_rot,
         dn_inv) = ctx.saved_tensors
        N, V, T, dev = int(tetra_id.shape[0]), int(tetpoints.shape[0]), int(tets.shape[0]), tetpoints.device
        need = ctx.needs_input_grad
        f32 = torch.float32
        with _lib.device_guard(dev):
            g_means = torch.zeros((N, 3), dtype=f32, device=dev) if g_means is None else g_means.contiguous()
            g_cov = torch.zeros((N, 6), dtype=f32, device=dev) if g_cov is None else g_cov.contiguous()
            g_energy = None if (g_energy is None or dn_inv is None) else g_energy.contiguous()
            mode, pitch4 = DEFORM_BACKWARD, DEFORM_PITCH4
            if mode == "auto":
                mode, pitch4 = ("ordered", True) if (N > 0 and need[0] and _tet_order(tetra_id) is True) else ("padded", False)
            elif mode == "ordered" and N > 0 and need[0] and _tet_order(tetra_id) is None:
                mode = "padded"    # first sight of this tetra_id inside a graph capture: no read-back, no permutation
            pitch = 4 if (mode == "padded" or (mode == "ordered" and pitch4)) else 3
            g_tet = torch.zeros((V, pitch), dtype=f32, device=dev) if need[0] else None
            g_bary = torch.empty((N, 4), dtype=f32, device=dev) if (need[3] or need[7]) else None
            g_scales = torch.empty((N, 3), dtype=f32, device=dev) if (need[5] or need[8]) else None
            g_rot = torch.empty((N, 4), dtype=f32, device=dev) if (need[6] or need[9]) else None
            order = _tet_order(tetra_id) if (N > 0 and g_tet is not None and mode == "ordered") else None
            rows_sorted, order = (1, None) if order is True else (0, order)
            head = (N, V, T, tetpoints.data_ptr(), tets.data_ptr(), _lib.ptr(tetra_id), _lib.ptr(barys), _lib.ptr(delta_bary),
                    _lib.ptr(canonical_gradient), _lib.ptr(scales), _lib.ptr(delta_scale), int(ctx.log_scales),
                    _lib.ptr(rotation), _lib.ptr(delta_rot), _lib.ptr(dn_inv), _lib.ptr(g_means), _lib.ptr(g_cov),
                    _lib.ptr(g_energy) if g_tet is not None else None, _lib.ptr(g_tet), _lib.ptr(g_bary), _lib.ptr(g_scales),
                    _lib.ptr(g_rot))
            if DEFORM_SPARSE and order is None and N > 0 and mode != "plain":
                scratch = torch.empty(int(lib.d3ga_deform_backward_scratch_bytes(N)), dtype=torch.uint8, device=dev)
                _lib.check(lib.d3ga_deform_backward_fused_sparse(*head, rows_sorted, pitch, scratch.data_ptr(),
                                                                 _lib.raw_stream(dev)), "deform_backward_sparse")
            else:
                _lib.check(lib.d3ga_deform_backward_fused(*head, _lib.ptr(order), rows_sorted, pitch, _lib.raw_stream(dev)),
                           "deform_backward")
            if g_tet is not None and pitch == 4:
                g_tet = g_tet[:, :3]
        # a base tensor and its delta have the same gradient: one tensor serves both
        return (g_tet, None, None, g_bary if need[3] else None, None, g_scales if need[5] else None,
                g_rot if need[6] else None, g_bary if need[7] else None, g_scales if need[8] else None,
                g_rot if need[9] else None, None, None)


def deform_gaussians(tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, delta_bary=None,
                     delta_scale=None, delta_rot=None, log_scales=False, dn_inv=None):
    """Fused cage -> Gaussian deformation (models/cage_net.py:213-230).  Returns (means3D, cov3D_precomp).

    With the seven positional arguments: lines 218-230 on activated inputs (scales post-exp, rotation any non-zero
    quaternion).  The keyword arguments take lines 213-215 into the same kernel -- ``delta_bary``, ``delta_rot`` are added
    to ``barys`` / ``rotation``; ``log_scales=True`` says ``scales`` is the pre-activation parameter (exp applied inside,
    after adding ``delta_scale``) -- and ``dn_inv`` (inverse rest-pose edge matrices, lib/cage.py:312-313) makes the launch
    also evaluate the cage's neo-Hookean energy of line 225: the return value is then (means3D, cov3D_precomp,
    energy[T])."""
    return _DeformGaussians.apply(tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, delta_bary,
                                  delta_scale, delta_rot, log_scales, dn_inv)


class _FemEnergy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tetpoints, tets, dn_inv):
        lib = _lib.load()
        tetpoints = _check(tetpoints, "tetpoints", torch.float32)
        tets = _check(tets, "tets", torch.int64)
        dn_inv = _check(dn_inv, "dn_inv", torch.float32)
        T, V, dev = int(tets.shape[0]), int(tetpoints.shape[0]), tetpoints.device
        with _lib.device_guard(dev):
            energy = torch.empty((T,), dtype=torch.float32, device=dev)
            _lib.check(lib.d3ga_fem_energy_forward(T, V, tetpoints.data_ptr(), _lib.ptr(tets), _lib.ptr(dn_inv),
                                                   _lib.ptr(energy), _lib.raw_stream(dev)), "fem_energy_forward")
        ctx.save_for_backward(tetpoints, tets, dn_inv)
        return energy

    @staticmethod
    def backward(ctx, g):
        lib = _lib.load()
        tetpoints, tets, dn_inv = ctx.saved_tensors
        T, V, dev = int(tets.shape[0]), int(tetpoints.shape[0]), tetpoints.device
        with _lib.device_guard(dev):
            g_tet = torch.zeros((V, 3), dtype=torch.float32, device=dev)
            _lib.check(lib.d3ga_fem_energy_backward(T, V, tetpoints.data_ptr(), _lib.ptr(tets), _lib.ptr(dn_inv),
                                                    _lib.ptr(g.contiguous()), g_tet.data_ptr(), _lib.raw_stream(dev)),
                       "fem_energy_backward")
        return g_tet, None, None


def fem_energy(tetpoints, tets, dn_inv):
    """Per-tet stable neo-Hookean energy of CageBase.fem_energy (lib/cage.py:349-361), fused forward and backward.
    ``dn_inv`` = inverse rest-pose edge matrices ``inv(G(rest_points[tets]))`` (lib/cage.py:312-313).  Returns [T]."""
    return _FemEnergy.apply(tetpoints, tets, dn_inv)
