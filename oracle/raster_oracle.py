"""oracle/raster_oracle.py -- TEST INFRASTRUCTURE ONLY.

numpy/ctypes front end of ``raster_oracle.c`` (the CPU restatement of the rasterizer D3GA calls at
renderer.py:79-93,130-141).  PARITY UNPINNED: the upstream module's source is absent from /root/reference
(.SUBMODULES.json:16-22) and the reference holds no golden vectors for this path; see the C file's header.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import
this module.  The product package (``d3ga_b200``) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {}


def build(force: bool = False) -> None:
    """Compile the C restatement (gcc only).  Called by ``__graft_entry__.build()`` and lazily by the loader."""
    need = force or not all(
        os.path.exists(os.path.join(_HERE, n))
        for n in ("liboracle_raster_f32.so", "liboracle_raster_f64.so", "liboracle_tet.so")
    )
    if need:
        subprocess.check_call(["make", "-s", "-C", _HERE, "all"])


def _lib(dtype):
    key = np.dtype(dtype).name
    if key not in _LIBS:
        name = {"float32": "liboracle_raster_f32.so", "float64": "liboracle_raster_f64.so"}[key]
        path = os.path.join(_HERE, name)
        if not os.path.exists(path):
            build()
        lib = C.CDLL(path)
        assert lib.orc_sizeof_real() == np.dtype(dtype).itemsize
        lib.orc_inclusive_scan.restype = C.c_uint32
        lib.orc_sort_bits.restype = C.c_int
        _LIBS[key] = lib
    return _LIBS[key]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class RasterOracle:
    """Stage-by-stage CPU rasterizer.  ``dtype=float64`` exists only for finite-difference validation."""

    def __init__(self, dtype=np.float32):
        self.dt = np.dtype(dtype)
        self.lib = _lib(dtype)
        self.creal = C.c_float if self.dt == np.float32 else C.c_double

    def _a(self, x, shape=None):
        if x is None:
            return None
        a = np.ascontiguousarray(np.asarray(x, dtype=self.dt))
        if shape is not None:
            a = a.reshape(shape)
        return a

    # ------------------------------------------------------------------------------------------------ forward
    def forward(self, means3D, opacities, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, bg,
                colors_precomp=None, shs=None, sh_degree=0, scales=None, rotations=None, cov3D_precomp=None,
                scale_modifier=1.0, antialiasing=False, stop_after=None):
        lib, r = self.lib, self.creal
        means = self._a(means3D, (-1, 3))
        P = means.shape[0]
        opac = self._a(opacities, (-1,))
        view, proj, cam = self._a(viewmatrix, (16,)), self._a(projmatrix, (16,)), self._a(campos, (3,))
        colors = self._a(colors_precomp, (P, 3)) if colors_precomp is not None else None
        sh = self._a(shs) if shs is not None else None
        M = 0 if sh is None else sh.reshape(P, -1, 3).shape[1]
        sc = self._a(scales, (P, 3)) if scales is not None else None
        rot = self._a(rotations, (P, 4)) if rotations is not None else None
        cov_pre = self._a(cov3D_precomp, (P, 6)) if cov3D_precomp is not None else None
        assert (colors is None) != (sh is None)
        assert (cov_pre is None) != (sc is None or rot is None)
        st = dict(P=P, W=W, H=H, M=M, deg=sh_degree, means=means, opac=opac, view=view, proj=proj, cam=cam,
                  colors=colors, shs=sh, scales=sc, rots=rot, mod=float(scale_modifier), aa=int(bool(antialiasing)),
                  tanx=float(tanfovx), tany=float(tanfovy), bg=self._a(bg, (3,)))
        st["radii"] = np.zeros(P, np.int32)
        st["xy"] = np.zeros((P, 2), self.dt)
        st["depths"] = np.zeros(P, self.dt)
        st["cov3D"] = np.zeros((P, 6), self.dt)
        st["rgb"] = np.zeros((P, 3), self.dt)
        st["conic_opacity"] = np.zeros((P, 4), self.dt)
        st["tiles_touched"] = np.zeros(P, np.uint32)
        st["clamped"] = np.zeros((P, 3), np.uint8)
        lib.orc_preprocess(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means), _p(sc), r(scale_modifier), _p(rot),
                           _p(opac), _p(sh), _p(cov_pre), _p(colors), _p(view), _p(proj), _p(cam), C.c_int(W),
                           C.c_int(H), r(tanfovx), r(tanfovy), C.c_int(st["aa"]), _p(st["radii"]), _p(st["xy"]),
                           _p(st["depths"]), _p(st["cov3D"]), _p(st["rgb"]), _p(st["conic_opacity"]),
                           _p(st["tiles_touched"]), _p(st["clamped"]))
        if stop_after == "preprocess":
            return st
        st["offsets"] = np.zeros(P, np.uint32)
        D = int(lib.orc_inclusive_scan(C.c_int(P), _p(st["tiles_touched"]), _p(st["offsets"])))
        st["num_rendered"] = D
        st["keys_unsorted"] = np.zeros(D, np.uint64)
        st["vals_unsorted"] = np.zeros(D, np.uint32)
        lib.orc_duplicate_keys(C.c_int(P), _p(st["xy"]), _p(st["depths"]), _p(st["offsets"]), _p(st["radii"]),
                               C.c_int(W), C.c_int(H), _p(st["keys_unsorted"]), _p(st["vals_unsorted"]))
        st["sort_bits"] = int(lib.orc_sort_bits(C.c_int(W), C.c_int(H)))
        st["keys"] = np.zeros(D, np.uint64)
        st["point_list"] = np.zeros(D, np.uint32)
        lib.orc_sort_pairs(C.c_uint64(D), _p(st["keys_unsorted"]), _p(st["vals_unsorted"]), _p(st["keys"]),
                           _p(st["point_list"]), C.c_int(st["sort_bits"]))
        ntiles = ((W + 15) // 16) * ((H + 15) // 16)
        st["ranges"] = np.zeros((ntiles, 2), np.uint32)
        lib.orc_tile_ranges(C.c_uint64(D), _p(st["keys"]), C.c_int(ntiles), _p(st["ranges"]))
        if stop_after == "binning":
            return st
        st["final_T"] = np.zeros(H * W, self.dt)
        st["n_contrib"] = np.zeros(H * W, np.uint32)
        st["color"] = np.zeros((3, H, W), self.dt)
        st["invdepth"] = np.zeros((1, H, W), self.dt)
        lib.orc_render_forward(C.c_int(W), C.c_int(H), _p(st["ranges"]), _p(st["point_list"]), _p(st["xy"]),
                               _p(st["rgb"]), _p(st["conic_opacity"]), _p(st["depths"]), _p(st["bg"]),
                               _p(st["final_T"]), _p(st["n_contrib"]), _p(st["color"]), _p(st["invdepth"]))
        return st

    # ----------------------------------------------------------------------------------------------- backward
    def backward(self, st, dL_dcolor_img, dL_dinvdepth_img=None):
        lib, r = self.lib, self.creal
        P, W, H, M = st["P"], st["W"], st["H"], st["M"]
        dpix = self._a(dL_dcolor_img, (3, H, W))
        dinv = self._a(dL_dinvdepth_img, (H * W,)) if dL_dinvdepth_img is not None else None
        g = dict(mean2D=np.zeros((P, 3), self.dt), conic=np.zeros((P, 4), self.dt), opacity=np.zeros(P, self.dt),
                 colors=np.zeros((P, 3), self.dt), invdepths=np.zeros(P, self.dt))
        lib.orc_render_backward(C.c_int(W), C.c_int(H), _p(st["ranges"]), _p(st["point_list"]), _p(st["bg"]),
                                _p(st["xy"]), _p(st["conic_opacity"]), _p(st["rgb"]), _p(st["depths"]),
                                _p(st["final_T"]), _p(st["n_contrib"]), _p(dpix), _p(dinv), _p(g["mean2D"]),
                                _p(g["conic"]), _p(g["opacity"]), _p(g["colors"]), _p(g["invdepths"]))
        g["means3D"] = np.zeros((P, 3), self.dt)
        g["cov3D"] = np.zeros((P, 6), self.dt)
        g["sh"] = np.zeros((P, M, 3), self.dt) if st["shs"] is not None else None
        g["scales"] = np.zeros((P, 3), self.dt) if st["scales"] is not None else None
        g["rotations"] = np.zeros((P, 4), self.dt) if st["rots"] is not None else None
        lib.orc_preprocess_backward(C.c_int(P), C.c_int(st["deg"]), C.c_int(M), _p(st["means"]), _p(st["radii"]),
                                    _p(st["shs"]), _p(st["clamped"]), _p(st["scales"]), _p(st["rots"]), r(st["mod"]),
                                    _p(st["cov3D"]), _p(st["view"]), _p(st["proj"]), _p(st["cam"]), C.c_int(W),
                                    C.c_int(H), r(st["tanx"]), r(st["tany"]), C.c_int(st["aa"]), _p(st["opac"]),
                                    _p(g["mean2D"]), _p(g["conic"]), _p(g["invdepths"]) if dinv is not None else None,
                                    _p(g["opacity"]), _p(g["colors"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]),
                                    _p(g["scales"]), _p(g["rotations"]))
        return g
