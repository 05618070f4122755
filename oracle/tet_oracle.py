"""oracle/tet_oracle.py -- TEST INFRASTRUCTURE ONLY: ctypes front end of tet_oracle.c and of oracle/_ref.

``compute_bary`` restates tetra_sampler.compute_bary (submodules/tetra-sampler/tetra_sampler/barys.py:9-33);
``ref_barycentric`` / ``ref_point_in_tet`` call the REFERENCE's own header code (include/utils.h:394-451) through
oracle/_ref/libtetmath_ref.so when that file exists (built by `make -C oracle ref` in the authoring container).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
_ref = None


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle_tet.so")
        if not os.path.exists(path):
            from oracle import raster_oracle
            raster_oracle.build()
        _lib = C.CDLL(path)
    return _lib


def ref():
    """The reference-compiled library, or None when it is not present."""
    global _ref
    if _ref is None:
        path = os.path.join(_HERE, "_ref", "libtetmath_ref.so")
        if not os.path.exists(path):
            return None
        _ref = C.CDLL(path)
    return _ref


def compute_bary(points, tetras, triangles, tri_to_tetra):
    points = np.ascontiguousarray(points, np.float32)
    tetras = np.ascontiguousarray(tetras, np.float32)
    triangles = np.ascontiguousarray(triangles, np.float32)
    t2t = np.ascontiguousarray(tri_to_tetra, np.int32)
    N = points.shape[0]
    barys = np.zeros((N, 4), np.float32)
    tid = np.zeros(N, np.int32)
    active = np.zeros(N, np.int32)
    face = np.zeros(N, np.int32)
    lib().orc_compute_bary(C.c_int(N), C.c_int(tetras.shape[0]), C.c_int(triangles.shape[0]), _p(points), _p(tetras),
                           _p(triangles), _p(t2t), _p(barys), _p(tid), _p(active), _p(face))
    return barys, tid.astype(np.int64), active, face


def barycentric(points, tets):
    points = np.ascontiguousarray(points, np.float32)
    tets = np.ascontiguousarray(tets, np.float32)
    out = np.zeros((points.shape[0], 4), np.float32)
    for i in range(points.shape[0]):
        lib().orc_barycentric(_p(tets[i]), _p(points[i]), _p(out[i]))
    return out


def point_in_tet(points, tets):
    points = np.ascontiguousarray(points, np.float32)
    tets = np.ascontiguousarray(tets, np.float32)
    return np.array([lib().orc_point_in_tet(_p(tets[i]), _p(points[i])) for i in range(points.shape[0])], np.int32)


def ref_barycentric(points, tets):
    points = np.ascontiguousarray(points, np.float32)
    tets = np.ascontiguousarray(tets, np.float32)
    out = np.zeros((points.shape[0], 4), np.float32)
    ref().ref_barycentric(C.c_int(points.shape[0]), _p(points), _p(tets), _p(out))
    return out


def ref_point_in_tet(points, tets):
    points = np.ascontiguousarray(points, np.float32)
    tets = np.ascontiguousarray(tets, np.float32)
    out = np.zeros(points.shape[0], np.int32)
    ref().ref_point_in_tet(C.c_int(points.shape[0]), _p(points), _p(tets), _p(out))
    return out
