"""Generates tests/golden/tet_golden.npz from the REFERENCE's own tet math (include/utils.h:394-451), executed through
oracle/_ref/libtetmath_ref.so (built by `make -C oracle ref`; only possible where /root/reference exists).

    python tests/golden/make_tet_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import tet_oracle  # noqa: E402

assert tet_oracle.ref() is not None, "run `make -C oracle ref` first (needs /root/reference)"
rng = np.random.default_rng(7)
n = 512
tets = rng.normal(size=(n, 4, 3)).astype(np.float32)
# half the points inside (random convex weights), half anywhere
w = rng.dirichlet(np.ones(4), size=n).astype(np.float32)
inside = np.einsum("nk,nkc->nc", w, tets)
outside = rng.normal(size=(n, 3)).astype(np.float32) * 2
points = np.where((np.arange(n) % 2 == 0)[:, None], inside, outside).astype(np.float32)
# the probe of SURVEY.md F8: unit tet, p = (.1,.2,.3)
tets[0] = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32)
points[0] = np.array([0.1, 0.2, 0.3], np.float32)
np.savez(os.path.join(ROOT, "tests", "golden", "tet_golden.npz"), tets=tets, points=points,
         barys=tet_oracle.ref_barycentric(points, tets), inside=tet_oracle.ref_point_in_tet(points, tets))
print("wrote tet_golden.npz; unit tet bary =", tet_oracle.ref_barycentric(points[:1], tets[:1]))
