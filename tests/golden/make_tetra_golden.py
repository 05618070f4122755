"""Generates tests/golden/tetra_golden.npz by executing the REFERENCE's own tetra_sampler.Tetra.__init__ and .build()
(submodules/tetra-sampler/tetra_sampler/tetra.py:12-130) on CPU for small lattice cages.

The class reads a Medit file through meshio and wraps the surface in a trimesh; both packages are absent here and are
replaced by minimal stand-ins that hand the constructor the arrays a file would have held (the constructor's own code
-- face extraction, torch.unique, the python-dict topology loop, the surface split -- runs unmodified, apart from two
environment shims: `.cuda()` is a no-op and numpy 2 no longer has ndarray.tostring(), so the module source is loaded
with `.tostring()` spelled `.tobytes()`, its documented equivalent).

    python tests/golden/make_tetra_golden.py      (needs /root/reference)
"""
import os
import sys
import tempfile
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.environ.get("D3GA_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
from d3ga_b200 import synth  # noqa: E402

torch.Tensor.cuda = lambda self, *a, **k: self
torch.cuda.empty_cache = lambda: None
current = {}
meshio = types.ModuleType("meshio")
meshio.read = lambda path: types.SimpleNamespace(cells_dict=current["cells"], points=current["points"])
trimesh = types.ModuleType("trimesh")


class _Trimesh:
    def __init__(self, vertices=None, faces=None, process=False):
        self.vertices, self.faces = np.asarray(vertices), np.asarray(faces)


trimesh.Trimesh = _Trimesh
trimesh.repair = types.SimpleNamespace(fix_normals=lambda mesh: None)
sys.modules["meshio"], sys.modules["trimesh"] = meshio, trimesh

src_path = os.path.join(REF, "submodules", "tetra-sampler", "tetra_sampler", "tetra.py")
src = open(src_path).read().replace(".tostring()", ".tobytes()")
mod = types.ModuleType("ref_tetra")
exec(compile(src, src_path, "exec"), mod.__dict__)

out = {}
cases = [(2, 2, 2), (3, 4, 2), (1, 1, 1)]
for k, (nx, ny, nz) in enumerate(cases):
    pts, tets = synth.make_cage(nx, ny, nz)
    rng = np.random.default_rng(k)
    tets = tets[rng.permutation(tets.shape[0])]                      # arbitrary tet order
    tets = np.stack([rng.permutation(row) for row in tets])          # arbitrary vertex order inside a tet
    current["points"] = pts.astype(np.float64)
    current["cells"] = {"tetra": tets, "triangle": np.zeros((1, 3), np.int64)}
    with tempfile.TemporaryDirectory() as d:
        t = mod.Tetra(os.path.join(d, "cage.mesh"))
    out[f"{k}_points"], out[f"{k}_tetras"] = pts, tets
    out[f"{k}_tetra_faces"] = t.tetra_faces.numpy()
    out[f"{k}_triangle_to_tetra"] = t.triangle_to_tetra.numpy()
    out[f"{k}_surface_triangle"] = t.surface_triangle.numpy()
    out[f"{k}_surface_tetras"] = t.surface_tetras.numpy()
    X = torch.from_numpy(pts)[torch.from_numpy(tets)]
    out[f"{k}_gradient"] = t.gradient(X).numpy()
    out[f"{k}_n"] = np.array([t.n(), t.n_points()])
out["n_cases"] = np.array(len(cases))
np.savez(os.path.join(ROOT, "tests", "golden", "tetra_golden.npz"), **out)
print("wrote tetra_golden.npz", {k: out[f"{k}_tetra_faces"].shape for k in range(len(cases))})
