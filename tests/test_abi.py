"""CPU tests of the boundary: the C-ABI library loads without a GPU and exports every symbol include/*.h declares;
the Python drop-in packages keep the reference's interface (names, fields, error behaviour)."""
import glob
import inspect
import os
import re

import pytest
import torch

from d3ga_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names += re.findall(r"\b(d3ga_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()           # raises if the .so has not been built: there is no fallback
    decl = declared_symbols()
    assert len(decl) >= 13
    for name in decl:
        assert hasattr(lib, name), name
    assert sorted(_lib.EXPORTS) == decl
    assert lib.d3ga_version() >= 100
    assert lib.d3ga_error_string(-1) == b"invalid argument"


def test_size_queries_need_no_gpu():
    g, b, i = _lib.workspace_bytes(1000, 64, 48, 5000)
    assert g > 1000 * (4 + 8 + 16) and b >= 5000 * 24 and i >= 64 * 48 * 8
    lay = _lib.raster_layout(1000, 64, 48, 5000)
    assert lay["geom_header"] == 0 and lay["bin_keys1"] - lay["bin_keys0"] >= 5000 * 8
    with pytest.raises(_lib.D3GAError):
        _lib.workspace_bytes(10, 64, 48, 1 << 32)


def test_binning_path_choice_needs_no_gpu():
    """The direct tile scatter is the default wherever its scratch fits the binning workspace; the choice is a pure
    function of the sizes (both forward stages must agree on it)."""
    lib = _lib.load()
    if os.environ.get("D3GA_SORT_MODE", "direct") != "direct":
        pytest.skip("a radix mode is forced")
    assert lib.d3ga_raster_binning_path(1_000_000, 1024, 1024, 8_000_000) == 1     # the metric workload
    assert lib.d3ga_raster_binning_path(1_000_000, 2048, 1334, 16_000_000) == 1    # c3: bands of tile rows
    assert lib.d3ga_raster_binning_path(4_000_000, 3840, 2160, 1_600_000_000) == 1   # c4
    assert lib.d3ga_raster_binning_path(1000, 64, 48, 5000) == 1
    assert lib.d3ga_raster_binning_path(0, 64, 48, 5000) == 0                      # nothing to bin
    assert lib.d3ga_raster_binning_path(1_000_000, 1024, 1024, 100) == 0           # count rows would not fit: radix path
    assert lib.d3ga_raster_binning_path(1000, 16 * 4096, 16, 1 << 20) == 0         # 4096 tiles in one row: strip width limit
    assert lib.d3ga_raster_binning_path(1000, 8192, 8192, 1 << 24) == 0            # 262144 tiles: beyond the plan's shared memory
    assert lib.d3ga_raster_binning_path(-1, 64, 48, 5000) < 0


def test_settings_and_rasterizer_interface():
    from d3ga_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

    # the 13 fields renderer.py:79-93 passes by keyword
    fields = ["image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
              "sh_degree", "campos", "prefiltered", "debug", "antialiasing"]
    assert list(GaussianRasterizationSettings._fields) == fields
    rs = GaussianRasterizationSettings(image_height=8, image_width=8, tanfovx=1.0, tanfovy=1.0, bg=torch.zeros(3),
                                       scale_modifier=1.0, viewmatrix=torch.eye(4), projmatrix=torch.eye(4),
                                       sh_degree=0, campos=torch.zeros(3), prefiltered=False, debug=False,
                                       antialiasing=False)
    r = GaussianRasterizer(raster_settings=rs)
    sig = list(inspect.signature(r.forward).parameters)
    assert sig == ["means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"]
    x = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=x[:, :1], cov3D_precomp=torch.zeros(4, 6))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=x, means2D=x, opacities=x[:, :1], colors_precomp=x)
    with pytest.raises(RuntimeError, match="CUDA tensor"):   # CHECK_INPUT behaviour of the reference extensions
        r(means3D=x, means2D=x, opacities=x[:, :1], colors_precomp=x, cov3D_precomp=torch.zeros(4, 6))


def test_tetra_interface():
    from d3ga_b200 import synth
    from d3ga_b200.tetra_sampler import Tetra, compute_bary

    pts, tets = synth.make_cage(2, 2, 1)
    cage = Tetra.from_tensors(torch.from_numpy(pts), torch.from_numpy(tets), device="cpu")
    assert cage.n() == tets.shape[0] and cage.n_points() == pts.shape[0]
    assert cage.tetra_faces.shape[1] == 3 and cage.triangle_to_tetra.dtype == torch.int32
    X = cage.get_positions()
    Gm = cage.gradient(X)
    assert Gm.shape == (tets.shape[0], 3, 3)
    assert torch.allclose(Gm[:, :, 0], X[:, 3] - X[:, 0])   # edges as columns (tetra.py:74-80)
    assert cage.get_triangles().shape == (cage.tetra_faces.shape[0], 3, 3)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        compute_bary(torch.zeros(3, 3), X, cage.get_triangles(), cage.triangle_to_tetra, cage)
