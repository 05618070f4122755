"""Development helper: per-kernel times of the rasterizer forward (both stages) on one workload, through the C-ABI.
    python tools/direct_probe.py [workload] [repeats]        (D3GA_LIB / D3GA_SORT_MODE / D3GA_DIRECT_* pass through)"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from d3ga_b200 import _lib, synth  # noqa: E402
from d3ga_b200.tetra_sampler import deform_gaussians  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "metric"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
lib = _lib.load()
sc = synth.make_workload(name)
cam = synth.make_camera(sc.W, sc.H, 20, 10)
d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
means, cov6 = deform_gaussians(d(sc.tetpoints), d(sc.tets), d(sc.tetra_id), d(sc.barys), d(sc.canonical_gradient),
                               d(sc.scales), d(sc.rotation))
op, col = d(sc.opacity), d(sc.colors)
view, proj, campos, bg = d(cam.viewmatrix), d(cam.projmatrix), d(cam.campos), torch.ones(3, device="cuda")
P, W, H = sc.N, sc.W, sc.H
cap = int(os.environ.get("PROBE_CAP", 8 * P))
gb, bb, ib = _lib.workspace_bytes(P, W, H, cap)
geom, binning, image = (torch.zeros(n, dtype=torch.uint8, device="cuda") for n in (gb, bb, ib))
radii = torch.zeros(P, dtype=torch.int32, device="cuda")
host = torch.zeros(2, dtype=torch.int32).pin_memory()
color = torch.zeros((3, H, W), device="cuda")
stream = torch.cuda.current_stream().cuda_stream


def once():
    _lib.check(lib.d3ga_raster_forward_geometry(P, 0, 0, means.data_ptr(), None, 1.0, None, op.data_ptr(), None,
                                                cov6.data_ptr(), col.data_ptr(), view.data_ptr(), proj.data_ptr(),
                                                campos.data_ptr(), W, H, cam.tanfovx, cam.tanfovy, 0, 0, geom.data_ptr(),
                                                binning.data_ptr(), cap, radii.data_ptr(), host.data_ptr(), stream), "geo")
    if os.environ.get("PROBE_NO_RENDER"):
        return
    _lib.check(lib.d3ga_raster_forward_render(P, W, H, bg.data_ptr(), col.data_ptr(), geom.data_ptr(), binning.data_ptr(),
                                              cap, image.data_ptr(), color.data_ptr(), None, stream), "render")


for _ in range(3):
    once()
torch.cuda.synchronize()
lib.d3ga_profile_enable(1)
for _ in range(reps):
    once()
torch.cuda.synchronize()
n = lib.d3ga_profile_kernel_count()
ms = (C.c_double * n)()
cnt = (C.c_uint64 * n)()
lib.d3ga_profile_collect(ms, cnt)
lib.d3ga_profile_enable(0)
out = {lib.d3ga_profile_kernel_name(i).decode().replace("_kernel", ""): round(ms[i] / reps * 1e3, 1) for i in range(n) if cnt[i]}
print(name, "D", int(host[0]) & 0xFFFFFFFF, "path", lib.d3ga_raster_binning_path(P, W, H, cap), out, "sum", round(sum(out.values()), 1))
