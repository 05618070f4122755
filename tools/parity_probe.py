"""Development probe: one view of a workload through the C-ABI against the f32 AND the f64 oracle (how much of the
GPU-vs-oracle difference is the float32 noise of the problem itself), plus the blend kernels' times.
    python tools/parity_probe.py metric [cam_index]        (D3GA_LIB=... selects another build of the library)"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import gpu_util  # noqa: E402
import parity_util as util  # noqa: E402
from d3ga_b200 import _lib, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "metric"
ci = int(sys.argv[2]) if len(sys.argv) > 2 else 0
sc = synth.make_workload(name)
cam = synth.camera_ring(sc.W, sc.H, 160)[ci]
means, cov6 = util.scene_geometry(sc)
g = dict(means=means, cov6=cov6, opacity=sc.opacity, colors=sc.colors)
W, H = sc.W, sc.H
w = np.random.default_rng(9).uniform(-1, 1, size=(3, H, W)).astype(np.float32)
_lib.profile_enable(True)
got = gpu_util.forward_stages(g, cam, W, H, [1, 1, 1])
gg = gpu_util.backward(got, W, H, cam, w)
prof = {k: round(v[0] / v[1] * 1e3, 1) for k, v in _lib.profile_collect().items() if v[1]}
_lib.profile_enable(False)
res = {"lib": os.environ.get("D3GA_LIB", "default"), "workload": name, "D": got["num_rendered"], "kernel_us": prof}
for dt, tag in ((np.float32, "f32"), (np.float64, "f64")):
    orc, ref = util.oracle_forward({k: v.astype(dt) for k, v in g.items()}, cam, W, H, [1, 1, 1], dtype=dt)
    rg = orc.backward(ref, w.astype(dt))
    d = np.abs(got["color"].astype(np.float64) - ref["color"].astype(np.float64))
    nc = got["n_contrib"] != ref["n_contrib"]
    off = (d > 1e-4).any(axis=0).reshape(-1)
    res[tag] = {"lists_equal": bool(np.array_equal(got["point_list"], ref["point_list"])) if got["num_rendered"] == ref["num_rendered"] else False,
                "img_beyond_1e-4": int((d > 1e-4).sum()), "img_max": float(d.max()), "n_contrib_differs": int(nc.sum()),
                "offending_pixels": int(off.sum()), "offenders_with_same_contributors": int((off & ~nc).sum()),
                "grads": {k: util.rel_l2(gg[k], rg[k]) for k in ("means3D", "cov3D", "opacity", "colors")}}
print(json.dumps(res))
