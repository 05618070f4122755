#!/usr/bin/env python
"""bench.py -- fwd+bwd views/s of the D3GA render hot path on B200 (the metric of BASELINE.json).

    python bench.py --gpus N --steps K --warmup W [--workload metric|c1|c2|c3|c4] [--views V]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...          # the CPU restatement of the reference path on the host cores

A "step" is one multi-view training step on the hot path: every rank runs V camera views (cage deformation ->
rasterizer forward -> loss -> rasterizer backward -> deformation backward, gradients accumulated) on replicated
parameters, then ONE all-reduce of the flat gradient buffer (no collective when N = 1).  Cameras are sharded over
ranks (camera c -> rank c mod N), per-rank work is fixed => weak scaling.  value = views of all ranks / max-over-ranks
device time.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# stdout carries ONE JSON line: NCCL's version banner / warnings go to stderr
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
if "reference" in sys.argv:
    # the CPU arm uses every host thread it can (torchrun exports OMP_NUM_THREADS=1 for its workers); must be set
    # before numpy / torch / the oracle's libgomp are loaded
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)

import numpy as np  # noqa: E402

METRIC = "fwd+bwd views/s"
UNIT = "views/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="metric")
    ap.add_argument("--views", type=int, default=20, help="camera views per rank per step (C5: 160 cameras / 8 GPUs)")
    ap.add_argument("--cameras", type=int, default=160)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="eager per-view launches instead of the captured CUDA graph")
    ap.add_argument("--lanes", type=int, default=4, help="views in flight on separate streams (graph replay only)")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


# ---- clocks ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc, self.t_mark = gpu_index, [], None, 0.0

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")] + [time.perf_counter()])

    def mark(self):
        """The timed region starts now.  (The sampler is started before the >= 1 s pre-warm of the same work, so that
        nvidia-smi is already up; only samples from here on -- plus the one just before, taken under the same load --
        are reported.)"""
        self.t_mark = time.perf_counter() - 0.03

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for r in self.rows if r[-1] >= self.t_mark]
        sm = [float(r[1]) for r in rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in rows:
            if len(r) >= 9:
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---- the CPU restatement of the reference path (cpu_baseline and --impl reference) -----------------------------------------
def cpu_view(scene, cam, weights):
    """One view fwd+bwd on the host: torch CPU deformation (the reference's own form, models/cage_net.py:213-230)
    + the C rasterizer oracle (OpenMP).  Test infrastructure timed as the reported CPU baseline, never shipped."""
    import torch as th

    from oracle import deform_oracle
    from oracle.raster_oracle import RasterOracle

    t = th.from_numpy
    leaves = [t(a).clone().requires_grad_(True) for a in (scene.tetpoints, scene.barys, np.log(scene.scales), scene.rotation)]
    means, cov6 = deform_oracle.deform_gaussians(leaves[0], t(scene.tets), t(scene.tetra_id), leaves[1],
                                                 t(scene.canonical_gradient), th.exp(leaves[2]), leaves[3])
    orc = RasterOracle(np.float32)
    st = orc.forward(means.detach().numpy(), scene.opacity, cam.viewmatrix, cam.projmatrix, cam.campos, scene.W,
                     scene.H, cam.tanfovx, cam.tanfovy, [1, 1, 1], colors_precomp=scene.colors,
                     cov3D_precomp=cov6.detach().numpy())
    g = orc.backward(st, weights)
    th.autograd.backward([means, cov6], [t(g["means3D"]), t(g["cov3D"])])
    return st


def time_cpu_views(scene, cams, weights, n_views, warm=0):
    import torch as th

    th.set_num_threads(os.cpu_count() or 1)
    for i in range(warm):
        cpu_view(scene, cams[i % len(cams)], weights)
    t0 = time.perf_counter()
    for i in range(n_views):
        cpu_view(scene, cams[i % len(cams)], weights)
    dt = time.perf_counter() - t0
    return n_views / dt, dt


def run_reference(args):
    """`--impl reference`: the reference path's CPU implementation on the host cores, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from d3ga_b200 import synth
    from oracle import raster_oracle

    raster_oracle.build()
    scene = synth.make_workload(args.workload)
    cams = synth.camera_ring(scene.W, scene.H, min(args.cameras, 16))
    weights = np.random.default_rng(0).uniform(-1, 1, (3, scene.H, scene.W)).astype(np.float32)
    cores = os.cpu_count() or 1
    for i in range(args.warmup):
        cpu_view(scene, cams[i % len(cams)], weights)
    t0 = time.perf_counter()
    for i in range(args.steps):     # each step = a bounded sample of the workload: ONE camera view
        cpu_view(scene, cams[i % len(cams)], weights)
    dt = time.perf_counter() - t0
    v = args.steps / dt
    sample = "1 camera view of the same scene per step (fwd+bwd), all host threads"
    line = {"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(1, args.steps), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": workload_config(scene, args, 1, impl="reference"),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(scene, args, views, impl="ours"):
    base = {"workload": f"{scene.name}: {scene.N} Gaussians in {scene.tets.shape[0]} tets over {len(scene.meta['garments'])} "
                        f"cage(s), {scene.W}x{scene.H}, cov3D_precomp + colors_precomp path",
            "views_per_rank_per_step": views, "cameras": args.cameras}
    if impl == "reference":
        base.update(parallelism="rank 0 only, all host threads (OpenMP rasterizer oracle + torch CPU deformation)",
                    launch="CPU: one camera view (forward + backward) per step, no GPU work")
        return base
    return {**base, "parallelism": f"camera-sharded dp{args.gpus}",
            "launch": "eager" if args.no_graph else f"one CUDA-graph replay per view, {max(1, args.lanes)} view(s) in flight on separate streams",
            "l2": "working set (>= 150 MB of Gaussian state + binning buffers per view) exceeds the 126 MB L2; no flush",
            "binning": "direct tile scatter; inside the replayed graphs its rows are sized for the views in flight "
                       "(one view's scatter takes ~1/n of the GPU for n times as long)"}


# ---- the GPU arm -------------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch

    from d3ga_b200 import _lib, dist as ddist, synth, training

    rank, local, world = ddist.init_from_env()
    assert torch.cuda.is_available(), "bench.py (impl ours) needs a GPU; there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _lib.load()

    scene = synth.make_workload(args.workload)
    cams = synth.camera_ring(scene.W, scene.H, args.cameras)
    mine = ddist.shard_cameras(len(cams), rank, world)
    V = args.views
    avatar = training.Avatar(scene, dev)
    bank = training.CameraBank(cams, dev)
    bg = torch.ones(3, device=dev)
    rng = np.random.default_rng(0)
    # the target / weight image of the loss lives on the host as uint8 (what a camera frame is: lib/batch.py:86-88 uploads
    # uint8 and normalises on the GPU); the resident copy is the same image, normalised
    weights_u8 = torch.from_numpy(rng.integers(0, 256, (3, scene.H, scene.W), dtype=np.uint8)).pin_memory()
    weights_h = (weights_u8.float() * training.U8_SCALE + training.U8_OFFSET).pin_memory()
    weights = weights_h.to(dev)
    loss_h = torch.zeros((), dtype=torch.float32).pin_memory()

    def cam_ids(step):
        return [mine[(step * V + v) % len(mine)] for v in range(V)]

    # one view = one replay of the captured graph (training.GraphedViewStep); --no-graph launches every view eagerly
    gv = pv = None
    lanes = 1
    if not args.no_graph:
        training.multi_view_step(avatar, bank, [mine[0]], weights, bg)     # one eager view: settles the capacity hint
        lanes = training.PipelinedViews.lanes_that_fit(avatar, args.lanes)
        if lanes > 1:
            pv = training.PipelinedViews(avatar, cams[0], bg, lanes=lanes)
        else:
            gv = training.GraphedViewStep(avatar, cams[0], bg)
    if pv is not None or gv is not None:
        # capture now, so that a failure (memory, an unsupported driver) degrades to eager launches instead of
        # costing the whole run; said in config.launch
        try:
            (pv.capture if pv is not None else gv.capture)(bank.dev[mine[0]])
            torch.cuda.synchronize()
        except Exception as e:   # noqa: BLE001
            print(f"bench.py: CUDA-graph capture failed ({type(e).__name__}: {e}); falling back to eager launches",
                  file=sys.stderr, flush=True)
            gv = pv = None
            lanes = 0
            torch.cuda.synchronize()
    args.lanes = lanes
    args.no_graph = args.no_graph or lanes == 0

    def region(n_steps, first_step, eager=False, **kw):
        ddist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for s in range(n_steps):
            kw2 = {k: v for k, v in kw.items() if k != "from_host"}
            if pv is not None and not eager:
                training.multi_view_step_pipelined(avatar, bank, cam_ids(first_step + s), pv, weights=weights,
                                                   from_host=kw.get("from_host", False), **kw2)
            elif gv is not None and not eager:
                training.multi_view_step_graphed(avatar, bank, cam_ids(first_step + s), gv, weights=weights,
                                                 from_host=kw.get("from_host", False), **kw2)
            else:
                training.multi_view_step(avatar, bank, cam_ids(first_step + s), weights, bg, **kw)
        b.record()
        torch.cuda.synchronize()
        ddist.barrier()
        return ddist.max_over_ranks(a.elapsed_time(b), dev)

    # warm-up: W steps as asked (this also settles the workspace capacity hints and the allocator), then keep the
    # same work running until the GPU has been busy for about a second so clocks and power state have ramped
    ms_w = region(args.warmup, 0)                     # max over ranks: the same number on every rank
    extra = min(200, int(1000.0 / max(1e-3, ms_w / max(1, args.warmup))) + 1)
    clocks = ClockSampler(local) if rank == 0 else None
    if clocks:
        clocks.start()
    region(extra, 0)
    if clocks:
        clocks.mark()
    l0 = _lib.launch_count()
    ms = region(args.steps, args.warmup)
    clk = clocks.stop() if clocks else None
    launches = _lib.launch_count() - l0
    views_total = world * V * args.steps
    value = views_total / (ms * 1e-3)

    # same region again with per-kernel CUDA events (kept out of `value`): per-kernel durations for the roofline
    _lib.profile_enable(True)
    region(args.steps, args.warmup, eager=True)
    prof = _lib.profile_collect()
    _lib.profile_enable(False)

    # scene statistics for the algorithmic-bytes model
    from d3ga_b200 import diff_gaussian_rasterization as dgr
    stats = dict(dgr.last_stats)
    stats.update(dgr.sort_plan_of_last_forward())
    D = int(stats.get("num_rendered", 0))
    passes = int(stats.get("passes_executed", 6))
    N, P, Tl = scene.N, scene.W * scene.H, ((scene.W + 15) // 16) * ((scene.H + 15) // 16)
    bytes_per_launch = {
        "deform_forward_kernel": 120 * N, "geometry_kernel": 84 * N + 12 * D, "sort_histogram_kernel": 8 * D,
        "preprocess_kernel": 96 * N, "emit_sorted_kernel": 20 * N + 12 * D, "depth_plan_kernel": 0,
        "depth_sort_pass_kernel": 16 * N,
        "sort_pass_kernel": 24 * D, "tile_ranges_kernel": 8 * D + 8 * Tl, "blend_forward_kernel": 40 * D + 20 * P,
        "blend_backward_kernel": 80 * D + 20 * P, "preprocess_backward_kernel": 104 * N,
        # the deformation backward is two launches since the sparse form: the selection streams the incoming gradients and
        # writes the zero rows (36 + 44 B per Gaussian), the backward proper runs over the Gaussians that have a gradient;
        # together they do the 212 B per Gaussian of algorithmic work of the dense definition
        "deform_select_kernel": 80 * N,
        "deform_backward_kernel": (132 if prof.get("deform_select_kernel", (0, 0))[1] else 212) * N,
        "rect_grid_kernel": 12 * N, "tile_plan_kernel": 12 * Tl,
        # direct tile scatter (raster_direct.cu): gathers + records; ids written once, straight into their final slots
        # (the count rows and their prefix, 2-6 B x rows x tiles, are scratch traffic and not counted)
        "direct_count_kernel": 36 * N, "direct_plan_kernel": 12 * Tl, "direct_scatter_kernel": 16 * N + 4 * D,
        "pack_features_kernel": 28 * N}
    peak, peak_kind = measured_peak()
    kern = {}
    for name, (tms, cnt) in prof.items():
        if cnt:
            avg = tms / cnt
            kern[name] = {"launches_per_view": cnt / (V * args.steps), "avg_ms": avg, "total_ms": tms,
                          "GBps": bytes_per_launch.get(name, 0) / (avg * 1e-3) / 1e9 if avg > 0 else None}
    dominant = max(kern, key=lambda k: kern[k]["total_ms"]) if kern else None
    view_bytes = 540 * N + (148 + 24 * passes) * D + 40 * P + 8 * Tl
    if prof.get("direct_scatter_kernel", (0, 0))[1]:
        # no duplicate list is emitted, sorted or scanned for ranges (12 + 8 + 8 B per duplicate): the ids are written once
        passes = 0
        view_bytes = 540 * N + (120 + 4) * D + 40 * P + 8 * Tl
    roof = None
    traffic = traffic_src = issue = None
    try:   # DRAM bytes per launch and the issue-slot figures of the dominant kernel come from the committed `ncu --set
        #    full` capture (profiles/r2_traffic.json names the .csv it was read from and the commit that was profiled)
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        if args.workload == "metric" and dominant in tj["kernels"]:
            rec = tj["kernels"][dominant]
            traffic = rec["dram_bytes"]
            traffic_src = {"file": "profiles/r2_traffic.json", "from": tj.get("source"), "commit": tj.get("commit")}
            # the blend kernels are bound by issue slots, not by HBM: a second roofline says how full
            # the schedulers and the busiest pipes were in that capture
            issue = {"bound": "issue", "kernel": dominant, "achieved": rec.get("issue_active_pct"), "peak": 100.0,
                     "unit": "% of issue slots (while an SM sub-partition holds warps)",
                     "frac": rec.get("issue_active_pct", 0.0) / 100.0, "warp_instructions": rec.get("warp_instructions"),
                     "pipes_pct": rec.get("pipes_pct"), "warps_active_pct": rec.get("warps_active_pct")}
    except Exception:
        traffic = None
    if dominant:
        ach = kern[dominant]["GBps"]
        roof = {"bound": "hbm", "kernel": dominant, "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak if ach else None, "traffic": traffic, "traffic_source": traffic_src,
                "issue_roofline": issue, "peak_kind": peak_kind,
                "algorithmic_bytes_per_launch": bytes_per_launch.get(dominant),
                "avg_launch_ms": kern[dominant]["avg_ms"],
                "whole_view": {"bytes": view_bytes, "GBps": view_bytes * (V * args.steps) / (ms * 1e-3) / 1e9,
                               "frac": view_bytes * (V * args.steps) / (ms * 1e-3) / 1e9 / peak,
                               "frac_of_nominal_8000": view_bytes * (V * args.steps) / (ms * 1e-3) / 1e9 / 8000.0},
                "kernels": kern, "num_rendered": D, "sort_passes_executed": passes,
                "kernel_times_from": "an identical region of eager launches with per-kernel CUDA events (single-view sizing)"}

    # e2e: the same steps through the public API with this step's inputs in pinned HOST memory
    e2e = None
    if not args.no_e2e:
        feed = training.HostFeed(weights_u8, dev, cam_host=bank.host, tetpoints_host=avatar.tetpoints_host,
                                 slots=2 * max(1, args.lanes))
        region(1, 0, from_host=True, feed=feed, loss_host=loss_h)
        ms_e = region(args.steps, args.warmup, from_host=True, feed=feed, loss_host=loss_h)
        h2d = 35 * 4 + avatar.tetpoints.numel() * 4 + weights_u8.numel()
        e2e = {"value": views_total / (ms_e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d * V,
               "d2h_bytes_per_step": 4,
               "inputs_from_host": "per view: camera (35 floats), cage vertices tetpoints [V,3], target/weight image [3,H,W] as "
                                   "uint8, normalised on the device like lib/batch.py:86-88 "
                                   "(all uploaded ahead of the view on a side stream); the step's loss is read back once per step; "
                                   "Gaussian parameters stay resident like nn.Parameters"}

    # a D3GA training FRAME renders the same Gaussians twice (RGB + silhouette); frames/s with and
    # without sharing R1-R5 between the two renders (extra information, not the headline metric)
    pair = None
    if rank == 0 and world == 1 and not args.no_e2e:
        from d3ga_b200 import diff_gaussian_rasterization as dgr
        pair = {}
        for name, share, fused in (("frames_per_s_fused_pair", True, True), ("frames_per_s_shared_geometry", True, False),
                                   ("frames_per_s_independent_renders", False, False)):
            dgr.SHARE_GEOMETRY = share
            dgr._geom_cache.clear()
            for i in range(3):
                avatar.grads.zero_()
                training.frame_step(avatar, bank, mine[i % len(mine)], weights, bg, fused_pair=fused)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            nfr = 2 * V
            a.record()
            for i in range(nfr):
                training.frame_step(avatar, bank, mine[(3 + i) % len(mine)], weights, bg, fused_pair=fused)
            b.record()
            torch.cuda.synchronize()
            pair[name] = nfr / (a.elapsed_time(b) * 1e-3)
        dgr.SHARE_GEOMETRY = True
        # the fused frame captured into CUDA graphs and replayed with several frames in flight, like the headline's views
        if pv is not None:
            try:
                pvf = training.PipelinedViews(avatar, cams[0], bg, lanes=args.lanes, step=training.frame_pair_step)
                stepf = lambda ids: training.multi_view_step_pipelined(avatar, bank, ids, pvf, weights=weights)  # noqa: E731
                for s_ in range(3):
                    stepf(cam_ids(s_))
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for s_ in range(args.steps):
                    stepf(cam_ids(args.warmup + s_))
                b.record()
                torch.cuda.synchronize()
                pair["frames_per_s_fused_pair_graph_replay"] = V * args.steps / (a.elapsed_time(b) * 1e-3)
                del pvf
            except Exception as e:   # noqa: BLE001
                pair["frames_per_s_fused_pair_graph_replay"] = f"{type(e).__name__}: {e}"
        pair["note"] = ("1 frame = deformation block incl. the cage's FEM energy (one launch each way) + RGB render + silhouette "
                        "render + backward of all three terms (models/trainer.py:91-140, models/cage_net.py:213-230), eager "
                        "launches; fused_pair = both images from one GaussianRasterizer.forward_pair call (one binning, one "
                        "blend forward, one blend backward), shared_geometry = the reference's two render() calls with the "
                        "second one re-using the first one's binning; fused_pair_graph_replay = the fused frame as one CUDA-graph replay "
                        "per frame with the headline's number of frames in flight")

    # the cage-vertex skinning that produces tetpoints (Blueman.get), forward + backward to delta
    lbs = None
    if rank == 0 and world == 1 and not args.no_e2e:
        from d3ga_b200 import lbsmodel
        n_cage = int(avatar.tetpoints.shape[0])
        model_json, cfg, lbs_scale, motion = synth.make_skeleton(n_cage, seed=0)
        tmpl = np.asarray(model_json["SkinnedModel"]["RestPositions"], np.float32)
        mod = lbsmodel.LBSModule(model_json, cfg, tmpl, lbs_scale, np.float32(1.0)).to(dev)
        mot = torch.from_numpy(motion).to(dev)
        delta = torch.zeros(1, n_cage, 3, device=dev, requires_grad=True)
        rot180 = torch.diag(torch.tensor([-1.0, 1.0, -1.0, 1.0], device=dev))
        g_out = torch.randn(1, n_cage, 3, device=dev)

        def lbs_iter():
            delta.grad = None
            mod.cage_pose(mot, delta=delta, pre_mul=rot180).backward(g_out)

        for _ in range(5):
            lbs_iter()
        l0 = _lib.launch_count()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(200):
            lbs_iter()
        b.record()
        torch.cuda.synchronize()
        lbs_launches = (_lib.launch_count() - l0) / 200
        _lib.profile_enable(True)
        for _ in range(20):
            lbs_iter()
        kprof = {k: round(v[0] / v[1] * 1e3, 2) for k, v in _lib.profile_collect().items() if v[1] and k.startswith("lbs_")}
        _lib.profile_enable(False)
        lbs = {"us_per_pose_fwd_bwd": a.elapsed_time(b) * 1e3 / 200, "kernel_us": kprof, "kernel_launches_per_pose": lbs_launches,
               "joints": len(model_json["Skeleton"]["Bones"]), "cage_vertices": n_cage,
               "note": "Blueman.get (lib/blueman.py:139-167) + gradient to delta; the reference's joint loop alone is "
                       "~10 torch launches per joint (lbsmodel/body_model.py:327-345)"}
        if not args.no_cpu_baseline:
            from oracle import lbs_oracle
            sk = lbs_oracle.Skeleton(model_json, cfg, dtype=np.float32)
            t0 = time.perf_counter()
            for _ in range(3):
                lbs_oracle.cage_pose(sk, motion, lbs_scale, np.float32(1.0), tmpl[None], delta=np.zeros((1, n_cage, 3), np.float32))
            lbs["cpu_port_us_per_pose_fwd"] = (time.perf_counter() - t0) / 3 * 1e6

    # Extra information, not the headline: the same avatar with its Gaussians STORED in tet order (a one-time reordering
    # at initialisation; the order of the Gaussians means nothing to the model -- synth.sort_scene_by_tet).  The headline
    # keeps the layout SURVEY.md 8(d) specifies (tets drawn at random, i.e. Gaussians of a tet scattered in memory).
    tet_sorted = None
    if rank == 0 and world == 1 and not args.no_e2e and not args.no_graph:
        try:
            av2 = training.Avatar(synth.sort_scene_by_tet(scene), dev)
            training.multi_view_step(av2, bank, [mine[0]], weights, bg)
            if pv is not None:
                pv2 = training.PipelinedViews(av2, cams[0], bg, lanes=args.lanes)
                step2 = lambda ids: training.multi_view_step_pipelined(av2, bank, ids, pv2, weights=weights)  # noqa: E731
            else:
                gv2 = training.GraphedViewStep(av2, cams[0], bg)
                step2 = lambda ids: training.multi_view_step_graphed(av2, bank, ids, gv2, weights=weights)  # noqa: E731
            for s_ in range(3):
                step2(cam_ids(s_))
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for s_ in range(args.steps):
                step2(cam_ids(args.warmup + s_))
            b.record()
            torch.cuda.synchronize()
            tet_sorted = {"value": V * args.steps / (a.elapsed_time(b) * 1e-3), "unit": UNIT,
                          "note": "same workload, Gaussians stored in tet order (one-time reordering at initialisation)"}
            del av2
        except Exception as e:   # noqa: BLE001
            tet_sorted = {"error": f"{type(e).__name__}: {e}"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n = 2 if scene.N >= 500_000 else 4
        v_cpu, dt = time_cpu_views(scene, cams, weights_h.numpy(), n, warm=1)
        cpu = {"value": v_cpu, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
               "sample": f"{n} camera views of the same scene (fwd+bwd) after 1 warm-up view, {dt:.1f} s: torch CPU "
                         f"deformation + OpenMP C rasterizer oracle"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "ours",
                "config": workload_config(scene, args, V), "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
                "roofline": roof, "cpu_baseline": cpu, "training_frame_pair": pair, "cage_lbs": lbs, "tet_sorted_layout": tet_sorted,
                "allreduce_bytes_per_step": avatar.grads.nbytes() if world > 1 else 0}
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


def main():
    args = parse()
    # stdout carries ONE JSON line and nothing else: whatever a library prints to file descriptor 1 while the run lasts
    # (NCCL's version banner does, on some boxes, whatever NCCL_DEBUG_FILE says) goes to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
