"""Turns the files tools/profile_round.sh left in gpurun_out/ into the tracked summaries under profiles/ (round 2)."""
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.chdir(ROOT)
for rep, out in (("gpurun_out/r2_blend.ncu-rep", "profiles/r2_ncu_full_blend.csv"), ("gpurun_out/r2_small.ncu-rep", "profiles/r2_ncu_full_small.csv"),
                 ("gpurun_out/r2_deform.ncu-rep", "profiles/r2_ncu_full_deform.csv")):
    with open(out, "w") as f:
        subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=f, stderr=subprocess.DEVNULL, check=True)
for a, b in (("r2_bench_metric.json", "r2_bench_metric.json"), ("r2_bench_reference.json", "r2_bench_metric_reference_arm.json"),
             ("r2_parity.jsonl", "r2_parity.jsonl"), ("r2_launches.csv", "r2_launches.csv"), ("r2_tests.log", "r2_gpu_tests.log")):
    shutil.copy(os.path.join("gpurun_out", a), os.path.join("profiles", b))
rows = list(csv.reader(open("profiles/r2_ncu_full_blend.csv")))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
f = lambda r, k: float(r[idx[k]].replace(",", ""))  # noqa: E731
scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1}
out = {"source": "profiles/r2_ncu_full_blend.csv (ncu --set full --clock-control none, tools/one_view.py metric 3 20: third view, camera 20 of the ring)",
       "commit": subprocess.check_output(["git", "rev-parse", "--short", "HEAD"]).decode().strip(), "kernels": {}}
for r in rows[2:]:
    name = "blend_forward_kernel" if "blend_forward" in r[idx["Kernel Name"]] else "blend_backward_kernel"
    dram = f(r, "dram__bytes_read.sum") * scale[units[idx["dram__bytes_read.sum"]]] + f(r, "dram__bytes_write.sum") * scale[units[idx["dram__bytes_write.sum"]]]
    out["kernels"][name] = {
        "dram_bytes": dram, "duration_us": f(r, "gpu__time_duration.sum"), "warp_instructions": f(r, "smsp__inst_executed.sum"),
        "issue_active_pct": f(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "warps_active_pct": f(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        "pipes_pct": {p: f(r, f"sm__inst_executed_pipe_{p}.avg.pct_of_peak_sustained_active") for p in ("alu", "fma", "lsu", "xu")},
        "l1_hit_pct": f(r, "l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": f(r, "lts__t_sector_hit_rate.pct"),
        "shared_wavefronts_pct_of_peak": f(r, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
        "sm_cycles_active_avg": f(r, "sm__cycles_active.avg"), "sm_cycles_active_max": f(r, "sm__cycles_active.max"),
        "stalls_per_issue": {k.split("stalled_")[1].split("_per_issue")[0]: round(f(r, k), 3) for k in hdr
                             if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio") and f(r, k) >= 0.2},
        "registers": f(r, "launch__registers_per_thread")}
json.dump(out, open("profiles/r2_traffic.json", "w"), indent=1)
title = ("round 2, final code, `python bench.py --steps 1 --warmup 1 --views 2 --no-e2e --no-cpu-baseline` under ncu --metrics "
         "gpu__time_duration.sum --clock-control none (includes the eager capacity / warm-up / capture views of the four lanes and "
         "the eager per-kernel region besides the replayed views)")
with open("profiles/r2_launches_summary.txt", "w") as fo:
    subprocess.run([sys.executable, "tools/launch_summary.py", "profiles/r2_launches.csv", title], stdout=fo, check=True)
d = json.loads(open("profiles/r2_bench_metric.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "ms/step", d["ms_per_step"], d["clocks"])
print({k: round(v["avg_ms"] * 1e3, 1) for k, v in d["roofline"]["kernels"].items()})
print({k: d["roofline"][k] for k in ("kernel", "achieved", "frac")}, d["roofline"]["whole_view"]["frac"])
print(d["training_frame_pair"], d["tet_sorted_layout"], d["cpu_baseline"]["value"])
for k, v in out["kernels"].items():
    print(k, {a: v[a] for a in ("duration_us", "warp_instructions", "issue_active_pct", "dram_bytes", "pipes_pct", "sm_cycles_active_avg", "sm_cycles_active_max")})
print(open("profiles/r2_launches_summary.txt").read()[:1800])
