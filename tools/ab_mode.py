"""Development helper: bench.py views/s per workload for the binning modes.
    python tools/ab_mode.py <workload> [<workload> ...]"""
import json
import os
import subprocess
import sys

for w in sys.argv[1:]:
    for mode in ("direct", "depthfirst"):
        env = dict(os.environ, D3GA_SORT_MODE=mode)
        out = subprocess.run([sys.executable, "bench.py", "--workload", w, "--no-cpu-baseline", "--no-e2e"], env=env,
                             capture_output=True, text=True)
        try:
            d = json.loads(out.stdout.strip().splitlines()[-1])
            k = {n.replace("_kernel", ""): round(v["avg_ms"] * 1e3 * v["launches_per_view"], 1) for n, v in d["roofline"]["kernels"].items()}
            print(w, mode, round(d["value"], 1), "views/s", k, flush=True)
        except Exception:
            print(w, mode, "FAILED", out.stderr[-800:], flush=True)
