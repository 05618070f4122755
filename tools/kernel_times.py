"""Development helper: run bench.py (kernel-only legs) and print views/s plus the average duration of every kernel.
    python tools/kernel_times.py [label] [extra bench.py args...]     (environment variables pass through)"""
import json
import subprocess
import sys

label = sys.argv[1] if len(sys.argv) > 1 else ""
cmd = [sys.executable, "bench.py", "--steps", "5", "--warmup", "3", "--no-e2e", "--no-cpu-baseline"] + sys.argv[2:]
out = subprocess.run(cmd, capture_output=True, text=True)
try:
    j = json.loads(out.stdout.strip().splitlines()[-1])
except Exception:
    print(label, "FAILED", out.stderr[-2000:])
    sys.exit(1)
k = j["roofline"]["kernels"]
print(label, round(j["value"], 1), "views/s", {n.replace("_kernel", ""): round(v["avg_ms"] * 1e3, 1) for n, v in k.items()})
