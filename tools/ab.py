"""Development helper: alternate bench.py runs between library builds (variants/libd3ga_<name>.so) and print views/s.
    python tools/ab.py <rounds> <name> [<name> ...]        (name "default" = the in-tree library)"""
import json
import os
import subprocess
import sys

rounds, names = int(sys.argv[1]), sys.argv[2:]
for _ in range(rounds):
    for n in names:
        env = dict(os.environ)
        if n != "default":
            env["D3GA_LIB"] = os.path.join("variants", f"libd3ga_{n}.so")
        out = subprocess.run([sys.executable, "bench.py", "--no-cpu-baseline", "--no-e2e"], env=env, capture_output=True, text=True)
        try:
            d = json.loads(out.stdout.strip().splitlines()[-1])
            t = d.get("tet_sorted_layout") or {}
            print(n, round(d["value"], 1), "tet-sorted", t.get("value"), flush=True)
        except Exception:
            print(n, "FAILED", out.stderr[-500:], flush=True)
