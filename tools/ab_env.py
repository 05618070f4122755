"""Development helper: alternate bench.py runs between environment settings and print views/s.
    python tools/ab_env.py <rounds> VAR=val[,VAR2=val2] [VAR=val ...]     ("-" = no change)"""
import json
import os
import subprocess
import sys

rounds, settings = int(sys.argv[1]), sys.argv[2:]
for _ in range(rounds):
    for st in settings:
        env = dict(os.environ)
        if st != "-":
            for kv in st.split(","):
                k, v = kv.split("=")
                env[k] = v
        out = subprocess.run([sys.executable, "bench.py", "--no-cpu-baseline", "--no-e2e"], env=env, capture_output=True, text=True)
        try:
            d = json.loads(out.stdout.strip().splitlines()[-1])
            print(st, round(d["value"], 1), flush=True)
        except Exception:
            print(st, "FAILED", out.stderr[-500:], flush=True)
