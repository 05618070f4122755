"""Per-kernel shares of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X`).
    python tools/launch_summary.py profiles/r2_launches.csv "<header line>" > profiles/r2_launches_summary.txt"""
import csv
import re
import sys
from collections import defaultdict

path, title = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
lines = [l for l in open(path) if l.startswith('"')]
rows = list(csv.reader(lines))
hdr = rows[0]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = defaultdict(float)
cnt = defaultdict(int)
scatter = []
for r in rows[1:]:
    name = r[ik]
    us = float(r[iv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(r[iu], 1.0)
    m = re.search(r"(\w+_kernel)", name)
    ours = m and ("d3ga" in name or "unnamed" in name or any(k in name for k in ("blend_", "direct_", "deform_", "preprocess", "depth_", "sort_pass", "pack_features", "fem_", "ssim_", "lbs_")))
    key = m.group(1) if ours else "torch: " + re.sub(r"<.*", "", name.split("(")[0]).replace("void ", "").replace("at::", "")[:60]
    tot[key] += us
    cnt[key] += 1
    if key == "direct_scatter_kernel":
        scatter.append(round(us))
total = sum(tot.values())
print(title)
print(f"total {total:.1f} us over {sum(cnt.values())} launches (ncu serialised, cold cache: compare SHARES)")
print("direct_scatter_kernel launches, us:", scatter)
own = sum(v for k, v in tot.items() if not k.startswith("torch"))
print(f"share of this library's kernels: {100 * own / total:.1f}%")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"{v:10.1f} us {100 * v / total:5.1f}% n={cnt[k]:4d} {k}")
