"""Per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`): launches, summed and mean duration, share.

    python scripts/launch_summary.py gpurun_out/r02/launches_venice.csv
"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows:
    if r is hdr or r[ik] == "Kernel Name":
        continue
    try:
        v = float(r[iv].replace(",", ""))
    except ValueError:
        continue
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[iu].strip().split()[0] if r[iu].strip() else "ns", 1e-3)  # -> microseconds
    name = re.sub(r"\(.*", "", r[ik]).replace("void ba::", "").strip()
    tot[name] += v * scale
    cnt[name] += 1
all_us = sum(tot.values())
print(f"{'kernel':60s} {'launches':>8s} {'total us':>12s} {'mean us':>10s} {'share':>7s}")
for k in sorted(tot, key=lambda k: -tot[k]):
    print(f"{k[:60]:60s} {cnt[k]:8d} {tot[k]:12.1f} {tot[k] / cnt[k]:10.2f} {tot[k] / all_us:7.3f}")
