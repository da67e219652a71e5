"""Kernel-time shares from an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python tools/launch_shares.py FILE.csv [FIRST_FRACTION LAST_FRACTION]  (slice of the launch sequence)"""
import collections
import csv
import re
import sys

path = sys.argv[1]
lo, hi = (float(sys.argv[2]), float(sys.argv[3])) if len(sys.argv) > 3 else (0.0, 1.0)
rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
ix = {h: i for i, h in enumerate(rows[0])}
recs = [(r[ix["Kernel Name"]], float(r[ix["Metric Value"]].replace(",", "")))
        for r in rows[1:] if len(r) > ix["Metric Value"] and r[ix["Metric Name"]] == "gpu__time_duration.sum"]
unit = rows[1][ix["Metric Unit"]]
scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1.0)
recs = recs[int(lo * len(recs)):int(hi * len(recs))]
agg, cnt = collections.Counter(), collections.Counter()
for k, v in recs:
    k = re.sub(r"^void ", "", k)
    k = re.sub(r"\(.*", "", k)[:90]
    agg[k] += v * scale
    cnt[k] += 1
tot = sum(agg.values())
print(f"{len(recs)} launches, {tot / 1e3:.3f} ms of kernel time")
for k, v in agg.most_common(30):
    print(f"{v / tot * 100:5.1f}% {cnt[k]:5d} x {v / cnt[k]:8.2f} us  {k}")
