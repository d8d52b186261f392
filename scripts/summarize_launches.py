"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and shares."""
import csv
import re
import sys
from collections import defaultdict

path = sys.argv[1]
rows = []
with open(path) as f:
    lines = [ln for ln in f if not ln.startswith("==")]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").strip()
        rows.append((name, float(r["Metric Value"].replace(",", "")) / 1e3, r["Grid Size"], r["Block Size"]))
if len(sys.argv) > 2 and sys.argv[2] == "--last-step":
    # one train step = the launches from one preprocess_kernel up to the next one
    marks = [i for i, r in enumerate(rows) if "preprocess_kernel" in r[0]]
    if len(marks) >= 2:
        rows = rows[marks[-2]: marks[-1]]
tot = sum(t for _, t, _, _ in rows)
agg = defaultdict(lambda: [0, 0.0])
for n, t, _, _ in rows:
    agg[n][0] += 1
    agg[n][1] += t
print(f"launches: {len(rows)}   total device time: {tot / 1e3:.3f} ms (cold-cache, serialised under ncu: compare shares)")
print(f"{'kernel':70s} {'count':>6s} {'total us':>12s} {'share':>7s} {'avg us':>10s}")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n[:70]:70s} {c:6d} {t:12.1f} {100 * t / tot:6.1f}% {t / c:10.1f}")
