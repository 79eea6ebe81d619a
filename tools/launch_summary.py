"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and
share of the LAST `--step-launches` launches (one step of bench.py --profile)."""
import csv, collections, sys, re, argparse
ap = argparse.ArgumentParser()
ap.add_argument("csv"); ap.add_argument("--last", type=int, default=0)
ap.add_argument("--by-grid", action="store_true")
a = ap.parse_args()
lines = open(a.csv).readlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[start:]))
if a.last:
    rows = rows[-a.last:]
def short(n):
    n = re.sub(r"\(.*", "", n); n = re.sub(r"^void ", "", n)
    return n[:70]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
    k = short(r["Kernel Name"]) + (f" grid={r['Grid Size']}" if a.by_grid else "")
    agg[k][0] += 1; agg[k][1] += v
tot = sum(v[1] for v in agg.values())
print(f"launches {len(rows)}  total {tot:.2f} ms")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{t:9.3f} ms {100*t/tot:5.1f}%  n={n:5d}  avg {1e3*t/n:8.1f} us  {k}")
