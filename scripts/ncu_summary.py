"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
    python scripts/ncu_summary.py gpurun_out/launches.csv > profiles/rNN_launches_summary.txt"""
import collections, csv, re, sys

path = sys.argv[1]
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
tot, n = 0.0, 0
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    v = v / 1e3 if unit == "ns" else v * 1e3 if unit == "ms" else v
    name = re.sub(r"\(.*", "", row["Kernel Name"])
    agg[name][0] += 1; agg[name][1] += v; tot += v; n += 1
print(f"# {path}: {n} launches, {tot/1e3:.3f} ms of serialised cold-cache device time (compare SHARES, not absolutes)")
print(f"# {'ms':>9} {'share':>6} {'count':>5} {'avg us':>9}  kernel")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t/1e3:11.3f} {100*t/tot:5.1f}% {c:5d} {t/c:9.1f}  {k[:120]}")
