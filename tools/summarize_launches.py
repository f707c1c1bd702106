"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv, sys, re, collections
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
r = csv.DictReader(lines)
agg = collections.OrderedDict()
for row in r:
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = row["Kernel Name"]
    name = re.sub(r"\(.*", "", name)[:90]
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    if unit == "ns": v /= 1e3
    elif unit == "ms": v *= 1e3
    elif unit == "s": v *= 1e6
    c, t = agg.get(name, (0, 0.0))
    agg[name] = (c + 1, t + v)
tot = sum(t for _, t in agg.values())
print(f"# {sys.argv[1]}: {sum(c for c,_ in agg.values())} launches, {tot/1e3:.3f} ms total (cold-cache, serialised: compare shares)")
print(f"{'kernel':92s} {'launches':>8s} {'total_us':>12s} {'avg_us':>10s} {'share':>7s}")
for name, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:92s} {c:8d} {t:12.1f} {t/c:10.2f} {100*t/tot:6.2f}%")
