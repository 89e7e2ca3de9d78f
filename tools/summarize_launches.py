"""Summarise an ncu --csv launch list (gpu__time_duration.sum) by kernel name."""
import csv, sys, re, collections
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
r = csv.DictReader(lines)
tot = collections.OrderedDict()
for row in r:
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = row["Kernel Name"]
    name = re.sub(r"\(.*", "", name)
    grid = row.get("Grid Size", "")
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    if unit == "ns":
        v /= 1e3
    elif unit == "ms":
        v *= 1e3
    key = (name, grid)
    t = tot.setdefault(key, [0, 0.0])
    t[0] += 1
    t[1] += v
total = sum(t[1] for t in tot.values())
print(f"total {total:.1f} us over {sum(t[0] for t in tot.values())} launches")
for (name, grid), (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{us/total*100:5.1f}%  {us:10.1f} us  n={n:4d}  avg {us/n:8.2f} us  {name[:60]} grid={grid}")
