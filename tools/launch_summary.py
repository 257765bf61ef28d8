"""Per-kernel summary of an ncu `--metrics gpu__time_duration.sum --csv` launch list: the last step's launches in order."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = None
seq = []
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr is None:
        continue
    d = dict(zip(hdr, r))
    try:
        seq.append((d["Kernel Name"].split("(")[0].split("::")[-1][:48], d.get("Grid Size", ""), float(d["Metric Value"].replace(",", "")) / 1000.0))
    except ValueError:
        pass
agg = collections.OrderedDict()
for name, grid, us in seq:
    agg.setdefault((name, grid), []).append(us)
tot_last = 0.0
for (name, grid), v in agg.items():
    print(f"{name:50s} {grid:14s} n={len(v):3d}  last-step us: {v[-1]:9.1f}   mean {sum(v)/len(v):9.1f}")
steps = min(len(v) for v in agg.values()) if agg else 0
print(f"# {len(seq)} launches of our kernels captured")
