"""Summarise an ncu launch list (gpu__time_duration.sum per launch) by kernel: python tools/launch_summary.py CSV [last_n]"""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr, rows = rows[0], rows[1:]
ix = {k: i for i, k in enumerate(hdr)}
last_n = int(sys.argv[2]) if len(sys.argv) > 2 else len(rows)
rows = rows[-last_n:]
unit = rows[0][ix["Metric Unit"]]
agg = collections.OrderedDict()
for r in rows:
    name = r[ix["Kernel Name"]].split("(")[0].replace("void <unnamed>::", "").replace("<unnamed>::", "")
    key = (name, r[ix["Grid Size"]], r[ix["Block Size"]])
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += float(r[ix["Metric Value"]].replace(",", ""))
tot = sum(a[1] for a in agg.values())
print(f"# {sys.argv[1]}: last {len(rows)} launches, gpu__time_duration.sum [{unit}], cold-cache serialised (ncu); total {tot:.1f}")
for (name, grid, blk), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:55s} grid={grid:14s} block={blk:14s} launches={n:4d} total={t:12.1f} avg={t / n:10.1f} share={t / tot:6.1%}")
