"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
    python scripts/launch_summary.py gpurun_out/launches.csv [top]"""
import collections, csv, io, sys
txt = open(sys.argv[1]).read()
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
rd = csv.DictReader(io.StringIO(txt[txt.find('"ID"'):]))
agg, tot, n = collections.OrderedDict(), 0.0, 0
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    a = agg.setdefault(r["Kernel Name"][:90], [0, 0.0]); a[0] += 1; a[1] += v; tot += v; n += 1
print(f"total {tot:.1f} us over {n} launches")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{t:10.1f} us {100 * t / tot:5.1f}% x{c:3d}  {k}")
