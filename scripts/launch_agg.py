"""Per-kernel averages of an ncu launch list (gpu__time_duration, dram bytes): python scripts/launch_agg.py file.csv"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for r in rows:
    if len(r) > 10 and r[0] == "ID":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    d = dict(zip(hdr, r))
    k, m, u = d["Kernel Name"][:44], d["Metric Name"], d["Metric Unit"]
    v = float(d["Metric Value"].replace(",", ""))
    a = agg[k]
    if m == "gpu__time_duration.sum":
        a[0] += 1
        a[1] += v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(u, 1.0)
    else:
        v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
        a[2 if "read" in m else 3] += v
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-46s n=%3d avg %8.1f us  rd %8.1f MB  wr %8.1f MB" % (k, a[0], a[1] / a[0], a[2] / a[0], a[3] / a[0]))
