"""Summarise an ncu --set full capture of k_eval_groups into profiles/: raw metrics CSV, a short text
summary and ncu_traffic.json (read by bench.py for roofline.traffic).
usage: python scripts/ncu_extract.py gpurun_out/prof_eval_TAG.ncu-rep TAG workload batch"""
import csv, io, json, subprocess, sys
rep, tag, workload, batch = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows0 = list(csv.reader(io.StringIO(raw)))
kname = rows0[-1][rows0[0].index("Kernel Name")].split("<")[0].split("(")[0].replace("void ", "").strip()
open("profiles/%s_%s_ncu_raw.csv" % (tag, kname), "w").write(raw)
rows = list(csv.reader(io.StringIO(raw)))
h, units, r = rows[0], rows[1], rows[-1]
def val(name):
    i = h.index(name); v = float(r[i].replace(",", "")); u = units[i]
    return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1, "us": 1e-6, "ms": 1e-3, "ns": 1e-9}.get(u, 1)
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "launch__registers_per_thread",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"]
out = {k: val(k) for k in keys if k in h}
out["kernel"] = r[h.index("Kernel Name")]
dram = out["dram__bytes_read.sum"] + out["dram__bytes_write.sum"]
entry = {"workload": workload, "batch": batch, "dram_bytes_per_launch": dram, "kernel": out["kernel"], "capture": rep.split("/")[-1],
         "note": "one ncu --set full --clock-control none capture; under ncu caches are flushed between replays"}
try:
    allent = json.load(open("profiles/ncu_traffic.json"))["entries"]
except (OSError, KeyError, ValueError):
    allent = []
allent = [e for e in allent if not (e["workload"] == workload and e["batch"] == batch)] + [entry]
json.dump({"entries": allent}, open("profiles/ncu_traffic.json", "w"), indent=1)
with open("profiles/%s_%s_summary.txt" % (tag, kname), "w") as fh:
    for k, v in out.items():
        fh.write("%-70s %s\n" % (k, v))
    fh.write("dram bytes per launch %.0f = %.1f B per candidate (batch %d)\n" % (dram, dram / batch, batch))
print(open("profiles/%s_%s_summary.txt" % (tag, kname)).read())
