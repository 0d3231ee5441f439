"""Summarise one exported ncu --set full capture (the `--page raw --csv` export written on the GPU box by
scripts/gpu_cap.sh) into profiles/: the raw CSV, a short text summary and - for the dominant kernel of a bench
workload - the entry of profiles/ncu_traffic.json that bench.py reads for roofline.traffic and roofline.cache_level.
usage: python scripts/ncu_summary.py gpurun_out/prof_X_raw.csv TAG NAME [workload batch]"""
import csv
import json
import shutil
import sys

raw, tag, name = sys.argv[1], sys.argv[2], sys.argv[3]
workload = sys.argv[4] if len(sys.argv) > 4 else None
batch = int(sys.argv[5]) if len(sys.argv) > 5 else None
rows = list(csv.reader(open(raw)))
h, units, r = rows[0], rows[1], rows[-1]
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "byte": 1, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1,
         "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1}


def val(k):
    i = h.index(k)
    return float(r[i].replace(",", "")) * SCALE.get(units[i], 1)


KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "launch__registers_per_thread",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"]
out = {k: val(k) for k in KEYS if k in h}
out["kernel"] = r[h.index("Kernel Name")]
shutil.copy(raw, "profiles/%s_%s_ncu_raw.csv" % (tag, name))
dram = out["dram__bytes_read.sum"] + out["dram__bytes_write.sum"]
with open("profiles/%s_%s_summary.txt" % (tag, name), "w") as fh:
    for k, v in out.items():
        fh.write("%-75s %s\n" % (k, v))
    if batch:
        fh.write("dram bytes per launch %.0f = %.1f B per candidate (batch %d)\n" % (dram, dram / batch, batch))
        fh.write("warp instructions per candidate %.1f\n" % (out["smsp__inst_executed.sum"] / batch))
print(open("profiles/%s_%s_summary.txt" % (tag, name)).read())
if workload:
    entry = {"workload": workload, "batch": batch, "dram_bytes_per_launch": dram, "kernel": out["kernel"],
             "capture": "profiles/%s_%s_ncu_raw.csv" % (tag, name),
             "kernel_us_under_ncu": out["gpu__time_duration.sum"] * 1e6,
             "l1tex_throughput_pct": out.get("l1tex__throughput.avg.pct_of_peak_sustained_active"),
             "lsu_wavefronts_pct": out.get("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
             "l2_throughput_pct": out.get("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
             "dram_throughput_pct": out.get("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
             "issue_active_pct": out.get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
             "l1_hit_pct": out.get("l1tex__t_sector_hit_rate.pct"),
             "note": "one ncu --set full --clock-control none capture; under ncu caches are flushed between replays"}
    try:
        allent = json.load(open("profiles/ncu_traffic.json"))["entries"]
    except (OSError, KeyError, ValueError):
        allent = []
    allent = [e for e in allent if not (e["workload"] == workload and e["batch"] == batch)] + [entry]
    json.dump({"entries": allent}, open("profiles/ncu_traffic.json", "w"), indent=1)
