"""Opcode histogram + hottest SASS lines per kernel from `ncu --page source --csv --print-source sass`, and a few raw metrics:
python scripts/ncu_ops.py source.csv [raw.csv]"""
import csv, collections, re, sys
rows = list(csv.reader(open(sys.argv[1])))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
for si, h in enumerate(starts):
    end = starts[si + 1] - 1 if si + 1 < len(starts) else len(rows)
    name = rows[h - 1][1] if h > 0 else "?"
    hdr = rows[h]
    ie, ss = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
    op, st, tot = collections.Counter(), collections.Counter(), 0
    sec = [r for r in rows[h + 1:end] if len(r) > ie and r[ie].isdigit()]
    if not sec or not any(int(r[ie]) for r in sec):
        continue
    for r in sec:
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[1])
        o = m.group(2).split(".")[0] if m else "?"
        op[o] += int(r[ie]); tot += int(r[ie]); st[o] += int(r[ss]) if r[ss].isdigit() else 0
    print("==", name[:90], "total warp instructions", tot)
    for o, c in op.most_common(14):
        print("   %-12s %10d %5.1f%%  stall samples %d" % (o, c, 100 * c / tot, st[o]))
    for r in sorted(sec, key=lambda r: -(int(r[ss]) if r[ss].isdigit() else 0))[:8]:
        print("      ", r[1][:72], r[ss], r[ie])
if len(sys.argv) > 2:
    rows = list(csv.reader(open(sys.argv[2])))
    hdr = rows[0]
    keys = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print()
        for k in keys:
            if k in d:
                print("%-78s %s" % (k, d[k]))
