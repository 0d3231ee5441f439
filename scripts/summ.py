"""Print the headline fields of a bench JSON line (helper for reading gpurun_out/)."""
import json, sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f)
    if d.get("impl") == "reference":
        print("  reference arm: %.3g evals/s on %d cores" % (d["value"], d["cpu_baseline"]["cores"])); continue
    r = d["roofline"]
    print("  value %.4g evals/s  %.3f ms/step  launches %s  clocks %s" % (d["value"], d["ms_per_step"], d["gpu_launches"], d["clocks"]))
    if d.get("e2e"): print("  e2e %.4g evals/s  %.3f ms/step" % (d["e2e"]["value"], d["e2e"]["ms_per_step"]))
    print("  roofline %.1f GB/s frac %.3f  kernel %.3f ms  %.0f B/cand" % (r["achieved"], r["frac"], r["kernel_ms"], r["algorithmic_bytes_per_candidate"]))
    if "cpu_baseline" in d: print("  cpu", d["cpu_baseline"])
    if "time_to_reward" in d:
        t = d["time_to_reward"]; print("  solve gpu %.2fs cpu %.2fs same %s" % (t["gpu_wall_s"], t.get("cpu_reference_wall_s", -1), t.get("same_result")))
