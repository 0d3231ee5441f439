mkdir -p gpurun_out
( time timeout 1500 python bench.py > gpurun_out/b1_bench.json 2> gpurun_out/b1_bench.err ) 2> gpurun_out/b1_time.txt; echo "bench exit $?" >> gpurun_out/b1_time.txt
tail -5 gpurun_out/b1_bench.err; cat gpurun_out/b1_time.txt
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/b1_bench.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "ms_per_call", "gpu_launches")})
    print("e2e", d["e2e"]["value"] if d.get("e2e") else None)
    r = d["roofline"]; print({k: r[k] for k in ("kernel", "kernel_ms", "bound", "achieved", "peak", "frac", "working_set_bytes", "share_of_step", "phases_ms_per_call", "peaks_measured_gbs", "bytes_per_candidate")})
    for w, x in d.get("workloads", {}).items():
        r = x["roofline"]; print(w, x["value"], x["ms_per_step"], {k: r[k] for k in ("kernel", "kernel_ms", "bound", "achieved", "peak", "frac", "share_of_step", "phases_ms_per_call")})
    print("ttr", {k: d["time_to_reward"][k] for k in ("gpu_wall_s", "cpu_reference_wall_s", "same_result")})
    print("ttr vm1084", d.get("time_to_reward_vm1084"))
    print("cpu", d.get("cpu_baseline"))
except Exception as e:
    print("parse failed", e)
PY
