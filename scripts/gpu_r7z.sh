mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r7z_gpu_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r7z_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r7z_bench.json 2> gpurun_out/r7z_bench.err; echo "bench exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r7z_bench.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]), d["roofline"].get("phases_ms_per_call"))
for k, v in d.get("workloads", {}).items(): print(k, v.get("value"))
t = d.get("time_to_reward"); print(t["gpu_wall_s"], t["cpu_reference_wall_s"], t["multi_start"]["gpu_runs_per_s"], t["multi_start"]["dense"]["gpu_runs_per_s"])
t = d.get("time_to_reward_dop_s16"); print(t["gpu_wall_s"], t["cpu_reference_wall_s"], t["multi_start"]["gpu_runs_per_s"], t["multi_start"]["dense"]["gpu_runs_per_s"])
print(d.get("full_neighbourhood"))
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r7z_bench_ref.json 2> gpurun_out/r7z_bench_ref.err; echo "ref exit $?"; tail -c 400 gpurun_out/r7z_bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r7z_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh --also synth20k > gpurun_out/r7z_ncu.log 2>&1; echo "ncu exit $?"
