mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r4z_gpu_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/r4z_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r4z_bench.json 2> gpurun_out/r4z_bench.err; echo "bench exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r4z_bench.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]), d["roofline"].get("phases_ms_per_call"))
for k, v in d.get("workloads", {}).items(): print(k, v.get("value"))
print(d.get("time_to_reward"), d.get("time_to_reward_dop_s16"))
PY
