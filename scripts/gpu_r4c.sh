mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_best.py -m gpu -x -q > gpurun_out/r4c_best_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4c_best_tests.log
tail -5 gpurun_out/r4c_best_tests.log
( time timeout 1500 python bench.py > gpurun_out/r4c_bench.json 2> gpurun_out/r4c_bench.err ) 2> gpurun_out/r4c_bench.time; echo "bench exit $?" >> gpurun_out/r4c_bench.err
tail -5 gpurun_out/r4c_bench.err; cat gpurun_out/r4c_bench.time
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4c_bench.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]))
print("ttr", json.dumps(d.get("time_to_reward"))[:1500])
print("fn", json.dumps(d.get("full_neighbourhood"))[:1800])
for w, s in d.get("workloads", {}).items():
    print(w, "%.4g" % s["value"], s["dtype"][:20], s["roofline"].get("kernel"), s["roofline"].get("frac"), json.dumps(s.get("full_neighbourhood"))[:900])
print("vm1084 solve", json.dumps(d.get("time_to_reward_vm1084"))[:600])
PY
