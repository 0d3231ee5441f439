mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "device_side" > gpurun_out/t3_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/t3_tests.log
tail -4 gpurun_out/t3_tests.log
( time timeout 1500 python bench.py --also "" --no-solve --no-cpu > gpurun_out/t3_bench.json 2> gpurun_out/t3_bench.err ) 2> gpurun_out/t3_time.txt; tail -3 gpurun_out/t3_bench.err; grep real gpurun_out/t3_time.txt
python - <<'PY'
import json
d = json.loads(open("gpurun_out/t3_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "ms_per_call", "gpu_launches")})
e = d["e2e"]; print("e2e", e["value"], "propose", e["propose_eval"]["value"], e["propose_eval"]["frac_of_value"], e["propose_eval"]["ms_per_step"])
PY
