mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_synthetic.py tests/test_best.py tests/test_gpu_parity.py -m gpu -x -q -k "not sequence and not vns_run and not edge_cases" 2>&1 | tail -3
for w in krob200 synth20k vm1084; do
timeout 600 python bench.py --workload $w --also "" --no-e2e --no-cpu --no-solve --no-full-nbh > gpurun_out/r6a_bench_$w.json 2> gpurun_out/r6a_bench_$w.err; echo "exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r6a_bench_$w.json").read().strip().splitlines()[-1])
r = d["roofline"]; print("$w", "%.4g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], r["phases_ms_per_call"])
PY
done
