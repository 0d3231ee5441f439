mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/t2_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/t2_tests.log
tail -5 gpurun_out/t2_tests.log
for w in synth20k; do
timeout 600 python bench.py --workload $w --also "" --no-cpu --no-e2e --no-solve > gpurun_out/t2_bench_$w.json 2> gpurun_out/t2_bench_$w.err
python - $w <<'PY'
import json, sys
d = json.loads(open("gpurun_out/t2_bench_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]; print(sys.argv[1], "%.3g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], {k: r[k] for k in ("kernel", "kernel_ms", "frac", "phases_ms_per_call")})
PY
done
( time timeout 1500 python bench.py > gpurun_out/t2_bench.json 2> gpurun_out/t2_bench.err ) 2> gpurun_out/t2_time.txt; tail -3 gpurun_out/t2_bench.err; grep real gpurun_out/t2_time.txt
python - <<'PY'
import json
d = json.loads(open("gpurun_out/t2_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "ms_per_call", "gpu_launches")})
e = d["e2e"]; print("e2e", e["value"], "propose", e["propose_eval"]["value"], e["propose_eval"]["frac_of_value"], e["propose_eval"]["ms_per_step"])
PY
