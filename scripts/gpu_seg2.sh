mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_synthetic.py -m gpu -x -q -k "eval_batch or synthetic or 20k or host_call" > gpurun_out/seg2_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/seg2_tests.log
tail -4 gpurun_out/seg2_tests.log
for w in synth20k vm1084 synth_dop64k; do
timeout 600 python bench.py --workload $w --also "" --no-cpu --no-e2e --no-solve > gpurun_out/seg2_bench_$w.json 2> gpurun_out/seg2_bench_$w.err
python - $w <<'PY'
import json, sys
d = json.loads(open("gpurun_out/seg2_bench_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]; print(sys.argv[1], "%.3g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], {k: r[k] for k in ("kernel", "kernel_ms", "frac", "phases_ms_per_call")})
PY
done
