mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_synthetic.py -m gpu -x -q -k "20k_instance_local" > gpurun_out/t5_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/t5_tests.log
tail -4 gpurun_out/t5_tests.log
VNSGPU_VEC_STAGED=1 timeout 300 python bench.py --workload synth20k --also "" --no-e2e --no-solve > gpurun_out/t5_bench_synth20k.json 2> gpurun_out/t5_bench_synth20k.err
python - <<'PY'
import json, sys
d = json.loads(open("gpurun_out/t5_bench_synth20k.json").read().strip().splitlines()[-1])
r = d["roofline"]; print("staged bulk", "%.3g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], {k: r[k] for k in ("kernel", "kernel_ms", "frac", "phases_ms_per_call")}, d.get("cpu_baseline", {}).get("bit_identical_to_gpu"))
PY
