mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_synthetic.py tests/test_gpu_parity.py -m gpu -x -q -k "synthetic or 20k or device_side or eval_batch" > gpurun_out/t4_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/t4_tests.log
tail -5 gpurun_out/t4_tests.log
for st in 1 0; do
VNSGPU_VEC_STAGED=$st timeout 600 python bench.py --workload synth20k --also "" --no-e2e --no-solve > gpurun_out/t4_bench_synth20k_st$st.json 2> gpurun_out/t4_bench_synth20k_st$st.err
python - $st <<'PY'
import json, sys
d = json.loads(open("gpurun_out/t4_bench_synth20k_st%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]; print("staged", sys.argv[1], "%.3g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], {k: r[k] for k in ("kernel", "kernel_ms", "frac", "phases_ms_per_call")}, d.get("cpu_baseline", {}).get("bit_identical_to_gpu"))
PY
done
( time timeout 1500 python bench.py --also "" --no-solve --no-cpu > gpurun_out/t4_bench.json 2> gpurun_out/t4_bench.err ) 2> gpurun_out/t4_time.txt; tail -3 gpurun_out/t4_bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/t4_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "ms_per_call", "gpu_launches")})
e = d["e2e"]; print("e2e", e["value"], "propose", e["propose_eval"]["value"], e["propose_eval"]["frac_of_value"], e["propose_eval"]["ms_per_step"])
PY
