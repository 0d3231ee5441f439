mkdir -p gpurun_out
for w in synth20k synth_dop64k vm1084 krob200; do
timeout 600 python bench.py --workload $w --also "" --no-e2e --no-cpu --no-solve --no-full-nbh > gpurun_out/r4x_bench_$w.json 2> gpurun_out/r4x_bench_$w.err; echo "exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r4x_bench_$w.json").read().strip().splitlines()[-1])
r = d["roofline"]; print("$w", "%.4g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], {k: r[k] for k in ("kernel", "kernel_ms", "phases_ms_per_call")})
PY
done
timeout 900 python -m pytest tests/test_synthetic.py tests/test_best.py tests/test_gpu_parity.py -m gpu -x -q -k "not sequence and not vns_run and not edge_cases" 2>&1 | tail -3
timeout 300 python scripts/time_greedy.py synth20k > gpurun_out/r4x_greedy_synth20k.log 2>&1; tail -4 gpurun_out/r4x_greedy_synth20k.log
