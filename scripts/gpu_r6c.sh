mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_synthetic.py tests/test_best.py tests/test_gpu_parity.py -m gpu -x -q -k "not sequence and not vns_run and not edge_cases" 2>&1 | tail -3
for w in synth20k synth_dop64k vm1084 krob200; do
timeout 600 python bench.py --workload $w --also "" --no-e2e --no-cpu --no-solve --no-full-nbh > gpurun_out/r6c_bench_$w.json 2> gpurun_out/r6c_bench_$w.err; echo "exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r6c_bench_$w.json").read().strip().splitlines()[-1])
r = d["roofline"]; print("$w", "%.4g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], r["phases_ms_per_call"])
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 300 --csv --log-file gpurun_out/r6c_launches_synth20k.csv python bench.py --workload synth20k --also "" --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh > gpurun_out/r6c_ncu.log 2>&1; echo "ncu exit $?"
