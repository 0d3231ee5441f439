mkdir -p gpurun_out
for wv in 2 3 4 6 8 16; do
for w in synth20k synth_dop64k; do
VNSGPU_VEC_WAVES=$wv timeout 600 python bench.py --workload $w --also "" --no-e2e --no-cpu --no-solve --no-full-nbh > gpurun_out/r4y_bench_${w}_$wv.json 2> gpurun_out/r4y_bench_${w}_$wv.err
python - <<PY
import json
d = json.loads(open("gpurun_out/r4y_bench_${w}_$wv.json").read().strip().splitlines()[-1])
r = d["roofline"]; print("waves $wv $w", "%.4g evals/s" % d["value"], "ms/call %.3f" % d["ms_per_call"], r["phases_ms_per_call"])
PY
done
done
