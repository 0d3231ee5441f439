# r4l (eight GPUs): the bench line at N = 8 and the sharded full-neighbourhood modes
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 > gpurun_out/r4l_scale_n8.json 2> gpurun_out/r4l_scale_n8.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4l_scale_n8.json").read().strip().splitlines()[-1])
print("n_gpus", d["n_gpus"], "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "ms/step", d["ms_per_step"])
for w, s in d.get("workloads", {}).items(): print(w, "%.4g" % s["value"])
print(json.dumps(d.get("race_n_gpus"))[:400])
PY
timeout 600 python scripts/time_greedy_sharded.py synth20k 8 > gpurun_out/r4l_sharded_synth20k_n8.log 2>&1; tail -3 gpurun_out/r4l_sharded_synth20k_n8.log
