# r4e: two GPUs - the NCCL exchange steps behind the C ABI, the command line's replica mode, the bench line at N = 2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r4e_multi_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4e_multi_gpu_tests.log
tail -5 gpurun_out/r4e_multi_gpu_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > gpurun_out/r4e_scale_n2.json 2> gpurun_out/r4e_scale_n2.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4e_scale_n2.json").read().strip().splitlines()[-1])
print("n_gpus", d["n_gpus"], "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "ms/step", d["ms_per_step"])
for w, s in d.get("workloads", {}).items(): print(w, "%.4g" % s["value"])
print(json.dumps(d.get("race_n_gpus"))[:600])
PY
