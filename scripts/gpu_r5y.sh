# r5y (two GPUs): NCCL tests through the C ABI and the bench line at N = 2 at the round's last kernel commit
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r5y_multi_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r5y_multi_gpu_tests.log
tail -4 gpurun_out/r5y_multi_gpu_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 > gpurun_out/r5y_scale_n2.json 2> gpurun_out/r5y_scale_n2.err; echo "bench exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r5y_scale_n2.json").read().strip().splitlines()[-1])
print("N=2 value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]))
for k, v in d.get("workloads", {}).items(): print(k, v.get("value"))
PY
