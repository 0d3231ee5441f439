# r6y (eight GPUs): the bench line at N = 8 at the round's last kernel commit
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 > gpurun_out/r6y_scale_n8.json 2> gpurun_out/r6y_scale_n8.err; echo "bench exit $?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r6y_scale_n8.json").read().strip().splitlines()[-1])
print("N=8 value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]))
for k, v in d.get("workloads", {}).items(): print(k, v.get("value"))
PY
