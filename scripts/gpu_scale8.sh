# 8-GPU box: default workload at N = 1, 4, 8 (+ the reference arm at N = 8), config 5 (synth20k) sweep at N = 8
bash scripts/gpu_scale.sh 4
bash scripts/gpu_scale.sh 8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --workload synth20k --no-cpu --no-e2e --no-solve --sweep > gpurun_out/scale_synth20k_n8.json 2> gpurun_out/scale_synth20k_n8.err
timeout 600 python bench.py --workload synth20k --no-cpu --no-e2e --no-solve --sweep > gpurun_out/scale_synth20k_n1.json 2> gpurun_out/scale_synth20k_n1.err
