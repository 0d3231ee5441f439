mkdir -p gpurun_out
timeout 600 python scripts/time_greedy_sharded.py synth20k 2 > gpurun_out/r7b_sharded_synth20k_n2.log 2>&1; tail -6 gpurun_out/r7b_sharded_synth20k_n2.log
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3
