# r4k (two GPUs): NCCL tests incl. the sharded full-neighbourhood modes, their timing on the 20 k-node instance
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_best.py -m gpu -x -q > gpurun_out/r4k_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4k_tests.log
tail -6 gpurun_out/r4k_tests.log
timeout 600 python scripts/time_greedy_sharded.py synth20k 2 > gpurun_out/r4k_sharded_synth20k.log 2>&1; cat gpurun_out/r4k_sharded_synth20k.log | tail -4
