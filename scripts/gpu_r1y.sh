# r1y (2 GPUs): sharded device entry point with a global key; parity suite, then 1- and 2-GPU bench + reference arm
TAG=r1y
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -5 gpurun_out/gpu_tests_$TAG.log
bash scripts/gpu_scale.sh 2
