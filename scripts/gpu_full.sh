mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/full_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/full_gpu_tests.log
tail -4 gpurun_out/full_gpu_tests.log
