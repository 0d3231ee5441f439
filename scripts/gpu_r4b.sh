# best.cu tests, then the whole GPU suite at this commit
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_best.py -m gpu -x -q > gpurun_out/r4b_best_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4b_best_tests.log
tail -15 gpurun_out/r4b_best_tests.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r4b_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4b_gpu_tests.log
tail -8 gpurun_out/r4b_gpu_tests.log
