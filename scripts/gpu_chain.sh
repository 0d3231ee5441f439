timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_ch.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_ch.log
timeout 300 python scripts/time_ls.py krob200 > gpurun_out/time_ls.log 2>&1
timeout 300 python scripts/time_ls.py vm1084 >> gpurun_out/time_ls.log 2>&1
