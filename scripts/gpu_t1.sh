mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "device_side or real_valued or dop_s16 or edge_cases or eval_batch" > gpurun_out/t1_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/t1_tests.log
tail -12 gpurun_out/t1_tests.log
