timeout 900 python -m pytest tests -m gpu -x -q -k "multi_start or full_vns" > gpurun_out/gpu_tests_q.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_q.log
timeout 600 python bench.py --steps 5 --no-cpu > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "exit $?" >> gpurun_out/bench_q.err
