timeout 900 python -m pytest tests -m gpu -x -q -k "cli or multi_start or abi" > gpurun_out/gpu_tests_q.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_q.log
timeout 600 python bench.py --steps 10 --no-cpu --no-solve > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "exit $?" >> gpurun_out/bench_q.err
