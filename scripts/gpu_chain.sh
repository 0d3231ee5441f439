timeout 900 python -m pytest tests -m gpu -x -q -k "sequence or full_vns" > gpurun_out/gpu_tests_ch.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_ch.log
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_solve.json 2> gpurun_out/bench_solve.err
