timeout 900 python -m pytest tests/test_synthetic.py -m gpu -x -q > gpurun_out/gpu_tests_q.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_q.log
timeout 600 python bench.py --steps 10 --no-cpu --no-solve > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "exit $?" >> gpurun_out/bench_q.err
timeout 600 python bench.py --workload synth2k --steps 10 --cpu-sample 262144 > gpurun_out/bench_q_synth2k.json 2> gpurun_out/bench_q_synth2k.err; echo "exit $?" >> gpurun_out/bench_q_synth2k.err
timeout 900 python bench.py --workload synth20k --steps 5 --batch 1048576 --cpu-sample 8192 > gpurun_out/bench_q_synth20k.json 2> gpurun_out/bench_q_synth20k.err; echo "exit $?" >> gpurun_out/bench_q_synth20k.err
