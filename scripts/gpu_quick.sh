timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_q.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_q.log
timeout 600 python bench.py --steps 10 --no-solve > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "exit $?" >> gpurun_out/bench_q.err
timeout 600 python bench.py --workload vm1084 --steps 10 --no-solve --cpu-sample 262144 > gpurun_out/bench_q_vm1084.json 2> gpurun_out/bench_q_vm1084.err; echo "exit $?" >> gpurun_out/bench_q_vm1084.err
timeout 600 python bench.py --workload synth2k --steps 10 --cpu-sample 262144 > gpurun_out/bench_q_synth2k.json 2> gpurun_out/bench_q_synth2k.err; echo "exit $?" >> gpurun_out/bench_q_synth2k.err
timeout 600 python bench.py --workload synth20k --steps 10 --no-cpu > gpurun_out/bench_q_synth20k.json 2> gpurun_out/bench_q_synth20k.err; echo "exit $?" >> gpurun_out/bench_q_synth20k.err
timeout 900 python bench.py --workload vm1084 --steps 3 --no-cpu --no-e2e --no-solve-cpu --multi-start 0 > gpurun_out/bench_q_vm1084_solve.json 2> gpurun_out/bench_q_vm1084_solve.err; echo "exit $?" >> gpurun_out/bench_q_vm1084_solve.err
