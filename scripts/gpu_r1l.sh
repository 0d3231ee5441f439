# r1l: endpoint vectors v2 (k_vec_uw + k_vec_pairs), seg build ring
TAG=r1l
timeout 900 python -m pytest tests -m gpu -x -q -k "eval_batch or synthetic or edge or real_valued" > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
timeout 600 python bench.py --no-solve --cpu-sample 1048576 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
timeout 600 python bench.py --workload synth20k --cpu-sample 16384 --vector-cache 2 > gpurun_out/bench_synth20k_vec_$TAG.json 2> gpurun_out/bench_synth20k_vec_$TAG.err
timeout 600 python bench.py --workload vm1084 --no-solve --cpu-sample 65536 > gpurun_out/bench_vm1084_$TAG.json 2> gpurun_out/bench_vm1084_$TAG.err
BC="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --vector-cache 2"
timeout 300 $BC > gpurun_out/plain_s20k.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth20k_vec_$TAG.csv $BC > gpurun_out/ncu_s20k.log 2>&1
