# r1k: fused table build, pipelined segment products, race mode
TAG=r1k
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
timeout 600 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
timeout 600 python bench.py --workload synth20k --no-cpu --vector-cache 2 > gpurun_out/bench_synth20k_vec_$TAG.json 2> gpurun_out/bench_synth20k_vec_$TAG.err
timeout 600 python bench.py --workload synth20k --no-cpu > gpurun_out/bench_synth20k_$TAG.json 2> gpurun_out/bench_synth20k_$TAG.err
timeout 600 python bench.py --workload dop_s16 --no-cpu --no-solve > gpurun_out/bench_dop_s16_$TAG.json 2> gpurun_out/bench_dop_s16_$TAG.err
BC="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --vector-cache 2"
timeout 300 $BC > gpurun_out/plain_s20k.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth20k_vec_$TAG.csv $BC > gpurun_out/ncu_s20k.log 2>&1
