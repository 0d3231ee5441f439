# r1o: 64-bit block offsets, synthetic DOP flavour (64k samples, 33 GB table), 2D grids of the vector builds
TAG=r1o
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
timeout 600 python bench.py --workload synth20k --no-cpu --vector-cache 2 > gpurun_out/bench_synth20k_vec_$TAG.json 2> gpurun_out/bench_synth20k_vec_$TAG.err
timeout 600 python bench.py --workload synth_dop4k --vector-cache 2 > gpurun_out/bench_synth_dop4k_$TAG.json 2> gpurun_out/bench_synth_dop4k_$TAG.err
timeout 900 python bench.py --workload synth_dop64k --vector-cache 2 --sweep > gpurun_out/bench_synth_dop64k_$TAG.json 2> gpurun_out/bench_synth_dop64k_$TAG.err; echo "exit $?" >> gpurun_out/bench_synth_dop64k_$TAG.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv > gpurun_out/mem_$TAG.txt
timeout 600 python bench.py --no-solve --cpu-sample 1048576 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
BC="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e --vector-cache 2"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth_dop64k_$TAG.csv $BC > gpurun_out/ncu_dop64k.log 2>&1
