# r1x: uint16 table copy for the vector builds, always-16-byte staging, 3 CTAs / SM for the four-register form
TAG=r1x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -5 gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py --no-solve --cpu-sample 1048576 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
tail -3 gpurun_out/bench_$TAG.err
timeout 600 python bench.py --workload vm1084 --no-solve --cpu-sample 262144 > gpurun_out/bench_vm1084_$TAG.json 2> gpurun_out/bench_vm1084_$TAG.err
timeout 900 python bench.py --workload synth_dop64k --sweep > gpurun_out/bench_synth_dop64k_$TAG.json 2> gpurun_out/bench_synth_dop64k_$TAG.err; echo "exit $?" >> gpurun_out/bench_synth_dop64k_$TAG.err
timeout 600 python bench.py --workload synth20k --cpu-sample 16384 --sweep > gpurun_out/bench_synth20k_$TAG.json 2> gpurun_out/bench_synth20k_$TAG.err
BD="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth_dop64k_$TAG.csv $BD > gpurun_out/ncu_dop64k.log 2>&1
BS="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-e2e --no-solve"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth20k_$TAG.csv $BS > gpurun_out/ncu_s20k.log 2>&1
bash scripts/gpu_cap.sh prof_vecuw_$TAG k_vec_uw 1 -- $BD
du -sh gpurun_out
