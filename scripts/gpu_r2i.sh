# r2i: final validation set of the round on the last kernel commit
TAG=r2i
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -5 gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
tail -3 gpurun_out/bench_$TAG.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
BC="python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e"
timeout 300 $BC > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $BC > gpurun_out/ncu1.log 2>&1
bash scripts/gpu_cap.sh prof_eval_$TAG k_eval_vec 1 -- $BC
timeout 900 python bench.py --workload synth_dop64k --sweep > gpurun_out/bench_synth_dop64k_$TAG.json 2> gpurun_out/bench_synth_dop64k_$TAG.err
timeout 600 python bench.py --workload synth20k --cpu-sample 16384 --sweep > gpurun_out/bench_synth20k_$TAG.json 2> gpurun_out/bench_synth20k_$TAG.err
BD="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth_dop64k_$TAG.csv $BD > gpurun_out/ncu_dop64k.log 2>&1
BS="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-e2e --no-solve"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth20k_$TAG.csv $BS > gpurun_out/ncu_s20k.log 2>&1
bash scripts/gpu_cap.sh prof_evalhbm_$TAG k_eval_vec 1 -- $BD
timeout 600 python bench.py --workload vm1084 --no-solve --cpu-sample 262144 > gpurun_out/bench_vm1084_$TAG.json 2> gpurun_out/bench_vm1084_$TAG.err
timeout 600 python bench.py --workload dop_s16 --no-solve --cpu-sample 262144 > gpurun_out/bench_dop_s16_$TAG.json 2> gpurun_out/bench_dop_s16_$TAG.err
du -sh gpurun_out
