# r1q: warp-row k_seg_build; parity tests, both bench arms at the default workload, launch lists, full captures as CSV
TAG=r1q
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -4 gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
tail -2 gpurun_out/bench_$TAG.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
BC="python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e"
timeout 300 $BC > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $BC > gpurun_out/ncu1.log 2>&1
bash scripts/gpu_cap.sh prof_eval_$TAG k_eval_vec 1 -- $BC
timeout 900 python bench.py --workload synth_dop64k --vector-cache 2 --sweep > gpurun_out/bench_synth_dop64k_$TAG.json 2> gpurun_out/bench_synth_dop64k_$TAG.err; echo "exit $?" >> gpurun_out/bench_synth_dop64k_$TAG.err
timeout 600 python bench.py --workload synth20k --no-cpu --vector-cache 2 --sweep > gpurun_out/bench_synth20k_vec_$TAG.json 2> gpurun_out/bench_synth20k_vec_$TAG.err
BD="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e --vector-cache 2"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth_dop64k_$TAG.csv $BD > gpurun_out/ncu_dop64k.log 2>&1
bash scripts/gpu_cap.sh prof_vecuw_$TAG k_vec_uw 1 -- $BD
bash scripts/gpu_cap.sh prof_vecpairs_$TAG k_vec_pairs 1 -- $BD
bash scripts/gpu_cap.sh prof_evallat_$TAG k_eval_vec 1 -- $BD
du -sh gpurun_out
