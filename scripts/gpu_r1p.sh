# r1p: HEAD validation after the container restore: parity tests, both bench arms at the default workload,
# ncu launch list + one full capture of the dominant kernel (k_eval_vec) of the same command
TAG=r1p
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
BC="python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e"
timeout 300 $BC > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $BC > gpurun_out/ncu1.log 2>&1
timeout 300 $BC > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_eval_vec -s 1 -c 1 -o gpurun_out/prof_eval_$TAG $BC > gpurun_out/ncu2.log 2>&1
BD="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e --vector-cache 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_vec_uw -s 1 -c 1 -o gpurun_out/prof_vecuw_$TAG $BD > gpurun_out/ncu3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_seg_build -s 1 -c 1 -o gpurun_out/prof_seg_$TAG $BD > gpurun_out/ncu4.log 2>&1
