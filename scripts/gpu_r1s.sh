# r1s: unrolled inner product of k_eval_vec, host call as a CUDA graph: all parity tests, both bench arms, launch list, capture
TAG=r1s
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -15 gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
tail -3 gpurun_out/bench_$TAG.err
VNSGPU_HOST_GRAPH=0 timeout 600 python bench.py --no-cpu --no-solve > gpurun_out/bench_${TAG}_nograph.json 2> gpurun_out/bench_${TAG}_nograph.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
BC="python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e"
timeout 300 $BC > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $BC > gpurun_out/ncu1.log 2>&1
bash scripts/gpu_cap.sh prof_eval_$TAG k_eval_vec 1 -- $BC
timeout 600 python bench.py --workload vm1084 --no-solve --cpu-sample 262144 > gpurun_out/bench_vm1084_$TAG.json 2> gpurun_out/bench_vm1084_$TAG.err
timeout 600 python bench.py --workload dop_s16 --no-solve --cpu-sample 262144 > gpurun_out/bench_dop_s16_$TAG.json 2> gpurun_out/bench_dop_s16_$TAG.err
du -sh gpurun_out
