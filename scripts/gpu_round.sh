# one gpurun call: parity tests, both bench arms, ncu launch list + one full capture of the top kernel
TAG=${1:-r1}
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
timeout 600 python bench.py --workload vm1084 --no-solve --cpu-sample 262144 > gpurun_out/bench_vm1084_$TAG.json 2> gpurun_out/bench_vm1084_$TAG.err
timeout 300 python bench.py --workload vm1084 --impl reference --ref-batch 65536 --steps 5 > gpurun_out/bench_ref_vm1084_$TAG.json 2> gpurun_out/bench_ref_vm1084_$TAG.err
timeout 600 python bench.py --workload synth20k --cpu-sample 16384 > gpurun_out/bench_synth20k_$TAG.json 2> gpurun_out/bench_synth20k_$TAG.err
timeout 600 python bench.py --workload synth20k --no-cpu --vector-cache 2 > gpurun_out/bench_synth20k_vec_$TAG.json 2> gpurun_out/bench_synth20k_vec_$TAG.err
timeout 600 python bench.py --workload synth20k --impl reference --ref-batch 16384 --steps 3 --warmup 1 > gpurun_out/bench_ref_synth20k_$TAG.json 2> gpurun_out/bench_ref_synth20k_$TAG.err
BC="python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e"
timeout 300 $BC > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $BC > gpurun_out/ncu1.log 2>&1
# full capture of the dominant kernel at bench.py's DEFAULT batch (roofline.traffic is per launch of that size)
timeout 300 $BC > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_eval_vec -s 1 -c 1 -o gpurun_out/prof_eval_$TAG $BC > gpurun_out/ncu2.log 2>&1
