# r1r: compact host call (vnsgpu_eval_batch_packed32): parity + default bench + synth20k launch list
TAG=r1r
timeout 1200 python -m pytest tests -m gpu -x -q -k "eval_batch or abi or synthetic" > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -4 gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py --no-solve > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?" >> gpurun_out/bench_$TAG.err
tail -3 gpurun_out/bench_$TAG.err
for c in 8 12; do VNSGPU_PIPE_CHUNKS=$c timeout 600 python bench.py --steps 20 --no-cpu --no-solve > gpurun_out/bench_${TAG}_c$c.json 2> gpurun_out/bench_${TAG}_c$c.err; done
BS="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-e2e --no-solve"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth20k_$TAG.csv $BS > gpurun_out/ncu_s20k.log 2>&1
