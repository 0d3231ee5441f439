BF="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e"
timeout 600 python bench.py --workload synth20k --steps 10 --cpu-sample 16384 > gpurun_out/bench_q_synth20k.json 2> gpurun_out/bench_q_synth20k.err; echo "exit $?" >> gpurun_out/bench_q_synth20k.err
timeout 600 python bench.py --workload synth20k --impl reference --ref-batch 16384 --steps 3 --warmup 1 > gpurun_out/bench_q_ref_synth20k.json 2> gpurun_out/bench_q_ref_synth20k.err
timeout 300 $BF > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_synth20k.csv $BF > gpurun_out/ncu1.log 2>&1
timeout 600 python bench.py --workload vm1084 --steps 3 --no-cpu --no-e2e > gpurun_out/bench_q_vm1084_solve.json 2> gpurun_out/bench_q_vm1084_solve.err; echo "exit $?" >> gpurun_out/bench_q_vm1084_solve.err
