# r2h: k_seg_build with unclamped block rows over zero-initialised shared memory
TAG=r2h
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -4 gpurun_out/gpu_tests_$TAG.log
timeout 900 python bench.py --workload synth_dop64k --no-e2e > gpurun_out/bench_synth_dop64k_$TAG.json 2> gpurun_out/bench_synth_dop64k_$TAG.err
timeout 600 python bench.py --workload synth20k --no-cpu --no-e2e > gpurun_out/bench_synth20k_$TAG.json 2> gpurun_out/bench_synth20k_$TAG.err
BD="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_seg_build -c 3 --csv --log-file gpurun_out/l_dop64k_$TAG.csv $BD > gpurun_out/ncu_d.log 2>&1
BS="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-e2e --no-solve"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_seg_build -c 3 --csv --log-file gpurun_out/l_s20k_$TAG.csv $BS > gpurun_out/ncu_s.log 2>&1
for f in gpurun_out/l_*_$TAG.csv; do echo -n "$f: "; grep "gpu__time_duration" $f | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' '; echo; done
