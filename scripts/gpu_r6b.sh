mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_eval_vec -s 2 -c 1 -o gpurun_out/r6b_eval python bench.py --workload synth20k --also "" --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh > gpurun_out/r6b_ncu.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/r6b_eval.ncu-rep --page raw --csv > gpurun_out/r6b_eval_raw.csv 2>/dev/null
ncu -i gpurun_out/r6b_eval.ncu-rep --page source --csv --print-source sass > gpurun_out/r6b_eval_source.csv 2>/dev/null
rm gpurun_out/r6b_eval.ncu-rep
