mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_vec_ -s 4 -c 2 -o gpurun_out/r5k_vec python bench.py --workload synth20k --also "" --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh > gpurun_out/r5k_ncu.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/r5k_vec.ncu-rep --page raw --csv > gpurun_out/r5k_vec_raw.csv 2>/dev/null
ncu -i gpurun_out/r5k_vec.ncu-rep --page source --csv --print-source sass > gpurun_out/r5k_vec_source.csv 2>/dev/null
rm gpurun_out/r5k_vec.ncu-rep
