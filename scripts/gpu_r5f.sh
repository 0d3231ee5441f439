mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_seg_cross -s 6 -c 2 -o gpurun_out/r5f_seg_cross python bench.py --workload synth20k --also "" --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh > gpurun_out/r5f_ncu.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/r5f_seg_cross.ncu-rep --page raw --csv > gpurun_out/r5f_seg_cross_raw.csv 2>/dev/null
ncu -i gpurun_out/r5f_seg_cross.ncu-rep --page source --csv --print-source sass > gpurun_out/r5f_seg_cross_source.csv 2>/dev/null
ls -la gpurun_out | grep r5f
