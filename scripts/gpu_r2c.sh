# r2c: HBM-sized tables: LAT form vs cache form of k_eval_vec, prefetch distance 1 / 2 (VNSGPU_HBM_FORM)
TAG=r2c
timeout 600 python -m pytest tests -m gpu -x -q -k "synthetic" > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -3 gpurun_out/gpu_tests_$TAG.log
for f in lat1 lat2 pf1 pf2; do
  BD="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e"
  VNSGPU_HBM_FORM=$f timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_eval_vec -c 3 --csv --log-file gpurun_out/l_dop64k_${TAG}_$f.csv $BD > gpurun_out/ncu_$f.log 2>&1
  BS="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-e2e --no-solve"
  VNSGPU_HBM_FORM=$f timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_eval_vec -c 3 --csv --log-file gpurun_out/l_s20k_${TAG}_$f.csv $BS > gpurun_out/ncu_s_$f.log 2>&1
done
grep -h "gpu__time_duration" gpurun_out/l_*_${TAG}_*.csv | awk -F'","' '{print $5, $NF}' | cut -c1-40,200-
for f in gpurun_out/l_*_${TAG}_*.csv; do echo $f; grep "gpu__time_duration" $f | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' '; echo; done
