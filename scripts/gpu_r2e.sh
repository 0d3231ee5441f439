# r2e: does the L1 prefetch pay on L2-resident tables (vm1084: 4.7 MB table + 36 MB of vectors)?
TAG=r2e
for w in vm1084 krob200 dop_s16; do for pf in 0 1; do
  VNSGPU_PF_ALL=$pf timeout 600 python bench.py --workload $w --no-cpu --no-solve --no-e2e > gpurun_out/bench_${TAG}_${w}_pf$pf.json 2> gpurun_out/bench_${TAG}_${w}_pf$pf.err
done; done
timeout 900 python bench.py --workload synth_dop64k --no-e2e --sweep > gpurun_out/bench_${TAG}_dop64k_sweep.json 2> gpurun_out/bench_${TAG}_dop64k_sweep.err
timeout 600 python -m pytest tests -m gpu -x -q -k "eval_batch or synthetic" > gpurun_out/gpu_tests_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests_$TAG.log
tail -3 gpurun_out/gpu_tests_$TAG.log
