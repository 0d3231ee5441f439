# r2e: does the L1 prefetch pay on L2-resident tables (vm1084: 4.7 MB table + 36 MB of vectors)?
TAG=r2e
for w in vm1084 krob200 dop_s16; do for pf in 0 1; do
  VNSGPU_PF_ALL=$pf timeout 600 python bench.py --workload $w --no-cpu --no-solve --no-e2e > gpurun_out/bench_${TAG}_${w}_pf$pf.json 2> gpurun_out/bench_${TAG}_${w}_pf$pf.err
done; done
