# r1v: chunk count of the host-buffer pipeline now that a chunk is one kernel (no sort, no unpack, no conversion)
TAG=r1v
for c in 8 12 16; do VNSGPU_PIPE_CHUNKS=$c timeout 600 python bench.py --steps 20 --no-cpu --no-solve > gpurun_out/bench_${TAG}_c$c.json 2> gpurun_out/bench_${TAG}_c$c.err; done
for w in vm1084 dop_s16; do VNSGPU_PIPE_CHUNKS=16 timeout 600 python bench.py --workload $w --steps 20 --no-cpu --no-solve > gpurun_out/bench_${TAG}_${w}_c16.json 2> gpurun_out/bench_${TAG}_${w}_c16.err; done
