# r2b: evidence captures for the kernels that had none: the whole-VNS chain kernel (berlin52 solve) and the fused table build
TAG=r2b
BB="python bench.py --workload berlin52 --steps 2 --warmup 1 --no-cpu --no-e2e --no-solve-cpu --multi-start 0"
timeout 300 $BB > gpurun_out/plain_b52.log 2>&1
bash scripts/gpu_cap.sh prof_chain_$TAG k_chain_vns_run 0 -- $BB
BS="python bench.py --workload synth20k --steps 2 --warmup 1 --no-cpu --no-e2e --no-solve"
bash scripts/gpu_cap.sh prof_table_$TAG k_table_pairs 0 -- $BS
BD="python bench.py --workload synth_dop64k --steps 2 --warmup 1 --no-e2e"
bash scripts/gpu_cap.sh prof_vecpairs_$TAG k_vec_pairs 1 -- $BD
bash scripts/gpu_cap.sh prof_seg_$TAG k_seg_build 1 -- $BD
du -sh gpurun_out
