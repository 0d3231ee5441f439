mkdir -p gpurun_out
B="python bench.py --workload synth20k --also '' --no-cpu --no-e2e --no-solve --steps 2 --warmup 3 --min-timed-s 0"
eval "$B" > gpurun_out/cap1_plain.json 2> gpurun_out/cap1_plain.err || exit 1
VNSGPU_VEC_STAGED=0 bash scripts/gpu_cap.sh cap1_evalvec_pf_synth20k "k_eval_vec" 3 -- python bench.py --workload synth20k --also "" --no-cpu --no-e2e --no-solve --steps 2 --warmup 3 --min-timed-s 0
VNSGPU_VEC_STAGED=1 bash scripts/gpu_cap.sh cap1_evalvec_st_synth20k "k_eval_vec_st" 3 -- python bench.py --workload synth20k --also "" --no-cpu --no-e2e --no-solve --steps 2 --warmup 3 --min-timed-s 0
ls -la gpurun_out/cap1*
