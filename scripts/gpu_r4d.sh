# r4d: ncu evidence at this commit - launch list of the default bench command, full capture of the dominant kernel at the
# bench's batch size (regenerates profiles/ncu_traffic.json's entry), full capture of the third chain form's kernel
TAG=r4d
mkdir -p gpurun_out
BC="python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh --also synth20k"
timeout 300 $BC > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $BC > gpurun_out/ncu1.log 2>&1
BK="python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh --also ''"
bash scripts/gpu_cap.sh prof_eval_$TAG k_eval_vec 1 -- python bench.py --steps 2 --warmup 1 --no-cpu --no-solve --no-e2e --no-full-nbh --also ""
LS="python scripts/time_ls.py krob200 c3"
timeout 300 $LS > gpurun_out/plain_ls.log 2>&1 && bash scripts/gpu_cap.sh prof_chain3_$TAG k_chain3_ls 3 -- $LS
FN="python scripts/time_greedy.py synth20k"
timeout 300 $FN > gpurun_out/greedy_synth20k_$TAG.log 2>&1; cat gpurun_out/greedy_synth20k_$TAG.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 3000 -c 60 --csv --log-file gpurun_out/launches_greedy_synth20k_$TAG.csv $FN > gpurun_out/ncu3.log 2>&1
du -sh gpurun_out; ls gpurun_out | grep $TAG
