mkdir -p gpurun_out
timeout 300 python scripts/time_greedy.py synth20k > gpurun_out/r6e_greedy_synth20k.log 2>&1; tail -3 gpurun_out/r6e_greedy_synth20k.log
timeout 300 python scripts/time_greedy.py vm1084 > gpurun_out/r6e_greedy_vm1084.log 2>&1; tail -3 gpurun_out/r6e_greedy_vm1084.log
