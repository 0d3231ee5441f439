mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_synthetic.py tests/test_best.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python scripts/time_greedy.py synth20k > gpurun_out/r6f_greedy_synth20k.log 2>&1; tail -3 gpurun_out/r6f_greedy_synth20k.log
