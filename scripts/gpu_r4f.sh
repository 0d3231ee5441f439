mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_best.py -m gpu -x -q > gpurun_out/r4f_best_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4f_best_tests.log
tail -12 gpurun_out/r4f_best_tests.log
timeout 300 python scripts/time_greedy.py synth20k > gpurun_out/r4f_greedy_synth20k.log 2>&1; cat gpurun_out/r4f_greedy_synth20k.log
VNSGPU_APPLY_PRODUCTS=0 timeout 300 python scripts/time_greedy.py synth20k > gpurun_out/r4f_greedy_synth20k_rescan.log 2>&1; cat gpurun_out/r4f_greedy_synth20k_rescan.log
timeout 300 python scripts/time_greedy.py vm1084 > gpurun_out/r4f_greedy_vm1084.log 2>&1; cat gpurun_out/r4f_greedy_vm1084.log
