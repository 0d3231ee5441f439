mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_synthetic.py -m gpu -x -q -k "sequence or full_vns or multi_start or synthetic or 20k" > gpurun_out/c2e_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2e_tests.log
tail -3 gpurun_out/c2e_tests.log
timeout 600 python scripts/time_ls.py vm1084 c2-1t,c2-1t,c2-8t > gpurun_out/c2e_ls_vm1084.log 2>&1
timeout 900 python scripts/time_ls_synth.py 20000 1t,0t 400000 > gpurun_out/c2e_ls_synth20k.log 2>&1
cat gpurun_out/c2e_ls_*.log | grep -v final
for w in krob200 dop_s12 opn_s08 berlin52; do timeout 300 python scripts/time_solve.py $w; done > gpurun_out/c2e_solve_small.log 2>&1
cat gpurun_out/c2e_solve_small.log
timeout 900 python scripts/time_solve.py vm1084 > gpurun_out/c2e_solve_vm1084.log 2>&1
cat gpurun_out/c2e_solve_vm1084.log
