# r4p: second chain form with the re-scan of the third form running beside the tree rebuild: parity of every chain mode, vm1084
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_synthetic.py -m gpu -x -q -k "sequence or edge_cases or local_search or vns_run or synthetic" > gpurun_out/r4p_chain_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4p_chain_tests.log
tail -6 gpurun_out/r4p_chain_tests.log
timeout 600 python scripts/time_ls.py vm1084 c2-1t > gpurun_out/r4p_ls_vm1084.log 2>&1; cat gpurun_out/r4p_ls_vm1084.log
timeout 900 python scripts/time_solve.py vm1084 > gpurun_out/r4p_solve_vm1084.json 2> gpurun_out/r4p_solve_vm1084.err; cut -c1-400 gpurun_out/r4p_solve_vm1084.json
