# chain2 + product tree: parity of every chain mode, then local-search timings with / without the tree
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "sequence or full_vns or multi_start" > gpurun_out/c2b_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2b_tests.log
tail -5 gpurun_out/c2b_tests.log
timeout 300 python scripts/time_ls.py krob200 c2-1,c2-1,c2-1t,v1 > gpurun_out/c2b_ls_krob200.log 2>&1
timeout 600 python scripts/time_ls.py vm1084 c2-1,c2-1,c2-1t,c2-8t,c2-24t,v1 > gpurun_out/c2b_ls_vm1084.log 2>&1
cat gpurun_out/c2b_ls_krob200.log gpurun_out/c2b_ls_vm1084.log | grep -v final
