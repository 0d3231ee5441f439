# chain2 first contact: parity of every chain mode, then local-search timings
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "sequence or full_vns or multi_start" > gpurun_out/c2a_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2a_tests.log
timeout 300 python scripts/time_ls.py krob200 c2-1,c2-2,c2-4,v1 > gpurun_out/c2a_ls_krob200.log 2>&1
timeout 300 python scripts/time_ls.py dop_s12 c2-1,v1 > gpurun_out/c2a_ls_dop_s12.log 2>&1
timeout 600 python scripts/time_ls.py vm1084 c2-1,c2-8,c2-24,c2-64,v1 > gpurun_out/c2a_ls_vm1084.log 2>&1
tail -5 gpurun_out/c2a_tests.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/c2a_noalloc_alone.log 2>&1; echo "exit $?" >> gpurun_out/c2a_noalloc_alone.log
tail -3 gpurun_out/c2a_noalloc_alone.log
