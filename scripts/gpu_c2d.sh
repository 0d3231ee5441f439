mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_synthetic.py -m gpu -x -q -k "sequence or full_vns or multi_start or synthetic or 20k" > gpurun_out/c2d_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2d_tests.log
tail -4 gpurun_out/c2d_tests.log
timeout 300 python scripts/time_ls.py krob200 c2-1,c2-1,c2-1t,v1 > gpurun_out/c2d_ls_krob200.log 2>&1
timeout 600 python scripts/time_ls.py vm1084 c2-1t,c2-1t,c2-8t > gpurun_out/c2d_ls_vm1084.log 2>&1
timeout 600 python scripts/time_ls_synth.py 2000 1t,4t 400000 > gpurun_out/c2d_ls_synth2k.log 2>&1
timeout 900 python scripts/time_ls_synth.py 20000 1t,16t,0t 400000 > gpurun_out/c2d_ls_synth20k.log 2>&1
cat gpurun_out/c2d_ls_*.log | grep -v final
