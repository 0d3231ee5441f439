# chain2 on the synthetic scaled instances: placements, 20k local search vs the C port, timings
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_synthetic.py -m gpu -x -q -s > gpurun_out/c2c_synth_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2c_synth_tests.log
grep -a "20k-node\|passed\|failed\|Error\|exit" gpurun_out/c2c_synth_tests.log | tail -12
timeout 600 python scripts/time_ls_synth.py 2000 1,1t,4t,0t 400000 > gpurun_out/c2c_ls_synth2k.log 2>&1
timeout 900 python scripts/time_ls_synth.py 20000 1t,16t,0t 400000 > gpurun_out/c2c_ls_synth20k.log 2>&1
cat gpurun_out/c2c_ls_synth2k.log gpurun_out/c2c_ls_synth20k.log
