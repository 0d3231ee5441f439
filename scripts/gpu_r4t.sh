# r4t: whole GPU suite + smoke at the round's last commit
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r4t_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4t_gpu_tests.log
tail -6 gpurun_out/r4t_gpu_tests.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
