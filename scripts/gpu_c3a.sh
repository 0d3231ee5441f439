# third chain form: parity of the new modes, per-round timing, seeded solves
mkdir -p gpurun_out
python -c "import sys; sys.path.insert(0,'.'); from vns_sop_b200 import build; build.build()" > gpurun_out/c3e_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "smem-tab" > gpurun_out/c3e_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/c3e_tests.log
tail -5 gpurun_out/c3e_tests.log
timeout 300 python scripts/time_ls.py krob200 c3,v1 > gpurun_out/c3e_ls_krob200.log 2>&1; cat gpurun_out/c3e_ls_krob200.log
timeout 300 python scripts/time_ls.py dop_s12 c3,v1 > gpurun_out/c3e_ls_dop_s12.log 2>&1; cat gpurun_out/c3e_ls_dop_s12.log
for inst in krob200 dop_s12 opn_s08 berlin52; do timeout 300 python scripts/time_solve.py $inst 0 -1 >> gpurun_out/c3e_solve.jsonl 2>> gpurun_out/c3e_solve.err; done
cat gpurun_out/c3e_solve.jsonl
