mkdir -p gpurun_out
timeout 300 python scripts/time_ls.py krob200 c3 2>&1 | tail -4
for inst in krob200 dop_s12 opn_s08; do timeout 300 python scripts/time_solve.py $inst 0 -1 2>/dev/null | cut -c1-200; done
