mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_best.py tests/test_synthetic.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
