mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_synthetic.py -m gpu -x -q 2>&1 | tail -5
