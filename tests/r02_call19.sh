timeout 500 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "fuzz" 2>&1 | tail -n 12
