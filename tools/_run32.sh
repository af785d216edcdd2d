(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -k "fuzz" 2>&1 | tail -5)
