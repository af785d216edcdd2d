set -x
(timeout 900 python -m pytest tests/test_gpu_search.py tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -k "search or packed_traceback" 2>&1 | tail -3)
timeout 600 python tools/search_bench.py --steps 3 2>&1 | tail -1 | cut -c150-420
