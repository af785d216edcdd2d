set -x
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_search.py -m gpu -x -q --timeout 900 -k "packed_traceback or search or fuzz or config5" 2>&1 | tail -15) > gpurun_out/r2n_pytest.log 2>&1
tail -6 gpurun_out/r2n_pytest.log
timeout 600 python tools/search_bench.py --steps 3 > gpurun_out/r2n_search.log 2>&1
tail -1 gpurun_out/r2n_search.log | cut -c1-700
