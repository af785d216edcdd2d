set -x
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_search.py -m gpu -x -q --timeout 900 -k "search or k2 or fuzz" 2>&1 | tail -15) > gpurun_out/r2o_pytest.log 2>&1
tail -6 gpurun_out/r2o_pytest.log
timeout 600 python tools/search_bench.py --steps 3 > gpurun_out/r2o_search.log 2>&1
tail -1 gpurun_out/r2o_search.log | cut -c1-700
SWB_NO_K2P=1 timeout 600 python tools/search_bench.py --steps 2 2>&1 | tail -1 | cut -c1-400
