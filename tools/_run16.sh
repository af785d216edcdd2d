set -x
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_search.py -m gpu -x -q --timeout 900 -k "packed_traceback or search or fuzz or config5 or glocal_batch" 2>&1 | tail -15) > gpurun_out/r2q_pytest.log 2>&1
tail -6 gpurun_out/r2q_pytest.log
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c1-330
NW_PAIRS=1024 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c1-330
timeout 600 python tools/search_bench.py --steps 3 2>&1 | tail -1 | cut -c1-500
