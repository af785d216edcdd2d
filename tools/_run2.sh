set -x
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_search.py -m gpu -q --timeout 600 -k "packed_traceback or native_search or search_sharded or search_overflow" 2>&1 | tail -40) > gpurun_out/r2b_pytest.log 2>&1
cat gpurun_out/r2b_pytest.log
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2b_nw10k_k3.log 2>&1; cat gpurun_out/r2b_nw10k_k3.log | tail -3
SWB_NO_K3=1 NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2b_nw10k_kw.log 2>&1; cat gpurun_out/r2b_nw10k_kw.log | tail -3
timeout 900 python bench.py > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -c 6000 gpurun_out/r2b_bench.json; tail -5 gpurun_out/r2b_bench.err
