set -x
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_search.py -m gpu -q --timeout 600 -k "packed_traceback or native_search or search_sharded or search_overflow or checkpointed or config4 or config5 or emboss_goldens or rust_kats or fuzz" 2>&1 | tail -30) > gpurun_out/r2d_pytest.log 2>&1
cat gpurun_out/r2d_pytest.log | tail -15
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | cut -c1-330
NW_PAIRS=1024 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | cut -c1-330
SWB_K3_TEAM=12 NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c1-330
SWB_K3_TEAM=4 NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c1-330
DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k 2>&1 | tail -3 | cut -c1-330
(cd ab/r1tree && DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k 2>&1 | tail -3 | cut -c1-330)
