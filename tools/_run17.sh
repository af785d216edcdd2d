set -x
mkdir -p gpurun_out
for w in 4 8 12 16; do
SWB_KT3_WARPS=$w NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c100-200
done
(timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_search.py -m gpu -x -q --timeout 900 -k "packed_traceback or native_search" 2>&1 | tail -4)
SWB_KT3_WARPS=16 timeout 600 python tools/search_bench.py --steps 3 2>&1 | tail -1 | cut -c150-420
SWB_KT3_WARPS=4 timeout 600 python tools/search_bench.py --steps 3 2>&1 | tail -1 | cut -c150-420
