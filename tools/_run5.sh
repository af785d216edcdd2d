set -x
(timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_search.py -m gpu -q --timeout 600 -k "packed_traceback or native_search" 2>&1 | tail -5)
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c1-330
NW_PAIRS=1024 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c1-330
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2f_plain.log 2>&1 && \
NW_PAIRS=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k3_fill -s 1 -c 1 -o gpurun_out/k3_r2f python tools/probe_configs.py nw10k > gpurun_out/r2f_ncu.log 2>&1
tail -3 gpurun_out/r2f_ncu.log
