set -x
mkdir -p gpurun_out
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2k_nw10k.log 2>&1 && \
NW_PAIRS=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:kt3_walk -s 1 -c 1 -o gpurun_out/kt3_r2k python tools/probe_configs.py nw10k > gpurun_out/r2k_ncu.log 2>&1
tail -2 gpurun_out/r2k_nw10k.log | cut -c1-400
tail -3 gpurun_out/r2k_ncu.log
