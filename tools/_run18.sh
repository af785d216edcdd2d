set -x
mkdir -p gpurun_out
for w in 1 2 8; do
SWB_KT3_WARPS=$w NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c100-200
done
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2r_plain.log 2>&1 && \
NW_PAIRS=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:kt3_walk -s 1 -c 1 -o gpurun_out/kt3_r2r python tools/probe_configs.py nw10k > gpurun_out/r2r_ncu.log 2>&1
tail -2 gpurun_out/r2r_ncu.log
