set -x
for rep in 1 2; do
NW_PAIRS=1024 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c100-200
SWB_LIB=$PWD/gpurun_ab_old.so NW_PAIRS=1024 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c100-200
done
NW_PAIRS=400 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c100-200
SWB_LIB=$PWD/gpurun_ab_old.so NW_PAIRS=400 timeout 600 python tools/probe_configs.py nw10k 2>&1 | tail -2 | head -1 | cut -c100-200
