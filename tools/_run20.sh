mkdir -p gpurun_out
cp gpurun_dbg_libswb200.so smithwaterman_b200/libswb200.so
NW_PAIRS=1 timeout 300 python tools/probe_configs.py nw10k > gpurun_out/r2s_dbg.log 2>&1
head -150 gpurun_out/r2s_dbg.log | cut -c1-140
