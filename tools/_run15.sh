set -x
mkdir -p gpurun_out
timeout 600 python tools/search_bench.py --steps 1 > gpurun_out/r2p_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k2p_search -s 1 -c 1 -o gpurun_out/k2p_r2p python tools/search_bench.py --steps 1 > gpurun_out/r2p_ncu.log 2>&1
tail -3 gpurun_out/r2p_ncu.log
