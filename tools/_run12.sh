set -x
mkdir -p gpurun_out
timeout 600 python tools/search_bench.py --steps 1 > gpurun_out/r2m_search.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2m_search_launches.csv python tools/search_bench.py --steps 1 > gpurun_out/r2m_ncu.log 2>&1
tail -1 gpurun_out/r2m_search.log | cut -c1-600
grep -v "sel_\|zero_kernel\|scan_kernel" gpurun_out/r2m_search_launches.csv | awk -F'","' '{print $5, $(NF)}' | tail -40
