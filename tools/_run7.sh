set -x
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -15) > gpurun_out/r2h_pytest.log 2>&1
tail -4 gpurun_out/r2h_pytest.log
timeout 1200 python bench.py > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err
tail -c 600 gpurun_out/r2h_bench.err
cut -c1-1500 gpurun_out/r2h_bench.json
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2h_nw10k.log 2>&1 && \
NW_PAIRS=256 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k3_fill -s 1 -c 1 -o gpurun_out/k3_r2h python tools/probe_configs.py nw10k > gpurun_out/r2h_ncu.log 2>&1
tail -2 gpurun_out/r2h_nw10k.log | cut -c1-400
tail -3 gpurun_out/r2h_ncu.log
