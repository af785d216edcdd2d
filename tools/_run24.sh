set -x
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -k "wide_tile_path" 2>&1 | tail -5)
DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k 2>&1 | tail -3 | head -2 | cut -c90-260
DNA_LEN=100000 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2x_launches.csv python tools/probe_configs.py dna30k > /dev/null 2>&1
grep "k3w_fill\|kt3_walk\|k3w_pick" gpurun_out/r2x_launches.csv | awk -F'","' '{print substr($5,1,40), $(NF)}' | tail -5
