set -x
mkdir -p gpurun_out
DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k > gpurun_out/r2y_plain.log 2>&1 && \
DNA_LEN=100000 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k3w_fill -s 1 -c 1 -o gpurun_out/k3w_r2y python tools/probe_configs.py dna30k > gpurun_out/r2y_ncu.log 2>&1
tail -3 gpurun_out/r2y_plain.log | head -2 | cut -c118-175
tail -2 gpurun_out/r2y_ncu.log
