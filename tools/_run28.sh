set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 --no-search > gpurun_out/r2z_plain.json 2> gpurun_out/r2z_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2z_launches.csv python bench.py --steps 3 --warmup 3 --no-search > gpurun_out/r2z_ncu.log 2>&1
tail -c 300 gpurun_out/r2z_plain.err
wc -l gpurun_out/r2z_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k1_local_kernel -s 3 -c 1 -o gpurun_out/k1_r2z python bench.py --steps 3 --warmup 3 --no-search > gpurun_out/r2z_ncu2.log 2>&1
tail -2 gpurun_out/r2z_ncu2.log
