set -x
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2c_plain1.log 2>&1 && \
NW_PAIRS=256 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2c_launches_nw10k.csv python tools/probe_configs.py nw10k > gpurun_out/r2c_ncu1.log 2>&1
DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k > gpurun_out/r2c_plain2.log 2>&1 && \
DNA_LEN=100000 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2c_launches_dna100k.csv python tools/probe_configs.py dna30k > gpurun_out/r2c_ncu2.log 2>&1
SWB_KT_THREAD=1 DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k > gpurun_out/r2c_dna100k_ktthread.log 2>&1; tail -3 gpurun_out/r2c_dna100k_ktthread.log
grep -c . gpurun_out/r2c_launches_nw10k.csv gpurun_out/r2c_launches_dna100k.csv
