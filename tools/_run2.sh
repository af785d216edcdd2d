set -x
(time timeout 1500 python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -60) > gpurun_out/r2b_pytest.log 2>&1
cat gpurun_out/r2b_pytest.log
NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2b_nw10k_k3.log 2>&1; cat gpurun_out/r2b_nw10k_k3.log | tail -3
SWB_NO_K3=1 NW_PAIRS=256 timeout 600 python tools/probe_configs.py nw10k > gpurun_out/r2b_nw10k_kw.log 2>&1; cat gpurun_out/r2b_nw10k_kw.log | tail -3
DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k > gpurun_out/r2b_dna100k.log 2>&1; cat gpurun_out/r2b_dna100k.log | tail -4
timeout 900 python bench.py > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -c 7000 gpurun_out/r2b_bench.json; tail -5 gpurun_out/r2b_bench.err
