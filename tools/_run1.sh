set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
(time python -m pytest tests -m gpu -q --timeout 1200 2>&1 | tail -40) > gpurun_out/r2a_pytest.log 2>&1
NOCHECK=1 bash tools/ab_k1.sh ab/lib_base.so ab/lib_u2.so ab/lib_vote.so > gpurun_out/r2a_ab_k1.txt 2>&1
cat gpurun_out/r2a_ab_k1.txt
python tools/search_bench.py --steps 2 > gpurun_out/r2a_search_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k2_search -s 1 -c 1 -o gpurun_out/k2_r2a python tools/search_bench.py --steps 1 > gpurun_out/r2a_ncu.log 2>&1
tail -3 gpurun_out/r2a_search_plain.log
tail -5 gpurun_out/r2a_ncu.log
cat gpurun_out/r2a_pytest.log
