set -x
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
(timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -6) > gpurun_out/r3a_pytest.log 2>&1
tail -3 gpurun_out/r3a_pytest.log
timeout 1200 python bench.py > gpurun_out/r3a_bench.json 2> gpurun_out/r3a_bench.err
tail -c 300 gpurun_out/r3a_bench.err
cut -c1-400 gpurun_out/r3a_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-600
