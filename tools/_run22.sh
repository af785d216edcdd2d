set -x
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -k "wide_tile_path" 2>&1 | tail -25) > gpurun_out/r2w_pytest.log 2>&1
tail -25 gpurun_out/r2w_pytest.log | cut -c1-300
DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k 2>&1 | tail -3 | cut -c1-330
