set -x
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -k "wide_tile_path or 100kb_properties" 2>&1 | tail -3)
DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k 2>&1 | tail -3 | head -2 | cut -c118-175
DNA_LEN=30000 timeout 600 python tools/probe_configs.py dna30k 2>&1 | tail -3 | head -2 | cut -c100-175
