set -x
for w in 8 12 16; do
SWB_KT3_WARPS=$w DNA_LEN=100000 timeout 600 python tools/probe_configs.py dna30k 2>&1 | tail -3 | head -2 | cut -c118-175
done
