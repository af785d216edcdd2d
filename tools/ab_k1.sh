# usage: bash tools/ab_k1.sh lib1.so lib2.so ...  -- device / e2e GCUPS of bench.py for differently compiled builds (SWB_LIB)
for so in "$@"; do
  for rep in 1 2; do
    SWB_LIB=$so SWB_BENCH_NOCHECK=$NOCHECK timeout 300 python bench.py --steps 10 --warmup 3 --no-extras --no-search --cpu-seconds 0.3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$so', 'device %.1f GCUPS  %.3f ms  frac %.4f  e2e %.1f' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value']))"
  done
done
