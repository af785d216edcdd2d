# usage: bash tools/e2e_sweep.sh  -- e2e of bench.py under a few pipeline settings (tuning aid)
for cfg in "131072 393216" "65536 262144" "65536 131072" "32768 131072" "32768 65536" "65536 65536"; do
  set -- $cfg
  echo "== first $1 max $2"
  SWB_CHUNK_FIRST=$1 SWB_CHUNK_PAIRS=$2 timeout 300 python bench.py --steps 3 --warmup 3 --no-extras --cpu-seconds 0.5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])"
done
SWB_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-extras --cpu-seconds 0.5 2>&1 >/dev/null | grep "\[swb\]" | tail -6
