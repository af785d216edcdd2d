for last in 0 65536 32768; do
  echo "== last $last"
  SWB_CHUNK_LAST=$last timeout 300 python bench.py --steps 3 --warmup 3 --no-extras --no-search --cpu-seconds 0.5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])"
done
