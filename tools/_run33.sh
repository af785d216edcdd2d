mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -k "20kb_default_routing" 2>&1 | tail -5)
timeout 900 python bench.py --steps 3 --warmup 3 --no-search > gpurun_out/r3b_bench.json 2> gpurun_out/r3b_bench.err; tail -c 200 gpurun_out/r3b_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r3b_bench.json").read().strip().splitlines()[-1])
for k,v in d["other_configs"].items(): print(k, {a:(round(b,2) if isinstance(b,float) else b) for a,b in v.items()})
PY
