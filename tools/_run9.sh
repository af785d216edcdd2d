# usage: bash tools/_run9.sh N   -- bench.py under torchrun on N GPUs of one box (search block included)
set -x
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2j_bench_n$N.json 2> gpurun_out/r2j_bench_n$N.err
echo "rc=$?"
tail -c 1500 gpurun_out/r2j_bench_n$N.err
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2j_bench_n$N.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "n_gpus", "ms_per_step", "e2e")})
    print(json.dumps(d.get("search"))[:1800])
except Exception as ex:
    print("no line", ex)
PY
