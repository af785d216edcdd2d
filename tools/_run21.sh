mkdir -p gpurun_out
timeout 300 python tools/h2d_bw.py > gpurun_out/r2v_h2d.jsonl 2>gpurun_out/r2v_h2d.err
for n in 2 4 8; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29544 tools/h2d_bw.py >> gpurun_out/r2v_h2d.jsonl 2>>gpurun_out/r2v_h2d.err
done
cat gpurun_out/r2v_h2d.jsonl; tail -3 gpurun_out/r2v_h2d.err
