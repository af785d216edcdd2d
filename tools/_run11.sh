set -x
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 1400 -k "100kb_score_and_end or glocal_batch_10kb" 2>&1 | tail -15) > gpurun_out/r2l_pytest.log 2>&1
tail -5 gpurun_out/r2l_pytest.log
timeout 600 python tools/mode_probe.py > gpurun_out/r2l_modes.jsonl 2>&1; cat gpurun_out/r2l_modes.jsonl | tail -4
