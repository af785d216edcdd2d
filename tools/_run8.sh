set -x
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 2>&1 | tail -25) > gpurun_out/r2i_noalloc_alone.log 2>&1
tail -25 gpurun_out/r2i_noalloc_alone.log
(timeout 1200 python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -25) > gpurun_out/r2i_pytest.log 2>&1
tail -6 gpurun_out/r2i_pytest.log
timeout 600 python tools/exercise_kernels.py > gpurun_out/r2i_exercise.log 2>&1; tail -15 gpurun_out/r2i_exercise.log
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/exercise_kernels.py > gpurun_out/r2i_memcheck.log 2>&1; echo "memcheck rc=$?"
tail -12 gpurun_out/r2i_memcheck.log
