set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_local.py tests/test_gpu_strict.py -x -q -m gpu 2>&1 | tail -30 > gpurun_out/O_local.txt
tail -5 gpurun_out/O_local.txt
