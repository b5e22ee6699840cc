set -u
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q -m gpu -k "suite and 8" 2>&1 | tail -5
