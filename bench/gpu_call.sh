set -u
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu -k "suite and 2" 2>&1 | tail -30 > gpurun_out/AE_multi2.txt; tail -3 gpurun_out/AE_multi2.txt
TR 2 29841 bench.py --gpus 2 --steps 40 --warmup 5 --no-e2e 2>gpurun_out/AE_bench2.err | grep "^{" > gpurun_out/AE_bench2.json
AAR_NATIVE_ENGINE=0 TR 2 29842 bench.py --gpus 2 --steps 40 --warmup 5 --no-e2e --no-nccl 2>gpurun_out/AE_bench2_py.err | grep "^{" > gpurun_out/AE_bench2_py.json
AAR_NATIVE_ENGINE=0 timeout 300 python -m pytest tests/test_gpu_local.py -x -q -m gpu -k "snapshot or strict or start_at or retires" 2>&1 | tail -3
