set -u
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu -k "suite and 2" 2>&1 | tail -40 > gpurun_out/X_multi2.txt; tail -3 gpurun_out/X_multi2.txt
TR 2 29831 bench/straggler_bench.py --rounds 48 --delays-us 250,500 --inflight 2 2>&1 | grep "^{" > gpurun_out/X_strag2_if2.json
TR 2 29832 bench/straggler_bench.py --rounds 48 --delays-us 250,500 --inflight 4 2>&1 | grep "^{" > gpurun_out/X_strag2_if4.json
