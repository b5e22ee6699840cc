set -u
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
timeout 600 python -m pytest tests/test_gpu_local.py tests/test_gpu_strict.py -x -q -m gpu 2>&1 | tail -5 > gpurun_out/T_local.txt
tail -2 gpurun_out/T_local.txt
TR 2 29801 bench/trace_bulk.py --split 16:72 > gpurun_out/T_trace2.txt 2>&1
i=0
for sp in 0:0 16:90 20:90; do
  i=$((i+1))
  TR 2 $((29710+i)) bench.py --gpus 2 --steps 50 --warmup 5 --no-e2e --no-nccl --algo bulk --split $sp 2>&1 | grep "^{" > gpurun_out/T_bench2_bulk_${sp/:/_}.json
done
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu -k "suite and 2" 2>&1 | tail -40 > gpurun_out/T_multi2.txt
tail -3 gpurun_out/T_multi2.txt
