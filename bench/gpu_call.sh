#!/bin/bash
# scratch: the command list of the current gpurun call
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
for rep in 1 2; do
for prio in default high; do
  AAR_STREAM_PRIORITY=$prio TR 4 2999$rep bench.py --gpus 4 --steps 30 --warmup 5 --no-e2e --no-nccl 2>/dev/null | grep "^{" | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$prio', d['ms_per_step'], d['algo'])"
done; done
