#!/bin/bash
# scratch: the command list of the current gpurun call
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/AR_multi2.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/AR_multi2.txt
grep -h "FAIL\|RESULT\|public API\|never hears" gpurun_out/multi_gpu_suite_2.log | head
TR 2 29981 bench.py --gpus 2 --steps 20 --warmup 5 2> gpurun_out/AR_bench2.err | grep "^{" > gpurun_out/AR_bench2.json; python -c "
import json; d=json.load(open('gpurun_out/AR_bench2.json')); e=d['e2e']; print(d['ms_per_step'], d['vs_nccl'], e['ms_per_step'], e['vs_h2d_floor'], e['vs_nccl'], d['verified'])"
