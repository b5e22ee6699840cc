#!/bin/bash
# scratch: the command list of the current gpurun call
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
TR 2 29971 bench/api_rate.py > gpurun_out/AQ_api_rate2.json 2> gpurun_out/AQ_api_rate2.err; tail -c 1500 gpurun_out/AQ_api_rate2.json; tail -3 gpurun_out/AQ_api_rate2.err
