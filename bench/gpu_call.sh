#!/bin/bash
# scratch: the command list of the current gpurun call
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
TR 4 29941 bench/gpt2_dp.py --steps 12 --warmup 4 --configs auto:0,auto:72 > gpurun_out/AM_gpt2_4gpu.json 2> gpurun_out/AM_gpt2_4gpu.err
tail -c 2500 gpurun_out/AM_gpt2_4gpu.json; tail -3 gpurun_out/AM_gpt2_4gpu.err
