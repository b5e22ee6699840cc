#!/bin/bash
# scratch: the command list of the current gpurun call
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
TR 2 29951 bench/gpt2_dp.py --steps 12 --warmup 4 --configs auto:0,auto:12:high:3:5,auto:16:high:4:7,auto:20:high:5:9,auto:24:high:6:11,auto:32:high:8:14,auto:48:high:12:22 > gpurun_out/AN_gpt2_2gpu_split.json 2> gpurun_out/AN_gpt2_2gpu_split.err
tail -c 3000 gpurun_out/AN_gpt2_2gpu_split.json; tail -3 gpurun_out/AN_gpt2_2gpu_split.err
