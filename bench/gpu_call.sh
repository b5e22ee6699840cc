set -u
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
TR 8 29921 bench/sweep.py --sizes 262144,1048576,2097152,4194304 --algos nvls_bulk,nvls_onehop 2>&1 | grep -E "^\{|rror" > gpurun_out/AJ_sweep8_onehop.jsonl
