set -u
mkdir -p gpurun_out
TR() { n=$1; port=$2; shift 2; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
nvidia-smi nvlink -gt d -i 0 2>&1 | head -8 > gpurun_out/AF_nvlink_raw.txt
TR 2 29861 bench/nvlink_bytes.py --algo bulk 2>&1 | grep "^\[" > gpurun_out/AF_nvlink_bulk2.json
TR 2 29862 bench/nvlink_bytes.py --algo nvls_bulk 2>&1 | grep "^\[" > gpurun_out/AF_nvlink_nvls2.json
