run() { n=$1; shift; echo "=== N=$n $*"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 30 --warmup 5 --no-e2e "$@" 2>&1 | grep "^{" | python -c "
import sys,json
d=json.loads(sys.stdin.read()); n=d['nccl'] or {}; print(d['algo'], 'ms', round(d['ms_per_step'],4), 'busbw', round(d['busbw_GBps'],1), 'nccl_ms', round(n.get('ms_per_step',0),4), 'vs_nccl', d['vs_nccl'] and round(d['vs_nccl'],3), 'verified', d['verified'], 'status', d['status_bits'])"; }
