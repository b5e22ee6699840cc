"""NVLink traffic per round, read from the link counters (`nvidia-smi nvlink -gt d`) around K rounds, against the model:
two-shot P2P moves 2 N (P-1)/P bytes out of (and into) every GPU per round, the in-switch path N (1 + 1/P).
    python -m torch.distributed.run --nproc-per-node N ... bench/nvlink_bytes.py [--algo bulk|nvls_bulk] [--rounds 200]"""
import argparse, json, os, re, subprocess, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.engine import RoundEngine
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=16 << 20)
ap.add_argument("--chunk", type=int, default=1 << 20)
ap.add_argument("--algo", default="auto")
ap.add_argument("--rounds", type=int, default=200)
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
sess = f"nvb_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
lay = L.compute_layout(world, a.n, a.chunk, 1, L.AAR_F32, 0, True)
symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), lay.regionBytes, lr, want_multicast=True)
eng = RoundEngine.from_symmetric(symm, lay, algo=a.algo)
xs = [torch.randn(a.n, device=f"cuda:{lr}") for _ in range(2)]


def counters():
    """Sum over this GPU's links of (tx KiB, rx KiB); None if the tool does not answer."""
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(lr)], capture_output=True, text=True, timeout=20).stdout
    except Exception:
        return None
    tx = sum(int(v) for v in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(v) for v in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return (tx, rx) if (tx or rx) else None


r = 0
for _ in range(6):
    eng.enqueue_round(r, xs[r % 2]); r += 1
eng.wait_all(); torch.cuda.synchronize(); rdv.barrier()
c0 = counters()
rdv.barrier()
for _ in range(a.rounds):
    eng.enqueue_round(r, xs[r % 2]); r += 1
eng.wait_all(); torch.cuda.synchronize(); rdv.barrier()
c1 = counters()
nbytes = a.n * 4
model = nbytes * (1 + 1 / world) if eng.algo in ("nvls", "nvls_bulk", "nvls_onehop") else 2 * nbytes * (world - 1) / world
res = {"rank": rank, "world": world, "algo": eng.algo, "bytes": nbytes, "rounds": a.rounds, "model_bytes_per_round_per_direction": model}
if c0 and c1:
    res["tx_bytes_per_round"] = (c1[0] - c0[0]) * 1024 / a.rounds
    res["rx_bytes_per_round"] = (c1[1] - c0[1]) * 1024 / a.rounds
    res["tx_over_model"] = round(res["tx_bytes_per_round"] / model, 3)
    res["rx_over_model"] = round(res["rx_bytes_per_round"] / model, 3)
else:
    res["error"] = "nvidia-smi nvlink -gt d gave no counters on this box"
allr = rdv.all_gather_bytes(json.dumps(res).encode())
if rank == 0:
    print(json.dumps([json.loads(b) for b in allr]))
rdv.barrier()
eng.close()
