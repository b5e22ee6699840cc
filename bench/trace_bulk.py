"""Multi-GPU: per-CTA timestamps (AAR_FLAG_TRACE) of one round of the bulk kernels, summarised per role.
    python -m torch.distributed.run --nproc-per-node N ... bench/trace_bulk.py [--algo bulk|nvls_bulk] [--split a:b]"""
import argparse, json, os, struct, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.engine import RoundEngine
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous

ap = argparse.ArgumentParser()
ap.add_argument("--elems", dest="n", type=int, default=16 << 20)
ap.add_argument("--chunk", type=int, default=1 << 20)
ap.add_argument("--algo", default="bulk")
ap.add_argument("--split", default="0:0")
ap.add_argument("--overlap", action="store_true")
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
sess = f"tb_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
lay = L.compute_layout(world, a.n, a.chunk, 1, L.AAR_F32, 0, True)
symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), lay.regionBytes, lr, want_multicast=True)
eng = RoundEngine.from_symmetric(symm, lay, algo=a.algo, overlap=a.overlap)
na, nb = (int(v) for v in a.split.split(":"))
eng.bulk_split = eng.nvls_bulk_split = na | (nb << 16)
xs = [torch.randn(a.n, device=f"cuda:{lr}") for _ in range(2)]
def sync():
    torch.cuda.synchronize(); rdv.barrier()
r = 0
for _ in range(6):
    eng.enqueue_round(r, xs[r % 2]); r += 1
eng.wait_all(); sync()
eng.base_flags |= L.FLAG_TRACE
for _ in range(2 if a.overlap else 1):
    eng.enqueue_round(r, xs[r % 2]); r += 1
eng.wait_all(); sync()
tr = symm.tensor(rank, lay.traceOff, L.TRACE_CTAS * L.TRACE_MARKS * 8, torch.int64).view(L.TRACE_CTAS, L.TRACE_MARKS).cpu().double()
n = int((tr[:, 0] != 0).sum())
tr = tr[:n]
t0 = float(tr[:, 0].min())
if a.algo == "nvls_onehop":   # stage CTAs first, then reduce CTAs: print every mark over all CTAs and stop
    out = {"world": world, "algo": eng.algo, "ctas": n}
    for k, nm in {0: "entry", 1: "stage done", 2: "all flags seen", 3: "reduce done", 5: "exit"}.items():
        col = tr[:, k]
        col = col[col > 0]
        if len(col):
            col = (col - t0) / 1e3
            out[nm] = [len(col), round(float(col.min()), 1), round(float(col.median()), 1), round(float(col.max()), 1)]
    res = rdv.all_gather_bytes(json.dumps(out).encode())
    if rank == 0:
        for b in res:
            print(b.decode())
    rdv.barrier()
    sys.exit(0)
split = eng.nvls_bulk_split if a.algo == "nvls_bulk" else eng.bulk_split
# the library's default split when 0 (keep in sync with aar_launch_round_bulk)
if split == 0:
    if a.algo == "nvls_bulk":
        nb = max(1, n // 8); na = (n - nb) * 2 // 5
    else:
        na, nb = n // 5, n * 2 // 5
names = {"push/stage": (0, na, {0: "entry", 6: "first unit", 1: "done", 5: "exit"}),
         "reduce": (na, na + nb, {0: "entry", 2: "first decision", 7: "first tile stored", 6: "last load issued", 3: "stores done", 5: "exit"}),
         "gather": (na + nb, n, {0: "entry", 6: "first unit", 4: "done", 5: "exit"})}
out = {"world": world, "algo": eng.algo, "ctas": n, "split": [na, nb], "t0_us_mod_1s": round((t0 % 1e9) / 1e3, 1)}
for role, (lo, hi, marks) in names.items():
    prof = {}
    for k, nm in marks.items():
        col = tr[lo:hi, k]
        col = col[col > 0]
        if len(col):
            col = (col - t0) / 1e3
            prof[nm] = [round(float(col.min()), 1), round(float(col.median()), 1), round(float(col.max()), 1)]
    out[role] = prof
res = rdv.all_gather_bytes(json.dumps(out).encode())
if rank == 0:
    for b in res:
        print(b.decode())
rdv.barrier()
