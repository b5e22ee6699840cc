import os, sys, struct, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.engine import RoundEngine
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank)); torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
sess = f"dr_{os.environ.get('MASTER_PORT','0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
n, chunk, maxlag = 1 << 24, 1 << 20, 4
lay = L.compute_layout(world, n, chunk, maxlag, L.AAR_F32, 0, False)
symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), lay.regionBytes, lr, want_multicast=False)
xs = [torch.randn(n, device=dev) for _ in range(2)]
def words(off, cnt, dt=torch.int32): return symm.tensor(rank, off, cnt * (4 if dt == torch.int32 else 8), dt).cpu().tolist()
def report(tag):
    torch.cuda.synchronize(); rdv.barrier()
    print(f"[{rank}] {tag}: peerBase={words(lay.peerBaseOff, world)} status={words(lay.statusOff, 2)} k3ctr={words(lay.k3CtrOff, lay.L)} k1ctr_sum={sum(words(lay.k1CtrOff, lay.L*world*lay.C))} k2ctr_sum={sum(words(lay.k2CtrOff, lay.L*lay.C))}", flush=True)
    rdv.barrier()
base = 0
for phase, (th, overlap) in enumerate([(0.75, None), (0.75, None), (0.75, False), (0.75, False), (1.0, None), (0.75, None)]):
    eng = RoundEngine.from_symmetric(symm, lay, th_reduce=th, th_complete=th, timeout_ms=1000.0, overlap=overlap)
    t = time.time()
    for r in range(base, base + 40):
        eng.enqueue_round(r, xs[r % 2])
    eng.wait_all(); torch.cuda.synchronize()
    dt = time.time() - t
    report(f"phase {phase} th={th} overlap={overlap} rounds {base}..{base+39} wall {dt:.3f}s")
    base += 40
    eng.close()
