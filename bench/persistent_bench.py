"""Small-message latency: resident doorbell-driven engine (K7) vs one launch per round vs NCCL.  Run under torch.distributed.run.
Host-timed for K7 (there is no launch to put events around), event-timed for the other two; all ranks synchronised on both sides."""
import json, os, struct, sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.engine import ONESHOT_MAX_BYTES, PersistentEngine, RoundEngine
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank)); torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
sess = f"pbn_{os.environ.get('MASTER_PORT','0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
def gmax(x): return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))
dist.init_process_group("nccl", device_id=dev)
res = {"world": world}
iters = 200
base = 0
for nbytes in (4096, 65536, 262144):
    n = nbytes // 4
    x = torch.randn(n, device=dev)
    buf = x.clone()
    for _ in range(10): dist.all_reduce(buf)
    torch.cuda.synchronize(); rdv.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): dist.all_reduce(buf)
    e1.record(); torch.cuda.synchronize(); rdv.barrier()
    nccl_us = gmax(e0.elapsed_time(e1) / iters * 1e3)
    lay = L.compute_layout(world, n, max(1, n // world), 1, L.AAR_F32, ONESHOT_MAX_BYTES // 4, False)
    symm = SymmetricMemory(UnixRendezvous(rank, world, f"{sess}_{nbytes}"), lay.regionBytes, lr, want_multicast=False)
    eng = RoundEngine.from_symmetric(symm, lay, algo="oneshot", timeout_ms=5000.0)
    r = 0
    for _ in range(10):
        eng.enqueue_round(r, x); r += 1
    eng.wait_all(); torch.cuda.synchronize(); rdv.barrier()
    e0.record()
    for _ in range(iters):
        eng.enqueue_round(r, x); r += 1
    eng.wait_all(); e1.record(); torch.cuda.synchronize(); rdv.barrier()
    launched_us = gmax(e0.elapsed_time(e1) / iters * 1e3)
    # resident engine: nothing below touches the device API while it runs
    torch.cuda.synchronize(); rdv.barrier()
    pe = PersistentEngine(eng, algo="oneshot", grid=8)
    for _ in range(10):
        pe.post(r, x); r += 1
    pe.wait(r - 1)
    rdv.barrier()
    t0 = time.perf_counter()
    for _ in range(iters):
        pe.post(r, x); r += 1
    pe.wait(r - 1)
    dt = time.perf_counter() - t0
    rdv.barrier()
    pe.stop()
    pers_us = gmax(dt / iters * 1e6)
    res[str(nbytes)] = {"nccl_us": round(nccl_us, 2), "launched_oneshot_us": round(launched_us, 2), "persistent_oneshot_us": round(pers_us, 2)}
    eng.close(); symm.close()
if rank == 0: print(json.dumps(res), flush=True)
rdv.barrier()
os._exit(0)
