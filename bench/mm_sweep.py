"""Sweep of the combined multimem.ld_reduce + multimem.st loop (grid x unroll x memory semantics)."""
import ctypes as C, json, os, struct, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank)); torch.cuda.set_device(lr)
sess = f"mm_{os.environ.get('MASTER_PORT','0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
nbytes = 64 << 20
symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), 2 * nbytes, lr, True)
lib = L.load(); s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
share = nbytes // world
mc = C.c_void_p(symm.mc_ptr + rank * share)
def gmax(x): return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))
res = {}
for mode, name in ((2, "relaxed"), (6, "weak")):
    for grid in (4, 8, 12, 16, 24, 32):
        for un in (4, 8, 16):
            L.check(lib.aar_bench_link(mode, grid, un, mc, None, None, share, 2, s)); torch.cuda.synchronize(); rdv.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); L.check(lib.aar_bench_link(mode, grid, un, mc, None, None, share, 5, s)); e1.record()
            torch.cuda.synchronize(); rdv.barrier()
            res[f"{name}_g{grid}_u{un}"] = round(gmax(e0.elapsed_time(e1) / 5 * 1e3), 1)
if rank == 0: print(json.dumps(res))
rdv.barrier()
