"""Two GPUs: can a fence-less flag written after cp.async.bulk.wait_group overtake the bulk-stored data on its way to the peer?
    python -m torch.distributed.run --nproc-per-node 2 ... bench/litmus_bulk_flag.py [--iters 20000]"""
import argparse, ctypes as C, json, os, struct, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=20000)
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
assert world == 2
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
sess = f"lit_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
GRID_MAX, BYTES_MAX = 64, 1 << 20
words_off = GRID_MAX * BYTES_MAX
symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), words_off + GRID_MAX * 32 * 4, lr, want_multicast=False)
lib = L.load()
s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
viol = torch.zeros(1, dtype=torch.int64, device=f"cuda:{lr}")
gave = torch.zeros(1, dtype=torch.int32, device=f"cuda:{lr}")
res = {}
for grid, nbytes in ((1, 32768), (8, 262144), (32, 262144), (64, 1 << 20), (64, 32768)):
    for mode, name in ((1, "no_wait_control"), (0, "wait_group_then_flag"), (2, "wait_group_fence_flag")):
        iters = a.iters if nbytes <= 262144 else a.iters // 4
        symm.tensor(rank, 0, words_off + GRID_MAX * 32 * 4, torch.uint8).zero_()
        viol.zero_(); gave.zero_()
        torch.cuda.synchronize(); rdv.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.check(lib.aar_bench_litmus_bulk_flag(1 if rank == 0 else 0, mode, grid, C.c_void_p(symm.ptrs[1 - rank]), C.c_void_p(symm.ptrs[rank]),
                                               C.c_void_p(symm.ptrs[1 - rank] + words_off), C.c_void_p(symm.ptrs[rank] + words_off), nbytes, iters,
                                               C.c_void_p(viol.data_ptr()), C.c_void_p(gave.data_ptr()), s))
        e1.record()
        torch.cuda.synchronize(); rdv.barrier()
        mine = struct.pack("qid", int(viol.item()), int(gave.item()), e0.elapsed_time(e1))
        allr = [struct.unpack("qid", b) for b in rdv.all_gather_bytes(mine)]
        res[f"g{grid}_b{nbytes}_{name}"] = {"rounds": iters * grid, "bad_words_seen_by_receiver": allr[1][0], "gave_up": allr[0][1] + allr[1][1],
                                           "us_per_round": round(allr[0][2] * 1e3 / iters, 2)}
if rank == 0:
    print(json.dumps(res))
rdv.barrier()
