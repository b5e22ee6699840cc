"""NVLink / NVLS micro-benchmark under torch.distributed.run: all ranks run the same loop at the same time.
Reports GB/s per GPU for multimem.ld_reduce, multimem.st, both, peer load, peer store and local copy over grid sizes."""
import argparse
import ctypes as C
import json
import os
import struct
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L  # noqa: E402
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mib", type=int, default=64)
ap.add_argument("--grids", default="16,32,64,148")
ap.add_argument("--unrolls", default="4,8")
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--bulk", action="store_true", help="also run the cp.async.bulk copy-role micro-benchmark (aar_bench_bulk_copy; first run on hardware pending)")
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
sess = f"lb_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
nbytes = a.mib << 20
symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), 2 * nbytes, lr, want_multicast=True)
lib = L.load()
peer = symm.ptrs[(rank + 1) % world]
local = symm.ptrs[rank]
s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
names = ["mm_ld_reduce", "mm_st", "mm_ld_reduce+st", "peer_ld", "peer_st", "local_copy"]
res = {"world": world, "MiB": a.mib, "multicast": symm.has_multicast}


def gmax(x):
    return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))


for mode in range(6):
    if mode <= 2 and not symm.has_multicast:
        continue
    # each rank reduces / stores only its 1/world share in the multimem modes (like the allreduce does)
    share = nbytes // world if mode <= 2 else nbytes
    for grid in [int(g) for g in a.grids.split(",")]:
        for un in [int(u) for u in a.unrolls.split(",")]:
            off = (rank * share) if mode <= 2 else 0
            mc = C.c_void_p(symm.mc_ptr + off) if symm.mc_ptr else None
            L.check(lib.aar_bench_link(mode, grid, un, mc, C.c_void_p(peer), C.c_void_p(local), share, 2, s))
            torch.cuda.synchronize(); rdv.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            L.check(lib.aar_bench_link(mode, grid, un, mc, C.c_void_p(peer), C.c_void_p(local), share, a.iters, s))
            e1.record()
            torch.cuda.synchronize(); rdv.barrier()
            us = gmax(e0.elapsed_time(e1) / a.iters * 1e3)
            res[f"{names[mode]}_g{grid}_u{un}"] = {"us": round(us, 1), "GBps_per_gpu": round(share / us / 1e3, 1)}
# all-to-all patterns: every GPU moves `nbytes` in total, nbytes / (world - 1) per peer
if world > 2 or True:
    others = [symm.ptrs[(rank + k) % world] for k in range(1, world)]
    arr = (C.c_void_p * len(others))(*others)
    per_peer = (nbytes // len(others)) // 4096 * 4096
    a2a = ["a2a_st_one_peer_per_cta", "a2a_st_all_peers_per_cta", "a2a_ld_one_peer_per_cta", "a2a_st32B_one_peer_per_cta"]
    for mode in range(4):
        for grid in [int(g) for g in a.grids.split(",")]:
            for un in (4, 8):
                L.check(lib.aar_bench_all2all(mode, grid, un, arr, len(others), C.c_void_p(local), per_peer, 2, s))
                torch.cuda.synchronize(); rdv.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                L.check(lib.aar_bench_all2all(mode, grid, un, arr, len(others), C.c_void_p(local), per_peer, a.iters, s))
                e1.record()
                torch.cuda.synchronize(); rdv.barrier()
                us = gmax(e0.elapsed_time(e1) / a.iters * 1e3)
                res[f"{a2a[mode]}_g{grid}_u{un}"] = {"us": round(us, 1), "GBps_per_gpu": round(per_peer * len(others) / us / 1e3, 1)}
# register-free copy role: one issuing thread per CTA, bulk async-copy engine, local -> local and local -> peer
if a.bulk:
    err = torch.zeros(1, dtype=torch.int32, device=f"cuda:{lr}")
    src_view = symm.tensor(rank, 0, nbytes, torch.int32)
    src_view.copy_(torch.arange(nbytes // 4, dtype=torch.int32, device=f"cuda:{lr}"))
    torch.cuda.synchronize(); rdv.barrier()
    for name, dst in (("bulk_local", local + nbytes), ("bulk_peer_st", peer + nbytes)):
        for stages, tile in ((4, 16384), (4, 32768), (8, 16384)):
            for grid in (16, 32, 74, 148, 296):
                args = (grid, stages, tile, C.c_void_p(local), C.c_void_p(dst), nbytes)
                L.check(lib.aar_bench_bulk_copy(*args, 1, s, C.c_void_p(err.data_ptr())))
                torch.cuda.synchronize(); rdv.barrier()
                ok = True
                if name == "bulk_local":   # verify once per configuration on the local destination
                    ok = bool(torch.equal(symm.tensor(rank, nbytes, nbytes, torch.int32), src_view))
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                L.check(lib.aar_bench_bulk_copy(*args, a.iters, s, C.c_void_p(err.data_ptr())))
                e1.record()
                torch.cuda.synchronize(); rdv.barrier()
                us = gmax(e0.elapsed_time(e1) / a.iters * 1e3)
                res[f"{name}_s{stages}_t{tile}_g{grid}"] = {"us": round(us, 1), "GBps_per_gpu": round(nbytes / us / 1e3, 1), "ok": ok, "gave_up": int(err.item())}
if rank == 0:
    print(json.dumps(res))
rdv.barrier()
