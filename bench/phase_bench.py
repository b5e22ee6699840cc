"""Multi-GPU phase timing (run under torch.distributed.run): each round's phases are launched as separate kernels with
CUDA events in between, for the two-shot P2P path and the NVLS path, plus the fused single-launch time and NCCL.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench/phase_bench.py [--n ..] [--chunk ..]
"""
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
from akka_allreduce_b200.parallel.engine import RoundEngine  # noqa: E402
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=16 * 1024 * 1024)
    ap.add_argument("--chunk", type=int, default=1024 * 1024)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--grids", default="0")
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--no-doorbell", action="store_true")
    ap.add_argument("--nreduce", default="24")
    ap.add_argument("--skip-split", action="store_true")
    ap.add_argument("--splits", default="0:0")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dtype = torch.float32 if a.dtype == "fp32" else torch.bfloat16
    es = 4 if a.dtype == "fp32" else 2
    sess = f"pb_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
    rdv = UnixRendezvous(rank, world, sess + "_c")
    lay = L.compute_layout(world, a.n, a.chunk, 1, L.AAR_F32 if a.dtype == "fp32" else L.AAR_BF16, 0, True)
    symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), lay.regionBytes, lr, want_multicast=True)
    xs = [torch.randn(a.n, device=dev).to(dtype) for _ in range(2)]
    lib = L.load()
    res = {"world": world, "n": a.n, "chunk": a.chunk, "bytes": a.n * es, "multicast": symm.has_multicast}

    def gmax(x):
        return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))

    def sync():
        torch.cuda.synchronize()
        rdv.barrier()

    rnd = [0]
    for grid in [int(g) for g in a.grids.split(",")]:
        variants = [("twoshot", 0)] + [("pipe", (int(x.split(":")[0]) | (int(x.split(":")[1]) << 16))) for x in a.splits.split(",")]
        variants += ([("nvls", int(x)) for x in a.nreduce.split(",")] if symm.has_multicast else [])
        for algo, nred in variants:
            tag = algo if algo == "twoshot" else (f"nvls_r{nred}" if algo == "nvls" else f"pipe_{nred & 0xffff}_{nred >> 16}")
            eng = RoundEngine.from_symmetric(symm, lay, grid=grid, algo="twoshot" if algo == "pipe" else algo, overlap=False)
            if a.no_doorbell:
                eng.params.doorbell = None
            launcher = lib.aar_launch_nvls if algo == "nvls" else None
            s = C.c_void_p(torch.cuda.current_stream().cuda_stream)

            def phases(mask):
                if algo == "nvls":
                    L.check(lib.aar_launch_nvls(eng._pref, eng._lay, mask | (nred << 8), grid, s))
                elif algo == "pipe" and mask == 7:
                    L.check(lib.aar_launch_round_pipelined(eng._pref, eng._lay, grid, nred, s))
                else:
                    fn = {1: lib.aar_launch_scatter, 2: lib.aar_launch_reduce_bcast, 4: lib.aar_launch_gather, 7: lib.aar_launch_round_fused}[mask]
                    L.check(fn(eng._pref, eng._lay, grid, s))

            def prep(r):
                p = eng.params
                p.round = r
                p.input = xs[r % 2].data_ptr()
                p.output = eng.outputs[r % 2].data_ptr()
                p.outCounts = eng.counts[r % 2].data_ptr()

            # fused
            for _ in range(3):
                prep(rnd[0]); phases(7); rnd[0] += 1
            sync()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.iters):
                prep(rnd[0]); phases(7); rnd[0] += 1
            e1.record()
            sync()
            res[f"{tag}_g{grid}_fused_us"] = gmax(e0.elapsed_time(e1) / a.iters * 1e3)
            # view-output fused
            eng.params.flags |= L.FLAG_VIEW_OUTPUT
            for _ in range(2):
                prep(rnd[0]); phases(7); rnd[0] += 1
            sync()
            e0.record()
            for _ in range(a.iters):
                prep(rnd[0]); phases(7); rnd[0] += 1
            e1.record()
            sync()
            res[f"{tag}_g{grid}_fused_view_us"] = gmax(e0.elapsed_time(e1) / a.iters * 1e3)
            eng.params.flags &= ~L.FLAG_VIEW_OUTPUT
            # traced fused round: where does each CTA spend its time?
            eng.params.flags |= L.FLAG_TRACE
            sync()
            prep(rnd[0]); phases(7); rnd[0] += 1
            sync()
            eng.params.flags &= ~L.FLAG_TRACE
            tr = symm.tensor(rank, lay.traceOff, L.TRACE_CTAS * L.TRACE_MARKS * 8, torch.int64).view(L.TRACE_CTAS, L.TRACE_MARKS).cpu()
            ncta = int((tr[:, 0] != 0).sum())
            tr = tr[:ncta].double()
            t0 = tr[:, 0].min()
            names = ["entry", "K1done", "K2firstDecision", "K2done", "K3decision", "exit"]
            prof = {}
            for k, nm in enumerate(names):
                col = (tr[:, k] - t0) / 1e3
                prof[nm] = [round(float(col.min()), 1), round(float(col.median()), 1), round(float(col.max()), 1)]
            res[f"{tag}_g{grid}_trace_us_min_med_max"] = prof
            symm.tensor(rank, lay.traceOff, L.TRACE_CTAS * L.TRACE_MARKS * 8, torch.int64).zero_()
            # split phases
            acc = {1: 0.0, 2: 0.0, 4: 0.0}
            for it in range(0 if a.skip_split else a.iters + 2):
                prep(rnd[0])
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                sync()
                ev[0].record(); phases(1); ev[1].record(); phases(2); ev[2].record(); phases(4); ev[3].record()
                rnd[0] += 1
                sync()
                if it >= 2:
                    for k, m in enumerate((1, 2, 4)):
                        acc[m] += ev[k].elapsed_time(ev[k + 1]) * 1e3 / a.iters
            for m, name in ((1, "K1"), (2, "K2"), (4, "K3")):
                if not a.skip_split:
                    res[f"{tag}_g{grid}_{name}_us"] = gmax(acc[m])
            eng.close()

    # NCCL
    import torch.distributed as dist

    dist.init_process_group("nccl", device_id=dev)
    b = [x.clone() for x in xs]
    for i in range(5):
        dist.all_reduce(b[i % 2])
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.iters):
        dist.all_reduce(b[i % 2])
    e1.record()
    sync()
    res["nccl_us"] = gmax(e0.elapsed_time(e1) / a.iters * 1e3)
    if rank == 0:
        print(json.dumps(res))
    rdv.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
