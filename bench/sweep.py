"""Bandwidth sweep (BASELINE.json config 5): allreduce bus GB/s vs message size, this framework (auto-selected kernel and
every forced kernel that applies) against NCCL on the same box.  Run under torch.distributed.run.

    ... bench/sweep.py --sizes 4096,65536,1048576,16777216,67108864,268435456,1073741824 [--chunks 1048576]
"""
import argparse
import json
import os
import struct
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L  # noqa: E402
from akka_allreduce_b200.parallel.engine import ONESHOT_MAX_BYTES, RoundEngine  # noqa: E402
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", default="4096,16384,65536,262144,1048576,4194304,16777216,67108864,268435456,1073741824", help="bytes")
ap.add_argument("--chunks", default="1048576", help="maxChunkSize values (elements)")
ap.add_argument("--algos", default="auto,oneshot,twoshot,pipelined,nvls")
ap.add_argument("--dtype", default="fp32")
ap.add_argument("--grid", type=int, default=0, help="CTAs per launch (0 = one per SM)")
ap.add_argument("--write-table", default="", help="JSON file to merge this sweep's winners into (akka_allreduce_b200/parallel/selector_table.json format)")
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
sess = f"sw_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
dtype = torch.float32 if a.dtype == "fp32" else torch.bfloat16
es = 4 if a.dtype == "fp32" else 2
code = L.AAR_F32 if a.dtype == "fp32" else L.AAR_BF16


def gmax(x):
    return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))


def sync():
    torch.cuda.synchronize()
    rdv.barrier()


import torch.distributed as dist  # noqa: E402

dist.init_process_group("nccl", device_id=dev)
rows = []
k = 0
for nbytes in [int(s) for s in a.sizes.split(",")]:
    n = nbytes // es
    iters = 50 if nbytes <= (1 << 22) else (20 if nbytes <= (1 << 28) else 8)
    xs = [torch.randn(n, device=dev).to(dtype) * (rank + 1) for _ in range(2)]
    ref = None
    # NCCL
    bufs = [x.clone() for x in xs]
    for i in range(5):
        dist.all_reduce(bufs[i % 2])
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        dist.all_reduce(bufs[i % 2])
    e1.record()
    sync()
    nccl_us = gmax(e0.elapsed_time(e1) / iters * 1e3)
    row = {"bytes": nbytes, "nccl_us": round(nccl_us, 2), "nccl_busbw": round(nbytes / nccl_us / 1e3 * 2 * (world - 1) / world, 1)}
    # expected sum for verification: every rank holds the same random pattern scaled by (rank+1)? no -- independent tensors,
    # so verify through NCCL itself on a fresh copy
    want = xs[0].clone()
    dist.all_reduce(want)
    for chunk in [int(c) for c in a.chunks.split(",")]:
        chunk_eff = max(1, min(chunk, (n + world - 1) // world))
        lay = L.compute_layout(world, n, chunk_eff, 1, code, ONESHOT_MAX_BYTES // es, True)
        k += 1
        symm = SymmetricMemory(UnixRendezvous(rank, world, f"{sess}_m{k}"), lay.regionBytes, lr, want_multicast=True)
        for algo in a.algos.split(","):
            try:   # "name:noov" = that kernel with one round in flight (no use of the maxLag window)
                eng = RoundEngine.from_symmetric(symm, lay, algo=algo.split(":")[0], timeout_ms=5000.0, grid=a.grid,
                                                 overlap=False if algo.endswith(":noov") else None)
            except ValueError:
                continue
            try:
                base = getattr(symm, "_next_round", 0)
                for r in range(base, base + 4):
                    out, counts = eng.enqueue_round(r, xs[r % 2])
                eng.wait_all(); sync()
                e0.record()
                for r in range(base + 4, base + 4 + iters):
                    out, counts = eng.enqueue_round(r, xs[r % 2])
                eng.wait_all()
                e1.record()
                sync()
                us = gmax(e0.elapsed_time(e1) / iters * 1e3)
                symm._next_round = base + 4 + iters
                # verify round with input xs[0]
                r = symm._next_round
                if r % 2 == 1:
                    eng.enqueue_round(r, xs[1]); r += 1
                out, counts = eng.enqueue_round(r, xs[0])
                eng.wait_all(); sync()
                symm._next_round = r + 1
                tol = 1e-5 if dtype == torch.float32 else 2e-2
                ok = bool(torch.allclose(out.float(), want.float(), rtol=tol, atol=tol * float(want.float().abs().max()))) and int(counts.min()) == world
                key = f"{algo}_c{chunk}" if len(a.chunks.split(",")) > 1 else algo
                row[key] = {"us": round(us, 2), "busbw": round(nbytes / us / 1e3 * 2 * (world - 1) / world, 1), "algo": eng.algo,
                            "vs_nccl": round(nccl_us / us, 3), "ok": ok, "status": eng.status()}
            except L.AarError as e:
                row[algo] = {"error": str(e)[:120]}
            eng.close()
        symm.close()
    rows.append(row)
    if rank == 0:
        print(json.dumps(row), flush=True)
if rank == 0:
    print("SWEEP " + json.dumps({"world": world, "dtype": a.dtype, "rows": rows}))
    if a.write_table:
        # the selection table: for every measured size the fastest kernel family that verified; a row covers sizes up to the
        # geometric mean with the next measured size (the last row covers everything above)
        best = []
        for row in rows:
            cands = [(v["us"], k.split(":")[0]) for k, v in row.items() if isinstance(v, dict) and v.get("ok") and "us" in v and ":" not in k and k != "auto"]
            if cands:
                best.append((row["bytes"], min(cands)[1]))
        table_rows = []
        for i, (nb, algo) in enumerate(best):
            hi = int((nb * best[i + 1][0]) ** 0.5) if i + 1 < len(best) else 1 << 40
            if table_rows and table_rows[-1][1] == algo:
                table_rows[-1][0] = hi
            else:
                table_rows.append([hi, algo, 0])
        path = Path(a.write_table)
        try:
            table = json.loads(path.read_text())
        except (OSError, ValueError):
            table = {}
        table.setdefault(str(world), {})[a.dtype] = table_rows
        path.write_text(json.dumps(table, indent=1))
        print("TABLE " + json.dumps({str(world): {a.dtype: table_rows}}))
rdv.barrier()
dist.destroy_process_group()
