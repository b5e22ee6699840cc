"""Straggler mode (BASELINE.json config 3): N GPUs, thReduce = thComplete = 0.75, maxLag = 4, one rank delayed per round
(the delayed rank rotates: round r delays rank r % N by `--delay-us` before it launches that round).

Reports rounds/sec without and with the injected straggler for (a) this framework at thresholds 0.75, (b) this framework
at thresholds 1.0 (every round waits for the straggler), (c) NCCL allreduce with the same injected delay.
Run under torch.distributed.run."""
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

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=16 * 1024 * 1024)
ap.add_argument("--chunk", type=int, default=1024 * 1024)
ap.add_argument("--rounds", type=int, default=64)
ap.add_argument("--warmup", type=int, default=8)
ap.add_argument("--delays-us", default="100,250,500")
ap.add_argument("--max-lag", type=int, default=4)
ap.add_argument("--th", type=float, default=0.75)
ap.add_argument("--inflight", type=int, default=2, help="rounds of one rank in flight (internal streams), at most maxLag + 1")
ap.add_argument("--algo", default="auto")
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
sess = f"sb_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
lib = L.load()


def gmax(x):
    return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))


def sync():
    torch.cuda.synchronize()
    rdv.barrier()


xs = [torch.randn(a.n, device=dev) for _ in range(2)]
res = {"world": world, "n": a.n, "chunk": a.chunk, "max_lag": a.max_lag, "threshold": a.th, "rounds": a.rounds}
lay = L.compute_layout(world, a.n, a.chunk, a.max_lag, L.AAR_F32, 0, True)
symm = SymmetricMemory(UnixRendezvous(rank, world, sess + "_m"), lay.regionBytes, lr, want_multicast=True)


def run_engine(th, delay_us, tag):
    # more than two rounds in flight only with a partial quorum: at thresholds 1.0 every round waits for the delayed rank, and
    # later rounds' kernels holding SMs while they wait can starve the round everybody is waiting for
    eng = RoundEngine.from_symmetric(symm, lay, th_reduce=th, th_complete=th, timeout_ms=2000.0, max_inflight=a.inflight if th < 1.0 else 2, algo=a.algo)
    eng.clear_status()   # the region is reused from run to run and its status bits are sticky
    import time
    base = run_engine.next_round
    tw = time.time()
    for r in range(base, base + a.warmup):
        eng.enqueue_round(r, xs[r % 2])
    eng.wait_all(); sync()
    warm_s = time.time() - tw
    st_warm = eng.status()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    cmin, cmax = 1 << 30, 0
    first = base + a.warmup
    for r in range(first, first + a.rounds):
        d = int(delay_us * 1000) if (delay_us and r % world == rank) else 0
        out, counts = eng.enqueue_round(r, xs[r % 2], delay_ns=d)
    eng.wait_all()
    e1.record()
    sync()
    ms = gmax(e0.elapsed_time(e1) / a.rounds)
    cmin, cmax = int(counts.min()), int(counts.max())
    # what the partial quorum cost in the last maxLag + 1 rounds: chunks flushed as zeros (count 0) and the mean contributor count
    recent = torch.stack(eng.counts).float()
    zero_frac = float((recent == 0).float().mean())
    mean_cnt = float(recent.mean())
    res[tag] = {"zero_chunk_fraction_last_rounds": round(zero_frac, 4), "mean_contributors_last_rounds": round(mean_cnt, 3), "rounds_in_flight": eng.inflight,"ms_per_round": round(ms, 4), "rounds_per_sec": round(1e3 / ms, 1), "last_round_counts_min_max": [cmin, cmax], "status_bits": eng.status(), "status_after_warmup": st_warm, "warmup_wall_s": round(warm_s, 3), "algo": eng.algo}
    run_engine.next_round = first + a.rounds
    eng.close()
    # the region keeps its flags: continue with later round numbers so epochs stay monotone


run_engine.next_round = 0
delays = [int(d) for d in a.delays_us.split(",")]
run_engine(a.th, 0, f"aar_th{a.th}_no_straggler")
for d in delays:
    run_engine(a.th, d, f"aar_th{a.th}_delay{d}us")
run_engine(1.0, 0, "aar_th1.0_no_straggler")
for d in delays:
    run_engine(1.0, d, f"aar_th1.0_delay{d}us")

import torch.distributed as dist  # noqa: E402

dist.init_process_group("nccl", device_id=dev)
bufs = [x.clone() for x in xs]
s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for d in [0] + delays:
    for i in range(a.warmup):
        dist.all_reduce(bufs[i % 2])
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(a.rounds):
        if d and r % world == rank:
            L.check(lib.aar_launch_delay(d * 1000, s))
        dist.all_reduce(bufs[r % 2])
    e1.record()
    sync()
    ms = gmax(e0.elapsed_time(e1) / a.rounds)
    res[f"nccl_delay{d}us"] = {"ms_per_round": round(ms, 4), "rounds_per_sec": round(1e3 / ms, 1)}
base_rate = res[f"aar_th{a.th}_no_straggler"]["rounds_per_sec"]
for d in delays:
    res[f"retained_th{a.th}_delay{d}us"] = round(res[f"aar_th{a.th}_delay{d}us"]["rounds_per_sec"] / base_rate, 3)
    res[f"retained_th1.0_delay{d}us"] = round(res[f"aar_th1.0_delay{d}us"]["rounds_per_sec"] / res["aar_th1.0_no_straggler"]["rounds_per_sec"], 3)
    res[f"retained_nccl_delay{d}us"] = round(res[f"nccl_delay{d}us"]["rounds_per_sec"] / res["nccl_delay0us"]["rounds_per_sec"], 3)
if rank == 0:
    print(json.dumps(res))
rdv.barrier()
dist.destroy_process_group()
