"""Single-GPU kernel timing: the three phases of one rank of an emulated P-rank cluster (all regions on one device,
so "NVLink" traffic is local HBM traffic) and the real single-rank fused round.  CUDA-event timed, warm-up first,
buffers larger than L2.  Used to tune tile shapes before spending multi-GPU minutes.

    python bench/kernel_bench.py [--P 8] [--n 16777216] [--chunk 1048576] [--dtype fp32] [--iters 10]
"""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L  # noqa: E402
from akka_allreduce_b200.parallel.engine import RoundEngine  # noqa: E402
from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster  # noqa: E402


def timed(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return ts[len(ts) // 2] * 1e3  # median, microseconds


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=8)
    ap.add_argument("--n", type=int, default=16 * 1024 * 1024)
    ap.add_argument("--chunk", type=int, default=1024 * 1024)
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--grid", type=int, default=0)
    ap.add_argument("--impl", default="reg", choices=["reg", "bulk"])
    ap.add_argument("--self-direct", action="store_true")
    ap.add_argument("--split", default="", help="nA:nB for the role-specialised single-rank round")
    a = ap.parse_args()
    dtype = torch.float32 if a.dtype == "fp32" else torch.bfloat16
    es = 4 if dtype == torch.float32 else 2
    res = {"P": a.P, "n": a.n, "chunk": a.chunk, "dtype": a.dtype, "bytes": a.n * es, "grid": a.grid, "impl": a.impl, "self_direct": a.self_direct}

    cl = LocalGpuCluster(a.P, a.n, a.chunk, max_lag=1, dtype=dtype, grid=a.grid, timeout_ms=5000, impl=a.impl, self_direct=a.self_direct)
    xs = [torch.randn(a.n, device="cuda:0").to(dtype) for _ in range(a.P)]
    lay = C.byref(cl.layout)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rnd = [0]

    # full emulated rounds (P x (K1 + K2 + K3)) so flags stay consistent; per-phase time measured inside
    def full():
        cl.run_round(rnd[0], xs)
        rnd[0] += 1

    res["emulated_round_us_all_ranks"] = timed(full, a.iters)

    # phase timing for rank 0 only: replay phases of a fresh round (other ranks' phases run untimed around it)
    def phase_times():
        r = rnd[0]
        rnd[0] += 1
        cl._prep(r, xs, False, ())
        out = {}
        e = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        for k in range(1, a.P):
            cl._phase(k, 1, stream)
        e[0].record()
        cl._phase(0, 1, stream)
        e[1].record()
        for k in range(1, a.P):
            cl._phase(k, 2, stream)
        e[2].record()
        cl._phase(0, 2, stream)
        e[3].record()
        for k in range(1, a.P):
            cl._phase(k, 4, stream)
        e[4].record()
        cl._phase(0, 4, stream)
        e[5].record()
        torch.cuda.synchronize()
        out["K1_scatter_us"] = e[0].elapsed_time(e[1]) * 1e3
        out["K2_reduce_bcast_us"] = e[2].elapsed_time(e[3]) * 1e3
        out["K3_gather_us"] = e[4].elapsed_time(e[5]) * 1e3
        return out

    runs = [phase_times() for _ in range(a.iters)]
    for k in runs[0]:
        v = sorted(r[k] for r in runs)
        res[k] = v[len(v) // 2]
    nb = a.n * es
    # bytes each phase moves through HBM on one device in the emulation
    res["K1_GBps"] = 2 * nb / res["K1_scatter_us"] / 1e3
    res["K2_GBps"] = 2 * nb / res["K2_reduce_bcast_us"] / 1e3
    res["K3_GBps"] = 2 * nb / res["K3_gather_us"] / 1e3
    del cl

    eng = RoundEngine.single_process(a.n, a.chunk, dtype=dtype, grid=a.grid, algo="twoshot" if a.impl == "reg" else "bulk")
    if a.split:
        na, nb_ = (int(v) for v in a.split.split(":"))
        eng.bulk_split = na | (nb_ << 16)
    x = xs[0]
    k = [0]

    def one():
        eng.enqueue_round(k[0], x)
        k[0] += 1

    res["single_rank_fused_round_us"] = timed(one, a.iters)
    res["single_rank_fused_GBps_algo"] = nb / res["single_rank_fused_round_us"] / 1e3
    print(json.dumps(res))


if __name__ == "__main__":
    main()
