"""One GPU: what a copy role costs per CTA.  Register copy loop (512 threads, 8 x 16 bytes in flight per thread) against the
bulk async-copy engine (one issuing thread per CTA, shared-memory ring), both local -> local, over grid sizes.  The point
is GB/s PER CTA: the fused kernels give each role a fixed number of SMs, so a role that moves more per SM needs fewer."""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mib", type=int, default=64)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--grids", default="8,16,32,48,74,148,296")
a = ap.parse_args()
lib = L.load()
nbytes = a.mib << 20
dev = torch.device("cuda", 0)
# four src/dst pairs rotate so that consecutive launches do not hit L2 (4 x 2 x 64 MiB > 126 MB)
bufs = [torch.empty(2 * nbytes, dtype=torch.uint8, device=dev) for _ in range(4)]
for b in bufs:
    b[:nbytes].view(torch.int32).copy_(torch.arange(nbytes // 4, dtype=torch.int32, device=dev))
s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
err = torch.zeros(1, dtype=torch.int32, device=dev)
res = {"MiB": a.mib, "device": torch.cuda.get_device_name(0)}


def timed(fn):
    fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.iters):
        fn(i % 4)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.iters * 1e3


for grid in [int(g) for g in a.grids.split(",")]:
    if grid <= 148:
        us = timed(lambda k: L.check(lib.aar_bench_link(5, grid, 8, None, None, C.c_void_p(bufs[k].data_ptr()), nbytes, 1, s)))
        res[f"reg_copy_g{grid}"] = {"us": round(us, 1), "GBps": round(nbytes / us / 1e3, 1), "GBps_per_cta": round(nbytes / us / 1e3 / grid, 2)}
    for stages, tile in ((4, 16384), (4, 32768), (6, 32768), (8, 16384), (3, 65536)):
        if stages not in (2, 4, 8):
            continue

        def run(k):
            L.check(lib.aar_bench_bulk_copy(grid, stages, tile, C.c_void_p(bufs[k].data_ptr()), C.c_void_p(bufs[k].data_ptr() + nbytes), nbytes, 1, s,
                                            C.c_void_p(err.data_ptr())))

        us = timed(run)
        ok = bool(torch.equal(bufs[0][nbytes:], bufs[0][:nbytes]))
        res[f"bulk_s{stages}_t{tile}_g{grid}"] = {"us": round(us, 1), "GBps": round(nbytes / us / 1e3, 1), "GBps_per_cta": round(nbytes / us / 1e3 / grid, 2),
                                                 "ok": ok, "gave_up": int(err.item())}
# several issuing threads per CTA (each its own ring): does the per-copy issue cost overlap?
for grid in (8, 32):
    for warps, stages, tile in ((1, 4, 8192), (2, 4, 8192), (4, 4, 8192), (8, 2, 8192), (1, 4, 16384), (2, 4, 16384), (4, 2, 16384), (2, 2, 32768), (4, 2, 16384)):
        def run(k):
            L.check(lib.aar_bench_bulk_copy_multi(grid, warps, stages, tile, C.c_void_p(bufs[k].data_ptr()), C.c_void_p(bufs[k].data_ptr() + nbytes), nbytes, 1, s,
                                                  C.c_void_p(err.data_ptr())))
        us = timed(run)
        ok = bool(torch.equal(bufs[0][nbytes:], bufs[0][:nbytes]))
        res[f"multi_w{warps}_s{stages}_t{tile}_g{grid}"] = {"us": round(us, 1), "GBps_per_cta": round(nbytes / us / 1e3 / grid, 2), "ok": ok, "gave_up": int(err.item())}
print(json.dumps(res))
