"""Tiny single-GPU target for compute-sanitizer (racecheck / memcheck): 3 emulated ranks, unaligned sizes, 3 rounds, thresholds < 1."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster  # noqa: E402

for (P, n, chunk, th) in [(3, 1001, 100, 1.0), (4, 4096, 512, 0.75)]:
    cl = LocalGpuCluster(P, n, chunk, max_lag=1, th_reduce=th, th_complete=th, timeout_ms=60000, grid=4)
    for r in range(3):
        xs = [torch.randn(n, device="cuda:0") for _ in range(P)]
        outs, counts = cl.run_round(r, xs, skip=[1] if th < 1 else [])
    torch.cuda.synchronize()
print("sanitizer target ok")
