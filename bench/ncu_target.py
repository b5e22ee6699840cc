"""Small single-GPU target for Nsight Compute: 8 emulated ranks, the headline shape (16 Mi fp32, 1 Mi chunks), three rounds.
Round k launches 8 x K1, 8 x K2, 8 x K3 (24 launches of roundKernel<float,false>), so `-s 48 -c 24` captures the third round."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster  # noqa: E402

P, n, chunk = 8, 16 * 1024 * 1024, 1024 * 1024
cl = LocalGpuCluster(P, n, chunk, max_lag=1, timeout_ms=20000)
xs = [torch.randn(n, device="cuda:0") for _ in range(P)]
for r in range(3):
    cl.run_round(r, xs)
torch.cuda.synchronize()
want = torch.zeros(n, device="cuda:0")
for x in xs:
    want = want + x
assert torch.equal(cl.outputs[0], want)
print("ncu target ok")
