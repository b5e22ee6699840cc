"""Small single-GPU target for Nsight Compute: the headline shape (16 Mi fp32, 1 Mi chunks).

    python bench/ncu_target.py [reg|bulk] [P]

Emulated ranks (default 8): round k launches P x K1, P x K2, P x K3 -- 3P launches of roundKernel (reg) or bulkRoundKernel (bulk,
each phase = one role of the bulk kernel with the whole grid), three rounds, so `-s 6P -c 3P` captures the third round.
Then the real single-rank round three times: one launch each of the role-specialised bulk kernel (push / reduce / gather CTAs
waiting on each other inside one grid), the kernel bench.py --gpus 1 times."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.parallel.engine import RoundEngine  # noqa: E402
from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster  # noqa: E402

impl = sys.argv[1] if len(sys.argv) > 1 else "bulk"
P = int(sys.argv[2]) if len(sys.argv) > 2 else 8
n, chunk = 16 * 1024 * 1024, 1024 * 1024
cl = LocalGpuCluster(P, n, chunk, max_lag=1, timeout_ms=20000, impl=impl)
xs = [torch.randn(n, device="cuda:0") for _ in range(P)]
for r in range(3):
    cl.run_round(r, xs)
torch.cuda.synchronize()
want = torch.zeros(n, device="cuda:0")
for x in xs:
    want = want + x
assert torch.equal(cl.outputs[0], want)
del cl
eng = RoundEngine.single_process(n, chunk, algo="bulk" if impl == "bulk" else "twoshot")
for r in range(3):
    y = eng.allreduce(xs[r % P], r)
torch.cuda.synchronize()
assert torch.equal(y, xs[2 % P])
print("ncu target ok")
