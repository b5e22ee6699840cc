"""One GPU, profiling build (-DAAR_BULK_PROFILE): where thread 0 of the bulk reduce role spends its cycles."""
import argparse, ctypes as C, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster
ap = argparse.ArgumentParser()
ap.add_argument("--P", type=int, default=2)
ap.add_argument("--grid", type=int, default=48)
a = ap.parse_args()
n, chunk = 16 << 20, 1 << 20
cl = LocalGpuCluster(a.P, n, chunk, max_lag=1, impl="bulk", timeout_ms=5000, grid=a.grid)
xs = [torch.randn(n, device="cuda:0") for _ in range(a.P)]
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for rnd in range(3):
    cl._prep(rnd, xs, False, ())
    for ph in (1, 2, 4):
        for r in range(a.P):
            if rnd == 2 and r == 0 and ph == 2:
                cl.params[0].flags |= L.FLAG_TRACE
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                cl._phase(r, ph, st)
                e1.record()
                torch.cuda.synchronize()
                tr = cl._section(0, cl.layout.traceOff, L.TRACE_CTAS * L.TRACE_MARKS * 8, torch.int64).view(L.TRACE_CTAS, L.TRACE_MARKS).cpu()
                cl.params[0].flags &= ~L.FLAG_TRACE
                print("phase us", e0.elapsed_time(e1) * 1e3)
            else:
                cl._phase(r, ph, st)
torch.cuda.synchronize()
names = ["fill", "wait", "comp", "sync", "store", "read", "done", "tiles"]
for b in (0, 1, a.grid // 2, a.grid - 1):
    row = tr[b].tolist()
    tiles = max(1, row[7])
    print("cta", b, {nm: (round(v / tiles) if nm != "tiles" else v) for nm, v in zip(names, row)}, "cycles per tile")
