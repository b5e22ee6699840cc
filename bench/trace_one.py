"""One GPU: per-CTA phase timestamps (AAR_FLAG_TRACE) of rank 0's phases in an emulated round, bulk or register kernels."""
import argparse, ctypes as C, json, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.ops import lib as L
from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster

ap = argparse.ArgumentParser()
ap.add_argument("--P", type=int, default=2)
ap.add_argument("--n", type=int, default=16 << 20)
ap.add_argument("--chunk", type=int, default=1 << 20)
ap.add_argument("--impl", default="bulk")
ap.add_argument("--phase", type=int, default=2)
a = ap.parse_args()
cl = LocalGpuCluster(a.P, a.n, a.chunk, max_lag=1, impl=a.impl, timeout_ms=5000)
xs = [torch.randn(a.n, device="cuda:0") for _ in range(a.P)]
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for rnd in range(3):
    cl._prep(rnd, xs, False, ())
    if rnd == 2:
        cl.params[0].flags |= L.FLAG_TRACE
    for ph in (1, 2, 4):
        for r in range(a.P):
            if rnd == 2 and r == 0:
                cl._section(0, cl.layout.traceOff, L.TRACE_CTAS * L.TRACE_MARKS * 8, torch.int64).zero_()
                torch.cuda.synchronize()
            cl._phase(r, ph, st)
            if rnd == 2 and r == 0 and ph == a.phase:
                torch.cuda.synchronize()
                tr = cl._section(0, cl.layout.traceOff, L.TRACE_CTAS * L.TRACE_MARKS * 8, torch.int64).view(L.TRACE_CTAS, L.TRACE_MARKS).cpu()
torch.cuda.synchronize()
t0 = int(tr[:148, 0][tr[:148, 0] > 0].min())
rows = []
for b in range(148):
    if int(tr[b, 0]) == 0:
        continue
    rows.append([b] + [round((int(v) - t0) / 1e3, 1) if int(v) else None for v in tr[b]])
print("cta, marks 0..7 in us since the first CTA started (0 start, 1 after push, 2 first decision, 6 first tile loaded, 7 loop end, 3 reduce end, 4 gather, 5 end)")
for r in rows[:6] + rows[70:74] + rows[-4:]:
    print(r)
ends = [r[6] for r in rows if r[6] is not None]
print("ctas", len(rows), "max end", max(ends))
