"""Debug helper: run an emulated cluster for several rounds and report where outputs differ from the fp32 oracle."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tests"))
from test_gpu_local import make_inputs, oracle  # noqa: E402
from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster  # noqa: E402

P, n, chunk = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (8, (1 << 20) + 8, 40000)))
for trial in range(6):
    cl = LocalGpuCluster(P, n, chunk, max_lag=1)
    lay = cl.layout
    print(f"trial {trial}: step={lay.stepSize} blkStride={lay.blkStride} C={lay.C} Tm={lay.tilesPerChunk} tile={lay.tileElems}")
    for rnd in range(4):
        xs = make_inputs(P, n, torch.float32, 100 + rnd)
        outs, counts = cl.run_round(rnd, xs)
        want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.float32)
        torch.cuda.synchronize()
        for r in range(P):
            bad = (outs[r] != want).nonzero().flatten()
            if bad.numel():
                idx = bad.tolist()
                blocks = sorted({i // lay.stepSize for i in idx})
                first, last = idx[0], idx[-1]
                inblk = first % lay.stepSize
                print(f"  round {rnd} rank {r}: {bad.numel()} bad, first {first} (block {first // lay.stepSize}, chunk {inblk // chunk}, off {inblk % chunk}) last {last}, blocks {blocks}, "
                      f"got {outs[r][first].item()} want {want[first].item()} status {cl.status(r)}")
                d = bad[1:] - bad[:-1]
                starts = [idx[0]] + [idx[k + 1] for k in (d != 1).nonzero().flatten().tolist()]
                print(f"    runs start at {starts[:12]} ({len(starts)} runs)")
                owner = first // lay.stepSize
                slot = rnd % lay.L
                es = 4
                for src in range(P):
                    ring = cl._section(owner, lay.scatterOff + ((slot * P + src) * lay.blkStride) * es, lay.stepSize * es, torch.float32)
                    blk = xs[src][owner * lay.stepSize:(owner + 1) * lay.stepSize]
                    nb = int((ring[: blk.numel()] != blk).sum())
                    if nb:
                        print(f"    scatter ring of owner {owner} src {src}: {nb} elements differ from the input")
                red = cl._section(r, lay.reduceOff + ((slot * P + owner) * lay.blkStride) * es, lay.stepSize * es, torch.float32)
                wb = want[owner * lay.stepSize:(owner + 1) * lay.stepSize]
                print(f"    reduce ring of rank {r} owner {owner}: {int((red[: wb.numel()] != wb).sum())} elements differ from the oracle")
                # is the wrong value a partial sum? identify which sources are included
                s = torch.zeros((), device="cuda")
                for src in range(P):
                    s = s + xs[src][first]
                    if s.item() == outs[r][first].item():
                        print(f"    value equals the sum of sources 0..{src}")
                break
        else:
            print(f"  round {rnd}: ok")
