"""Body of the multi-GPU tests (launched by tests/test_gpu_multi.py under torch.distributed.run, one process per GPU).

Every check is bit-exact against a PyTorch fp32 restatement of the reference's ordered reduction over the contributor set
(AllreduceWorker.scala:240-251): the inputs of all ranks are generated from a shared seed, so every rank can compute the
oracle locally."""
import os
import struct
import sys
import threading
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput  # noqa: E402
from akka_allreduce_b200.ops import lib as L  # noqa: E402
from akka_allreduce_b200.parallel.engine import ONESHOT_MAX_BYTES, RoundEngine  # noqa: E402
from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
sess = f"mt_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
failures = []


def inputs_for(n, dtype, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return [(torch.randn(n, generator=g) * 3).to(dtype).to(dev) for _ in range(world)]


def ordered_sum(xs, srcs, dtype, normalize=False):
    acc = torch.zeros_like(xs[0], dtype=torch.float32)
    for s in srcs:
        acc = acc + xs[s].float()
    if normalize and len(srcs) > 1:
        acc = torch.div(acc, torch.full_like(acc, float(len(srcs))))
    return acc.to(dtype)


def check(name, cond):
    ok = all(b == b"\1" for b in rdv.all_gather_bytes(b"\1" if cond else b"\0"))
    if rank == 0:
        print(("PASS " if ok else "FAIL ") + name, flush=True)
    if not ok:
        failures.append(name)


k = [0]


def engine_case(n, chunk, dtype, algo, rounds=5, th=1.0, normalize=False, overlap=None, max_lag=1):
    code = L.AAR_F32 if dtype == torch.float32 else L.AAR_BF16
    es = 4 if dtype == torch.float32 else 2
    lay = L.compute_layout(world, n, chunk, max_lag, code, ONESHOT_MAX_BYTES // es, True)
    k[0] += 1
    symm = SymmetricMemory(UnixRendezvous(rank, world, f"{sess}_m{k[0]}"), lay.regionBytes, lr, want_multicast=True)
    name = f"{algo} n={n} chunk={chunk} {str(dtype)[6:]} th={th} overlap={overlap} mc={symm.has_multicast}"
    try:
        eng = RoundEngine.from_symmetric(symm, lay, algo=algo, th_reduce=th, th_complete=th, normalize=normalize, overlap=overlap, timeout_ms=5000.0)
    except ValueError as e:
        if rank == 0:
            print(f"SKIP {name}: {e}", flush=True)
        symm.close()
        return
    good = True
    keep = []
    first_pass = []
    for r in range(rounds):
        xs = inputs_for(n, dtype, 1000 + r)
        keep.append(xs)
        out, counts = eng.enqueue_round(r, xs[rank])
        eng.wait_round(r)
        torch.cuda.synchronize()
        want = ordered_sum(xs, list(range(world)), dtype, normalize)
        if eng.algo in ("nvls", "nvls_bulk", "nvls_onehop"):  # the switch does not promise peer-index order: equal up to fp32 reassociation
            good &= bool(torch.allclose(out.float(), want.float(), rtol=2e-5 if dtype == torch.float32 else 2e-2, atol=1e-4))
            # ... but every rank must hold the SAME bits (one multimem.st lands everywhere), and the same inputs must give the
            # same bits again (round r + rounds below repeats round r's inputs)
            digest = struct.pack("q", int(out.view(torch.int16 if dtype == torch.bfloat16 else torch.int32).to(torch.int64).sum().item()))
            good &= len(set(rdv.all_gather_bytes(digest))) == 1
            first_pass.append(out.clone())
        else:
            good &= bool(torch.equal(out, want))
        good &= int(counts.min()) == world and int(counts.max()) == world
        rdv.barrier()
    if first_pass:   # run-to-run determinism of the in-switch sum
        for r in range(rounds):
            out, counts = eng.enqueue_round(rounds + r, keep[r][rank])
            eng.wait_round(rounds + r)
            torch.cuda.synchronize()
            good &= bool(torch.equal(out, first_pass[r]))
            rdv.barrier()
    good &= eng.status() == 0
    check(name + f" -> {eng.algo}", good)
    eng.close()
    symm.close()


def threshold_case():
    """One rank never scatters (AAR_FLAG_SKIP_SCATTER): with thReduce = thComplete = 0.5 every chunk is reduced over whichever
    quorum its owner saw first, the counts say how many, and nobody hangs.  The contributor set of a chunk is read back
    from its owner's decision word (all regions are peer-mapped), so the flushed values can be checked bit for bit against
    the ordered sum over exactly that set -- "bit-exact against a CPU reduction of the contributing set"."""
    n, chunk = 1 << 16, 1 << 12
    lay = L.compute_layout(world, n, chunk, 1, L.AAR_F32, 0, False)
    k[0] += 1
    symm = SymmetricMemory(UnixRendezvous(rank, world, f"{sess}_m{k[0]}"), lay.regionBytes, lr, want_multicast=False)
    eng = RoundEngine.from_symmetric(symm, lay, algo="twoshot", th_reduce=0.5, th_complete=0.5, timeout_ms=5000.0, overlap=False)
    silent = world - 1
    xs = inputs_for(n, torch.float32, 77)
    out, counts = eng.enqueue_round(0, xs[rank], skip_scatter=(rank == silent))
    torch.cuda.synchronize()
    rdv.barrier()   # every owner has published its decisions
    ok = True
    geo_step, C = lay.stepSize, lay.C
    cmin, cmax = int(counts.min()), int(counts.max())
    ok &= cmax <= world - 1 or world == 1
    for owner in range(world):
        for c in range(C):
            s0 = owner * geo_step + c * chunk
            e0 = min(s0 + chunk, (owner + 1) * geo_step, n)
            if e0 <= s0:
                continue
            cnt = int(counts[owner * C + c])
            got = out[s0:e0]
            if cnt == 0:   # the chunk missed this rank's round decision (or was reduced over nobody): zero-filled
                ok &= bool((got == 0).all())
                continue
            word = int(symm.tensor(owner, lay.k2DecOff + (0 * C + c) * 8, 8, torch.int64)[0].item())
            mask = word & 0xFFFFFFFF
            srcs = [r for r in range(world) if (mask >> r) & 1]
            ok &= (word >> 32) == 1 and len(srcs) == cnt and silent not in srcs
            ok &= bool(torch.equal(got, ordered_sum([x[s0:e0] for x in xs], srcs, torch.float32)))
    check(f"threshold 0.5 with a silent rank: sums over the reported contributor sets (counts {cmin}..{cmax})", ok and eng.status() == 0)
    eng.close()
    symm.close()


def out_of_order_case():
    """AllreduceSpec.scala:675-677 on real peers: the last rank starts round 0 thirty milliseconds late (straggler injector), round
    1 is launched right behind it on the other stream.  At thresholds 1.0 round 0 waits for the late push, round 1 does not:
    every rank sees round 1 flushed while round 0 is still in flight, then round 0; both are bit-exact."""
    import time

    n, chunk = 1 << 18, 1 << 14
    lay = L.compute_layout(world, n, chunk, 1, L.AAR_F32, 0, False)
    k[0] += 1
    symm = SymmetricMemory(UnixRendezvous(rank, world, f"{sess}_m{k[0]}"), lay.regionBytes, lr, want_multicast=False)
    eng = RoundEngine.from_symmetric(symm, lay, algo="bulk", timeout_ms=5000.0)
    xs0, xs1 = inputs_for(n, torch.float32, 8100), inputs_for(n, torch.float32, 8101)
    rdv.barrier()
    out0, _ = eng.enqueue_round(0, xs0[rank], delay_ns=30_000_000 if rank == world - 1 else 0)
    out1, _ = eng.enqueue_round(1, xs1[rank])
    t0 = time.time()
    while not eng.round_flushed(1) and time.time() - t0 < 10:
        pass
    first = eng.round_flushed(1) and not eng.round_flushed(0)
    eng.synchronize()
    torch.cuda.synchronize()
    ok = first and eng.round_flushed(0) and eng.rounds_flushed() == 2
    ok &= bool(torch.equal(out0, ordered_sum(xs0, list(range(world)), torch.float32))) and bool(torch.equal(out1, ordered_sum(xs1, list(range(world)), torch.float32)))
    check("round 1 is flushed before round 0 when round 0's last contribution is late", ok and eng.status() == 0)
    eng.close()
    symm.close()


def self_start_case():
    """AllreduceSpec.scala:106-124 on real peers: the last worker never sees a StartAllreduce (its handler drops them); it starts
    every round by itself because its peers post the rounds they start on the shared board, and the job completes."""
    from akka_allreduce_b200 import _core as core

    n, chunk, rounds = 1 << 16, 1 << 12, 10
    port = int(os.environ.get("MASTER_PORT", "29500")) + 58
    master = None
    if rank == 0:
        master = AllreduceMaster(world, 1.0, 1.0, 1.0, 1, n, rounds - 1, chunk, pipeline_depth=1, port=port)
        master.start(120)
    rdv.barrier()
    seen = []

    def deafen(w):   # the master's StartAllreduce never reaches the last rank (fault injection, both round loops honour it)
        w.drop_start = w.id == world - 1

    def source(req):
        return AllReduceInput(inputs_for(n, torch.float32, 6000 + req.iteration)[worker.id])

    def sink(out):
        want = ordered_sum(inputs_for(n, torch.float32, 6000 + out.iteration), list(range(world)), torch.float32)
        in_switch = worker.engine.algo in ("nvls", "nvls_bulk", "nvls_onehop")
        seen.append((out.iteration, bool(torch.allclose(out.data, want, rtol=2e-5, atol=1e-4)) if in_switch else bool(torch.equal(out.data, want))))

    worker = AllreduceWorker(source, sink, backend="cuda", device=lr, session=f"{sess}_selfstart", on_init=deafen)
    ok = worker.run("127.0.0.1", port, timeout_sec=120)
    if master:
        master.join(30)
    m = worker.metrics()
    mine = m["self_started_rounds"] == rounds if worker.id == world - 1 else m["self_started_rounds"] == 0
    check("a worker that never hears StartAllreduce starts every round because its peers have", ok and mine and sorted(seen) == [(r, True) for r in range(rounds)])
    worker.close()


def api_case(native_loop=False):
    """AllreduceMaster + AllreduceWorker(backend=cuda) end to end, 12 rounds, pipelined master clock.  native_loop: the round loop
    in C++ (csrc/core/device_loop.cpp) launching through aar_engine_enqueue_ex directly; Python is the source and the sink only."""
    n, chunk, rounds = 1 << 18, 1 << 14, 12
    port = int(os.environ.get("MASTER_PORT", "29500")) + (59 if native_loop else 57)
    master = None
    if rank == 0:
        master = AllreduceMaster(world, 1.0, 1.0, 1.0, 1, n, rounds - 1, chunk, pipeline_depth=1, port=port)
        master.start(120)
    rdv.barrier()
    seen = []
    good = [True]

    def source(req):
        return AllReduceInput(inputs_for(n, torch.float32, 5000 + req.iteration)[worker.id])

    def sink(out):
        xs = inputs_for(n, torch.float32, 5000 + out.iteration)
        want = ordered_sum(xs, list(range(world)), torch.float32)
        in_switch = worker.engine.algo in ("nvls", "nvls_bulk", "nvls_onehop")   # the switch does not promise peer-index order
        ok = bool(torch.allclose(out.data, want, rtol=2e-5, atol=1e-4)) if in_switch else bool(torch.equal(out.data, want))
        good[0] &= ok and int(out.count.min()) == world
        seen.append(out.iteration)

    worker = AllreduceWorker(source, sink, backend="cuda", device=lr, session=f"{sess}_api{int(native_loop)}", native_loop=native_loop)
    ok = worker.run("127.0.0.1", port, timeout_sec=120)
    if master:
        master.join(30)
    m = worker.metrics()
    what = "public API, round loop in C++" if native_loop else "public API"
    check(f"{what}: master + cuda workers, every round once, sums exact",
          ok and good[0] and sorted(seen) == list(range(rounds)) and m["rounds_delivered"] == rounds and (not native_loop or worker._direct))
    worker.close()


only = os.environ.get("AAR_MULTI_ONLY", "")
if only == "api":   # quick check of the public API path alone
    api_case()
    api_case(native_loop=True)
    rdv.barrier()
    if rank == 0:
        print("MULTI-GPU RESULT:", "FAILED " + "; ".join(failures) if failures else "ALL PASSED", flush=True)
    sys.exit(1 if failures else 0)
for dtype in (torch.float32, torch.bfloat16):
    engine_case(1 << 20, 1 << 16, dtype, "twoshot", overlap=False)
    engine_case(1 << 20, 1 << 16, dtype, "pipelined")
    engine_case(1 << 20, 1 << 16, dtype, "nvls")
    engine_case(8192, 1024, dtype, "oneshot")
    engine_case(1 << 20, 1 << 16, dtype, "bulk")
    engine_case(1 << 22, 1 << 18, dtype, "bulk", rounds=8, overlap=False)
    engine_case(1 << 22, 1 << 18, dtype, "nvls_bulk", rounds=6)
    engine_case(1 << 18, 1 << 18, dtype, "nvls_bulk")
    engine_case(1 << 18, 1 << 18, dtype, "nvls_onehop", rounds=6)
    engine_case(1 << 20, 1 << 16, dtype, "nvls_onehop", rounds=4, normalize=True)
engine_case((1 << 20) + 8, 40000, torch.float32, "pipelined")      # unaligned blocks and chunks: scalar path
engine_case((1 << 20) + 8, 40000, torch.float32, "bulk")           # ... which the bulk engine hands to the register kernels
engine_case(1 << 24, 1 << 20, torch.float32, "bulk", rounds=6, max_lag=1)   # the headline geometry
engine_case(1 << 24, 1 << 20, torch.float32, "nvls_bulk", rounds=4, max_lag=1)
engine_case(1 << 22, 1 << 16, torch.float32, "bulk", rounds=12, max_lag=3, normalize=True)
engine_case(1 << 22, 1 << 18, torch.bfloat16, "nvls_bulk", rounds=4, normalize=True)   # gradient-bucket shape: mean over the ranks
engine_case(1003, 17, torch.float32, "twoshot", overlap=False)
engine_case(1 << 20, 1 << 16, torch.float32, "auto", normalize=True)
engine_case(1 << 22, 1 << 18, torch.float32, "auto", rounds=12, max_lag=3)
threshold_case()
out_of_order_case()
api_case()
api_case(native_loop=True)
self_start_case()
rdv.barrier()
if rank == 0:
    print("MULTI-GPU RESULT:", "FAILED " + "; ".join(failures) if failures else "ALL PASSED", flush=True)
sys.exit(1 if failures else 0)
