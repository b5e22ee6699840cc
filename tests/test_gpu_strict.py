"""GPU tests of the strict protocol variants on one device: the reduce-broadcast obeys the scatter's per-destination slot flow
control at thresholds < 1 (AAR_FLAG_BCAST_FC on the register kernels, always on in the bulk kernels), late-work elision, the
resumed round numbering and the native round engine (docs/PROTOCOL.md; model: csrc/tests/protocol_litmus.cpp).  First run on
B200 at the start of round 2: all green, so they are part of the regular suite.
"""
import pytest
import torch

from test_gpu_local import cluster, kernel_impl, make_inputs, oracle  # noqa: F401  (kernel_impl: autouse fixture, both kernel generations)

pytestmark = [pytest.mark.gpu]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bcast_flow_control_matches_the_default_path_when_nobody_lags(dtype):
    """Same rounds, same inputs, with and without the flag: identical outputs, counts and masks over two ring revolutions."""
    P, n, chunk = 4, 8192, 1024
    a = cluster(P, n, chunk, max_lag=1, th_reduce=0.75, th_complete=0.75, dtype=dtype)
    b = cluster(P, n, chunk, max_lag=1, th_reduce=0.75, th_complete=0.75, dtype=dtype, bcast_fc=True)
    for rnd in range(5):
        xs = make_inputs(P, n, dtype, 40 + rnd)
        skip = [rnd % P] if rnd % 2 else []
        oa, ca = a.run_round(rnd, xs, skip=skip)
        ob, cb = b.run_round(rnd, xs, skip=skip)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(oa[r], ob[r]) and torch.equal(ca[r], cb[r])
            assert b.status(r) == 0


def test_bcast_flow_control_full_rounds_bit_exact():
    P, n, chunk = 8, (1 << 18) + 8, 40000
    cl = cluster(P, n, chunk, max_lag=1, th_reduce=0.9, th_complete=0.9, bcast_fc=True)
    for rnd in range(4):
        xs = make_inputs(P, n, torch.float32, 300 + rnd)
        outs, counts = cl.run_round(rnd, xs)
        want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.float32)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want) and torch.equal(counts[r], wc)


def test_bcast_skips_a_destination_that_fell_out_of_the_window():
    """Rank 3 stops taking part (never retires anything).  After L rounds its slots are still occupied: the scatter drops the
    pushes to it (AAR_ST_PUSH_DROPPED) and, with the flag, the broadcast leaves its reduce ring alone as well: the rflag words
    of rank 3 for the dropped round keep their old epoch."""
    from akka_allreduce_b200.ops import lib as L

    P, n, chunk = 4, 4096, 512
    cl = cluster(P, n, chunk, max_lag=1, th_reduce=0.75, th_complete=0.75, timeout_ms=50.0, bcast_fc=True)
    live = [0, 1, 2]
    lay = cl.layout
    for rnd in range(4):
        xs = make_inputs(P, n, torch.float32, 500 + rnd)
        cl.run_phases(rnd, xs, scatter=live, reduce=live, gather=live)
        torch.cuda.synchronize()
    if cl.impl == "bulk":   # the bulk kernels publish per piece
        flags3 = cl._section(3, lay.pflagROff, lay.L * lay.P * lay.C * lay.piecesPerChunk * 8, torch.int64)
    else:
        flags3 = cl._section(3, lay.rflagOff, lay.L * lay.P * lay.C * 8, torch.int64)
    epochs = (flags3 >> 32).tolist()
    # rounds 0 and 1 fit the ring (nothing to wait for); rounds 2 and 3 would overwrite slots rank 3 never retired
    assert max(epochs) <= 2, epochs
    assert all(cl.status(r) & L.ST_PUSH_DROPPED for r in live)
    # the live ranks still complete every round among themselves
    want, _ = oracle(xs, P, n, chunk, lambda o, c: live, torch.float32, missing_owner_chunks=[(3, c) for c in range(2)])
    for r in live:
        assert torch.equal(cl.outputs[r], want)


def test_engine_start_at_continues_the_round_numbering():
    """RoundEngine.start_at (resume from a master checkpoint): a fresh engine begins at round 7 -- flow-control words say every
    earlier round is retired -- and runs rounds 7..12 bit-exactly (slots reused three times)."""
    from akka_allreduce_b200.parallel.engine import RoundEngine

    n = 1 << 16
    eng = RoundEngine.single_process(n, 1 << 12, max_lag=1, timeout_ms=2000.0)
    eng.start_at(7)
    for rnd in range(7, 13):
        x = make_inputs(1, n, torch.float32, 700 + rnd)[0]
        out, counts = eng.enqueue_round(rnd, x)
        eng.wait_round(rnd)
        torch.cuda.synchronize()
        assert torch.equal(out, x) and int(counts.min()) == 1
    assert eng.rounds_flushed() == 13 and eng.status() == 0
    eng.close()


def test_eager_decision_words_exist_in_the_layout():
    """AAR_FLAG_EAGER_DEC on the pipelined kernel is not reachable on one GPU (the emulation launches phases separately with the
    leader gather), so the single-GPU check is the layout: the decision words exist, sit behind every older section and start
    zeroed; the kernel path itself is covered by tests/multi_gpu_worker.py run with AAR_STRICT=1."""
    from akka_allreduce_b200.ops import lib as L

    lay = L.compute_layout(8, 1 << 20, 1 << 16, 1, L.AAR_F32, 0, True)
    assert lay.egDecOff > lay.traceOff and lay.egDecOff + lay.L * lay.P * lay.C * 8 <= lay.regionBytes


def test_late_rank_elides_work_nobody_will_read():
    """AAR_FLAG_ELIDE_LATE: ranks 0-2 run round 0 to the end without rank 3 (thresholds 0.75).  Rank 3 then runs the same round:
    its pushes to owners that decided without it are skipped (their arrival flags for source 3 stay unset, its election counters
    are back at zero), its broadcast goes to nobody but itself, and its own flushed vector is still exact: blocks of owners 0-2
    are the sums those owners decided (sources 0-2), its own block sums all four."""
    P, n, chunk = 4, 4096, 512
    cl = cluster(P, n, chunk, max_lag=1, th_reduce=0.75, th_complete=0.75, elide_late=True, impl="reg")   # elision lives in the register kernels
    xs = make_inputs(P, n, torch.float32, 900)
    live = [0, 1, 2]
    cl.run_phases(0, xs, scatter=live, reduce=live, gather=live)
    torch.cuda.synchronize()
    lay = cl.layout
    rflag_before = [cl._section(r, lay.rflagOff, lay.L * lay.P * lay.C * 8, torch.int64).clone() for r in live]
    cl.run_phases(0, xs, scatter=[3], reduce=[3], gather=[3])
    torch.cuda.synchronize()
    for r in live:
        sflags = cl._section(r, lay.sflagOff, lay.L * lay.C * 16 * 4, torch.int32).view(lay.L, lay.C, 16)
        assert int(sflags[0, :, 3].abs().sum()) == 0, "rank 3 pushed although the owner had decided without it"
        assert torch.equal(cl._section(r, lay.rflagOff, lay.L * lay.P * lay.C * 8, torch.int64), rflag_before[r]), "late broadcast reached a retired peer"
    ctr = cl._section(3, lay.k1CtrOff, lay.L * lay.P * lay.C * 4, torch.int32)
    assert int(ctr.abs().sum()) == 0, "election counters must be back at zero for the slot's next tenant"
    geo_block = n // P
    want_others, _ = oracle(xs, P, n, chunk, lambda o, c: live, torch.float32)
    want_own, _ = oracle(xs, P, n, chunk, lambda o, c: [0, 1, 2, 3], torch.float32)
    out3 = cl.outputs[3]
    assert torch.equal(out3[: 3 * geo_block], want_others[: 3 * geo_block])
    assert torch.equal(out3[3 * geo_block:], want_own[3 * geo_block:])
    assert cl.status(3) == 0


def test_native_engine_runs_the_same_rounds_as_the_python_engine(monkeypatch):
    """AAR_NATIVE_ENGINE=1 (csrc/cuda/aar_engine.cu): one C call per round instead of the Python stream / event choreography.
    Single rank, ring of two slots reused three times: bit-exact, completion word advances, counts filled."""
    from akka_allreduce_b200.parallel.engine import RoundEngine

    monkeypatch.setenv("AAR_NATIVE_ENGINE", "1")
    n = 1 << 16
    eng = RoundEngine.single_process(n, 1 << 12, max_lag=1, timeout_ms=2000.0)
    assert eng._native is not None
    for rnd in range(6):
        x = make_inputs(1, n, torch.float32, 800 + rnd)[0]
        out, counts = eng.enqueue_round(rnd, x)
        eng.wait_round(rnd)
        torch.cuda.synchronize()
        assert torch.equal(out, x) and int(counts.min()) == 1
    eng.synchronize()
    assert eng.rounds_flushed() == 6 and eng.launches == 6
    eng.close()
