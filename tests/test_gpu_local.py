"""GPU tests of the sm_100a kernels on ONE device (P emulated ranks, see akka_allreduce_b200/parallel/local_gpu.py).

Every result is compared bit-for-bit (torch.equal) with a plain PyTorch fp32 restatement of the reference's
reduce / flush (AllreduceWorker.scala:240-251,180-192) over the contributor set the scenario defines.
The scenarios mirror the reference spec's fault cases (missed scatter T9, missed reduce T11, cold catch-up
T14, ring reuse T7/T12) on the device data plane.
"""
import pytest
import torch

from akka_allreduce_b200 import _core as core

pytestmark = pytest.mark.gpu


def make_inputs(P, n, dtype, seed, device="cuda:0"):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return [(torch.randn(n, generator=g) * 3).to(dtype).to(device) for _ in range(P)]


def oracle(inputs, P, n, chunk, contributors, dtype, normalize=False, missing_owner_chunks=()):
    """contributors: callable (owner, c) -> sorted list of source ranks summed for that chunk."""
    geo = core.Geometry(P, n, chunk)
    out = torch.zeros(n, dtype=dtype, device=inputs[0].device)
    counts = []
    for owner in range(P):
        for c in range(geo.num_chunks(owner)):
            s = geo.block_start(owner) + c * chunk
            e = s + geo.chunk_len(owner, c)
            if (owner, c) in missing_owner_chunks:
                counts.append(0)
                continue
            srcs = contributors(owner, c)
            acc = torch.zeros(e - s, dtype=torch.float32, device=out.device)
            for src in srcs:  # peer-index order, from 0.0f
                acc = acc + inputs[src][s:e].float()
            if normalize and len(srcs) > 1:
                acc = torch.div(acc, torch.full_like(acc, float(len(srcs))))  # IEEE division (tensor / python scalar multiplies by 1/x)
            out[s:e] = acc.to(dtype)
            counts.append(len(srcs))
    return out, torch.tensor(counts, dtype=torch.int32, device=out.device)


# Every scenario runs on both generations of kernels: the register kernels (csrc/cuda/aar_kernels.cu) and the bulk async-copy
# engine kernels (csrc/cuda/aar_bulk.cu), the latter also with AAR_FLAG_SELF_DIRECT (own contribution read in place, own
# reduced block written straight to the output).  Geometries the bulk kernels cannot take (blocks or chunks that are not
# multiples of 16 bytes) fall back to the register kernels inside LocalGpuCluster, exactly as RoundEngine does.
_IMPL = {"impl": "reg", "self_direct": False}


@pytest.fixture(autouse=True, params=["reg", "bulk", "bulk_sd"])
def kernel_impl(request):
    _IMPL["impl"] = "bulk" if request.param.startswith("bulk") else "reg"
    _IMPL["self_direct"] = request.param == "bulk_sd"
    yield request.param
    _IMPL["impl"], _IMPL["self_direct"] = "reg", False


def cluster(*a, **kw):
    from akka_allreduce_b200.parallel.local_gpu import LocalGpuCluster

    kw.setdefault("impl", _IMPL["impl"])
    kw.setdefault("self_direct", _IMPL["self_direct"])
    return LocalGpuCluster(*a, **kw)


def test_bulk_kernels_really_run(kernel_impl):
    """Guard against a silent fallback: on an aligned geometry impl=bulk must launch aar_launch_round_bulk for every phase."""
    P, n, chunk = 4, 1 << 18, 1 << 14
    cl = cluster(P, n, chunk)
    xs = make_inputs(P, n, torch.float32, 3)
    outs, counts = cl.run_round(0, xs)
    want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.float32)
    torch.cuda.synchronize()
    for r in range(P):
        assert torch.equal(outs[r], want) and torch.equal(counts[r], wc) and cl.status(r) == 0
    assert getattr(cl, "bulk_launches", 0) == (3 * P if kernel_impl != "reg" else 0)


@pytest.mark.parametrize("P,n,chunk", [(2, 1 << 22, 1 << 18), (4, 1 << 22, 1 << 20), (8, 1 << 23, 1 << 18), (3, 3 * 70000 * 4, 100000)])
def test_large_rounds_bit_exact_many_tiles(P, n, chunk):
    """Sizes where a segment spans many tiles and a CTA owns several segments (pipelines run across segment and chunk
    boundaries), over three ring revolutions."""
    cl = cluster(P, n, chunk, max_lag=1)
    for rnd in range(5):
        xs = make_inputs(P, n, torch.float32, 1000 + rnd)
        outs, counts = cl.run_round(rnd, xs)
        want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.float32)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want), f"rank {r} round {rnd}"
            assert torch.equal(counts[r], wc) and cl.status(r) == 0


def test_in_place_round_output_aliases_input():
    """ThresholdDDP reduces gradient buckets in place (out = x): every phase must read a location before it is overwritten."""
    P, n, chunk = 4, 1 << 20, 1 << 16
    cl = cluster(P, n, chunk, max_lag=1)
    for rnd in range(3):
        xs = make_inputs(P, n, torch.float32, 2000 + rnd)
        want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.float32)
        for r in range(P):
            cl.params[r].output = xs[r].data_ptr()
        cl.outputs = xs
        outs, counts = cl.run_round(rnd, xs)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want) and torch.equal(counts[r], wc)


@pytest.mark.parametrize(
    "P,n,chunk",
    [(1, 10, 2), (2, 10, 2), (3, 9, 2), (4, 9, 2), (8, 7, 4), (2, 3, 2), (5, 1003, 17), (4, 4096, 1024), (8, 1 << 20, 1 << 16), (8, (1 << 20) + 8, 40000)],
)
def test_full_rounds_fp32_bit_exact(P, n, chunk):
    cl = cluster(P, n, chunk, max_lag=1)
    for rnd in range(4):  # slots are reused twice
        xs = make_inputs(P, n, torch.float32, 100 + rnd)
        outs, counts = cl.run_round(rnd, xs)
        want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.float32)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want), f"rank {r} round {rnd}"
            assert torch.equal(counts[r], wc)
            assert cl.status(r) == 0
        assert cl.peer_base(0) == [rnd + 1] * P and cl.rounds_flushed(0) == rnd + 1


@pytest.mark.parametrize("P,n,chunk", [(4, 4096, 512), (8, 1 << 19, 1 << 15), (3, 1001, 100)])
def test_full_rounds_bf16_bit_exact(P, n, chunk):
    cl = cluster(P, n, chunk, max_lag=2, dtype=torch.bfloat16)
    for rnd in range(4):
        xs = make_inputs(P, n, torch.bfloat16, 7 + rnd)
        outs, counts = cl.run_round(rnd, xs)
        want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.bfloat16)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want)
            assert torch.equal(counts[r], wc)


def test_missed_scatter_within_threshold_and_normalisation():
    """T9 analogue: one rank never scatters; thReduce .75 of 4 -> the other three are summed, count 3 reaches the sink."""
    P, n, chunk = 4, 8192, 1024
    for normalize in (False, True):
        cl = cluster(P, n, chunk, th_reduce=0.75, th_complete=0.75, normalize=normalize)
        xs = make_inputs(P, n, torch.float32, 5)
        outs, counts = cl.run_round(0, xs, skip=[2])
        want, wc = oracle(xs, P, n, chunk, lambda o, c: [0, 1, 3], torch.float32, normalize=normalize)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want) and torch.equal(counts[r], wc)
            assert cl.status(r) == 0
        assert cl.contributor_mask(1, 0, 0) == 0b1011


def test_exact_quorum_policy_uses_lowest_ids():
    """quorum-policy = exact reproduces the reference's `==` firing: exactly minRequired contributors."""
    P, n, chunk = 4, 2048, 256
    cl = cluster(P, n, chunk, th_reduce=0.5, exact_quorum=True)
    xs = make_inputs(P, n, torch.float32, 9)
    outs, counts = cl.run_round(0, xs)
    want, wc = oracle(xs, P, n, chunk, lambda o, c: [0, 1], torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(outs[3], want) and torch.equal(counts[3], wc)


def test_forced_round_with_no_arrivals_flushes_zeros():
    """T14 analogue (cold catch-up): nothing arrived, the host forces the round: zeros with count 0."""
    P, n, chunk = 4, 4096, 512
    cl = cluster(P, n, chunk)
    xs = make_inputs(P, n, torch.float32, 11)
    for o in cl.outputs:
        o.fill_(123.0)
    outs, counts = cl.run_round(0, xs, skip=range(P), force=True)
    torch.cuda.synchronize()
    for r in range(P):
        assert torch.count_nonzero(outs[r]) == 0 and torch.count_nonzero(counts[r]) == 0
        assert cl.status(r) == 0  # forced by the host, not a timeout


def test_missed_reduce_block_stays_zero():
    """T11 analogue: owner 3 never broadcasts; thComplete .75 completes the round, its block is zero with count 0."""
    P, n, chunk = 4, 4096, 512
    cl = cluster(P, n, chunk, th_complete=0.75)
    xs = make_inputs(P, n, torch.float32, 13)
    cl.run_phases(0, xs, scatter=range(P), reduce=[0, 1, 2], gather=range(P))
    geo = core.Geometry(P, n, chunk)
    missing = {(3, c) for c in range(geo.num_chunks(3))}
    want, wc = oracle(xs, P, n, chunk, lambda o, c: list(range(P)), torch.float32, missing_owner_chunks=missing)
    torch.cuda.synchronize()
    for r in range(P):
        assert torch.equal(cl.outputs[r], want) and torch.equal(cl.counts[r], wc)


def test_late_contributor_is_excluded_and_decision_fires_once():
    """The chunk decision is taken once per (round, chunk): a source that lands after it is not half-included, and
    re-running the reduce phase reproduces the same sum (AllreduceSpec.scala:375-385 semantics)."""
    P, n, chunk = 4, 4096, 1024
    cl = cluster(P, n, chunk, th_reduce=0.75, th_complete=1.0)
    xs = make_inputs(P, n, torch.float32, 17)
    cl.run_phases(0, xs, scatter=[0, 1, 2], reduce=range(P))
    cl.run_phases(0, xs, scatter=[3])              # late
    cl.run_phases(0, xs, reduce=range(P), gather=range(P))
    want, wc = oracle(xs, P, n, chunk, lambda o, c: [0, 1, 2], torch.float32)
    torch.cuda.synchronize()
    for r in range(P):
        assert torch.equal(cl.outputs[r], want) and torch.equal(cl.counts[r], wc)
        assert cl.contributor_mask(r, 0, 0) == 0b0111


def test_stale_slot_data_never_leaks_into_a_later_round():
    """Ring reuse without zeroing: round 2 shares slot 0 with round 0; a rank that skips round 2 must not contribute
    its round-0 data (epoch-tagged flags, replacing DataBuffer.up()'s re-zeroing, DataBuffer.scala:63-67)."""
    P, n, chunk = 4, 4096, 512
    cl = cluster(P, n, chunk, max_lag=1, th_reduce=0.5, th_complete=1.0)
    for rnd in range(2):
        cl.run_round(rnd, make_inputs(P, n, torch.float32, 20 + rnd))
    xs = make_inputs(P, n, torch.float32, 22)
    outs, counts = cl.run_round(2, xs, skip=[1])
    want, wc = oracle(xs, P, n, chunk, lambda o, c: [0, 2, 3], torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], want) and torch.equal(counts[0], wc)


def test_device_timeout_is_a_liveness_guard():
    """thReduce 1.0 but one rank is silent and nobody forces the round: after spin-timeout the kernels proceed with
    what they have and raise the status bit instead of hanging the stream."""
    from akka_allreduce_b200.ops import lib as L

    P, n, chunk = 4, 2048, 512
    cl = cluster(P, n, chunk, timeout_ms=20.0)
    xs = make_inputs(P, n, torch.float32, 31)
    outs, counts = cl.run_round(0, xs, skip=[0])
    want, wc = oracle(xs, P, n, chunk, lambda o, c: [1, 2, 3], torch.float32)
    torch.cuda.synchronize()
    assert torch.equal(outs[2], want) and torch.equal(counts[2], wc)
    assert cl.status(2) & L.ST_TIMEOUT_REDUCE


def test_rank_outside_the_lag_window_forces_itself_through():
    """Device-side catch-up (AllreduceWorker.scala:91-97 / :123-126): a rank that sees a peer has already FINISHED a round
    more than maxLag ahead stops waiting for its quorum and completes with what it has -- here nothing -- instead of
    sitting in its wait until the liveness timeout."""
    import time

    P, n, chunk = 4, 4096, 512
    cl = cluster(P, n, chunk, max_lag=1, th_reduce=1.0, th_complete=1.0, timeout_ms=20000.0)
    lay = cl.layout
    xs = make_inputs(P, n, torch.float32, 90)
    # peers 0..2 announce (through rank 3's copy of the retire words) that they finished rounds 0..4
    cl._section(3, lay.peerBaseOff, 16 * 4, torch.int32)[:3] = 5
    torch.cuda.synchronize()
    t0 = time.time()
    cl.run_phases(0, xs, reduce=[3], gather=[3])  # nobody scattered anything for round 0
    torch.cuda.synchronize()
    assert time.time() - t0 < 5.0, "the round waited for the timeout instead of forcing itself"
    assert torch.count_nonzero(cl.outputs[3]) == 0 and torch.count_nonzero(cl.counts[3]) == 0
    assert cl.status(3) == 0


@pytest.mark.parametrize("P,n,dtype", [(2, 10, torch.float32), (8, 2048, torch.float32), (4, 50000, torch.bfloat16), (8, 65536, torch.float32)])
def test_oneshot_bit_exact(P, n, dtype):
    cl = cluster(P, n, max(1, n // P), max_lag=1, dtype=dtype, oneshot=True)
    for rnd in range(3):
        xs = make_inputs(P, n, dtype, 40 + rnd)
        outs, counts = cl.run_oneshot(rnd, xs)
        acc = torch.zeros(n, dtype=torch.float32, device="cuda:0")
        for x in xs:
            acc = acc + x.float()
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], acc.to(dtype))
            assert int(counts[r].min()) == P and int(counts[r].max()) == P


def test_oneshot_threshold():
    P, n = 4, 4096
    cl = cluster(P, n, n // P, th_reduce=0.75, oneshot=True)
    xs = make_inputs(P, n, torch.float32, 50)
    outs, counts = cl.run_oneshot(0, xs, skip=[1])
    acc = (torch.zeros_like(xs[0]) + xs[0]) + xs[2] + xs[3]
    torch.cuda.synchronize()
    assert torch.equal(outs[1], acc) and int(counts[1][0]) == 3


def test_round_one_retires_before_round_zero(kernel_impl):
    """Retirement is per ring slot (AllreduceSpec.scala:675-677: round 1 completes before round 0): round 1 runs to the end while
    round 0 has not even started, then round 0, then the slots are reused by rounds 2 and 3 -- every round bit-exact, the slot
    words say which round each slot retired last, the newest-round words never go backwards."""
    P, n, chunk = 4, 1 << 16, 1 << 12
    cl = cluster(P, n, chunk, max_lag=1)
    inputs = {r: make_inputs(P, n, torch.float32, 4000 + r) for r in range(4)}
    for rnd in (1, 0, 3, 2):
        outs, counts = cl.run_round(rnd, inputs[rnd])
        want, wc = oracle(inputs[rnd], P, n, chunk, lambda o, c: list(range(P)), torch.float32)
        torch.cuda.synchronize()
        for r in range(P):
            assert torch.equal(outs[r], want) and torch.equal(counts[r], wc) and cl.status(r) == 0, f"round {rnd} rank {r}"
        if rnd == 1:
            assert cl.slot_base(0) == [[0] * P, [2] * P] and cl.peer_base(0) == [2] * P
        if rnd == 0:
            assert cl.slot_base(0) == [[1] * P, [2] * P] and cl.peer_base(0) == [2] * P, "the newest-round word is only raised"
    assert cl.slot_base(2) == [[3] * P, [4] * P] and cl.peer_base(2) == [4] * P


def test_device_ring_snapshot_restore(tmp_path):
    """Checkpoint of the device-resident ring: flags and payload of a rank survive save -> load -> restore into a fresh
    engine, which then continues the round numbering (the reference has no checkpointing)."""
    from akka_allreduce_b200.parallel.engine import RoundEngine

    n, chunk = 4096, 512
    a = RoundEngine.single_process(n, chunk, max_lag=2)
    xs = [torch.randn(n, device="cuda:0") for _ in range(3)]
    for r in range(3):
        a.enqueue_round(r, xs[r])
    snap = a.snapshot()
    path = str(tmp_path / "ring.ckpt")
    core.save_ring_snapshot(path, snap)
    back = core.load_ring_snapshot(path)
    assert (back.P, back.L, back.data_size, back.round) == (1, 3, n, 3)
    assert back.reduce_flags == snap.reduce_flags and back.scatter_flags == snap.scatter_flags and back.reduce_ring == snap.reduce_ring
    # slot r holds round r's reduced vector, its flag carries epoch r+1 and one contributor
    assert back.reduce_ring[:n] == xs[0].cpu().tolist() and back.reduce_flags[0] == (1 << 32) | 1
    b = RoundEngine.single_process(n, chunk, max_lag=2, timeout_ms=3000.0)  # device-side liveness guard: a bug fails, never hangs
    b.restore(back)
    assert b.snapshot().reduce_ring == snap.reduce_ring and b.rounds_enqueued == 3
    out, counts = b.enqueue_round(3, xs[1])
    torch.cuda.synchronize()
    assert torch.equal(out, xs[1]) and b.metrics()["rounds_enqueued"] == 4 and b.status() == 0
    a.close()
    b.close()


def test_oracle_agrees_with_cpu_protocol_engine():
    """The device data plane and the C++ message-passing engine (the CPU restatement of the reference) flush the
    same vector for the same inputs."""
    import numpy as np

    P, n, chunk = 3, 1001, 100
    xs = make_inputs(P, n, torch.float32, 60)
    host = [x.cpu().numpy() for x in xs]
    sinks = [[] for _ in range(P)]
    cfg = core.MasterConfig(total_workers=P, data_size=n, max_chunk_size=chunk, max_round=0)
    cpu = core.LocalCluster(cfg, P, lambda i: (lambda it: host[i]), lambda i: (lambda d, c, it: sinks[i].append((np.array(d), np.array(c)))))
    cpu.join_all()
    cpu.run()
    cl = cluster(P, n, chunk)
    outs, counts = cl.run_round(0, xs)
    torch.cuda.synchronize()
    for r in range(P):
        np.testing.assert_array_equal(outs[r].cpu().numpy(), sinks[r][0][0])
        np.testing.assert_array_equal(counts[r].cpu().numpy(), sinks[r][0][1])
