"""Behavioural regression suite of the worker state machine.

Every scenario of the reference spec (src/test/scala/AllreduceSpec.scala, T1-T16 in SURVEY.md
section 4) re-expressed against the C++ `Worker` with a fake cluster (tests/harness.py).
Line ranges of the original scenario are cited on each test.
"""
import numpy as np
import pytest

from harness import Harness, basic_source, custom_source


def test_flushed_output_sums_all_correct_data():
    """AllreduceSpec.scala:45-88 -- P=2, id=1, uneven blocks (2,1); sink sees input x P for rounds 0 and 1."""
    gen = lambda idx, it: idx + float(it)
    data_size, num = 3, 2
    expected = {r: [gen(i, r) * num for i in range(data_size)] for r in (0, 1)}
    seen = []

    def sink(data, count, iteration):
        assert list(data) == expected[iteration]
        seen.append(iteration)

    h = Harness(custom_source(data_size, gen), sink)
    h.init(num, 1, 1.0, 1.0, 5, data_size, 2, real_self=True)
    h.start(0)
    h.scatter([2], 0, 1, 0, 0)
    h.reduce([0, 2], 0, 1, 0, 0, 2)
    h.start(1)
    h.scatter([3], 0, 1, 0, 1)
    h.reduce([2, 4], 0, 1, 0, 1, 2)
    h.fish_complete(1, 0)
    h.fish_complete(1, 1)
    assert seen == [0, 1]
    # unlike the reference (count = Array(0)) the real contributor counts reach the sink
    assert list(h.flushed[0][2]) == [2, 2]


def test_early_receiving_reduce_then_completed_round_ignores_scatter():
    """AllreduceSpec.scala:90-132 -- future-round ReduceBlocks self-start the worker and complete round 3 first."""
    h = Harness(basic_source(8))
    h.init(4, 0, 1.0, 0.8, 5, 8, 2)
    h.start(0)
    future = 3
    h.reduce([12, 15], 0, 0, 0, future, 4)
    h.reduce([11, 10], 1, 0, 0, future, 4)
    h.reduce([10, 20], 2, 0, 0, future, 4)
    h.reduce([9, 10], 3, 0, 0, future, 4)
    h.fish_complete(0, future)
    assert h.worker.round == 0 and future in h.worker.completed
    h.outbox.clear()
    # "no longer act on completed scatter for that round"
    for i in range(4):
        h.scatter([2 * i, 2 * i], i, 0, 0, future)
    h.expect_no_msg()


def test_single_round_allreduce():
    """AllreduceSpec.scala:136-173"""
    h = Harness(basic_source(8))
    h.init(4, 0, 1.0, 0.75, 5, 8, 2)
    h.start(0)
    for i in range(4):
        h.expect_scatter([2 * i, 2 * i + 1], 0, i, 0, 0)
    for i in range(4):
        h.scatter([2 * i, 2 * i], i, 0, 0, 0)
    for i in range(4):
        h.expect_reduce([12, 12], 0, i, 0, 0, 4)
    h.reduce([12, 15], 0, 0, 0, 0, 4)
    h.reduce([11, 10], 1, 0, 0, 0, 4)
    h.reduce([10, 20], 2, 0, 0, 0, 4)
    h.reduce([9, 10], 3, 0, 0, 0, 4)
    h.expect_complete(0, 0)


def test_single_round_nasty_chunk_size():
    """AllreduceSpec.scala:176-220 -- chunks of (2,1); thReduce .9 -> 1, thComplete .8 -> 3 of 4."""
    h = Harness(basic_source(6))
    h.init(2, 0, 0.9, 0.8, 5, 6, 2)
    h.start(0)
    h.expect_scatter([0, 1], 0, 0, 0, 0)
    h.expect_scatter([2], 0, 0, 1, 0)
    h.expect_scatter([3, 4], 0, 1, 0, 0)
    h.expect_scatter([5], 0, 1, 1, 0)
    h.scatter([0, 1], 0, 0, 0, 0)
    h.scatter([2], 0, 0, 1, 0)
    h.scatter([0, 1], 1, 0, 0, 0)
    h.scatter([2], 1, 0, 1, 0)
    h.expect_reduce([0, 1], 0, 0, 0, 0, 1)
    h.expect_reduce([0, 1], 0, 1, 0, 0, 1)
    h.expect_reduce([2], 0, 0, 1, 0, 1)
    h.expect_reduce([2], 0, 1, 1, 0, 1)
    h.reduce([0, 2], 0, 0, 0, 0, 1)
    h.reduce([4], 0, 0, 1, 0, 1)
    h.reduce([6, 8], 1, 0, 0, 0, 1)
    h.expect_complete(0, 0)
    h.reduce([10], 1, 0, 1, 0, 1)
    h.expect_no_msg()


def test_single_round_nasty_chunk_size_contd():
    """AllreduceSpec.scala:222-285 -- P=3, chunk 1; thresholds .7 -> 2 and 6."""
    h = Harness(basic_source(9))
    h.init(3, 0, 0.7, 0.7, 5, 9, 1)
    h.start(0)
    for d in range(3):
        for c in range(3):
            h.expect_scatter([3 * d + c], 0, d, c, 0)
    for s in range(3):
        for c in range(3):
            h.scatter([c], s, 0, c, 0)
    for c in range(3):
        for d in range(3):
            h.expect_reduce([2 * c], 0, d, c, 0, 2)
    vals = iter(range(0, 27, 3))
    h.reduce([next(vals)], 0, 0, 0, 0, 2)
    h.reduce([next(vals)], 0, 0, 1, 0, 2)
    h.reduce([next(vals)], 0, 0, 2, 0, 2)
    h.reduce([next(vals)], 1, 0, 0, 0, 2)
    h.reduce([next(vals)], 1, 0, 1, 0, 2)
    h.reduce([next(vals)], 1, 0, 2, 0, 2)
    h.reduce([next(vals)], 2, 0, 0, 0, 2)
    h.expect_complete(0, 0)
    h.reduce([next(vals)], 2, 0, 1, 0, 2)
    h.reduce([next(vals)], 2, 0, 2, 0, 2)
    h.expect_no_msg()


def test_multi_round_allreduce():
    """AllreduceSpec.scala:287-320 -- 10 rounds, thReduce .8 -> 3, thComplete .5 -> 2."""
    h = Harness(basic_source(8))
    h.init(4, 0, 0.8, 0.5, 5, 8, 2)
    for i in range(10):
        h.start(i)
        for d in range(4):
            h.expect_scatter([2 * d + i, 2 * d + 1 + i], 0, d, 0, i)
        for s in range(4):
            h.scatter([0 + i, 1 + i], s, 0, 0, i)
        for d in range(4):
            h.expect_reduce([0 + 3 * i, 3 + 3 * i], 0, d, 0, i, 3)
        h.reduce([1, 2], 0, 0, 0, i, 3)
        h.reduce([1, 2], 1, 0, 0, i, 3)
        h.expect_complete(0, i)
        h.reduce([1, 2], 2, 0, 0, i, 3)
        h.reduce([1, 2], 3, 0, 0, i, 3)
        h.expect_no_msg()


def test_multi_round_allreduce_v2():
    """AllreduceSpec.scala:322-356 -- P=2, two chunks per block, th .6 -> 1 and .8 -> 3."""
    h = Harness(basic_source(8))
    h.init(2, 0, 0.6, 0.8, 5, 8, 2)
    for i in range(10):
        h.start(i)
        h.expect_scatter([0 + i, 1 + i], 0, 0, 0, i)
        h.expect_scatter([2 + i, 3 + i], 0, 0, 1, i)
        h.expect_scatter([4 + i, 5 + i], 0, 1, 0, i)
        h.expect_scatter([6 + i, 7 + i], 0, 1, 1, i)
        h.scatter([0 + i, 1 + i], 0, 0, 0, i)
        h.scatter([2 + i, 3 + i], 0, 0, 1, i)
        h.scatter([10 + i, 11 + i], 1, 0, 0, i)
        h.scatter([12 + i, 13 + i], 1, 0, 1, i)
        h.expect_reduce([0 + i, 1 + i], 0, 0, 0, i, 1)
        h.expect_reduce([0 + i, 1 + i], 0, 1, 0, i, 1)
        h.expect_reduce([2 + i, 3 + i], 0, 0, 1, i, 1)
        h.expect_reduce([2 + i, 3 + i], 0, 1, 1, i, 1)
        h.reduce([1, 2], 0, 0, 0, i, 1)
        h.reduce([1, 2], 0, 0, 1, i, 1)
        h.reduce([1, 2], 1, 0, 0, i, 1)
        h.expect_complete(0, i)
        h.reduce([1, 2], 1, 0, 1, i, 1)
        h.expect_no_msg()


def test_missed_scatter():
    """AllreduceSpec.scala:358-392 -- reduce fires on the 3rd arrival; the 4th is not in the sum."""
    h = Harness(basic_source(4))
    h.init(4, 0, 0.75, 0.75, 5, 4, 2)
    h.start(0)
    for d in range(4):
        h.expect_scatter([d], 0, d, 0, 0)
    h.scatter([0], 0, 0, 0, 0)
    h.expect_no_msg()
    h.scatter([2], 1, 0, 0, 0)
    h.expect_no_msg()
    h.scatter([4], 2, 0, 0, 0)
    h.scatter([6], 3, 0, 0, 0)
    for d in range(4):
        h.expect_reduce([6], 0, d, 0, 0, 3)
    h.reduce([12], 0, 0, 0, 0, 3)
    h.reduce([11], 1, 0, 0, 0, 3)
    h.reduce([10], 2, 0, 0, 0, 3)
    h.expect_complete(0, 0)
    h.reduce([9], 3, 0, 0, 0, 3)
    h.expect_no_msg()


def test_future_scatter():
    """AllreduceSpec.scala:394-445 -- round 1 proceeds while round 0 stalls; late round-0 scatter completes it."""
    h = Harness(basic_source(4))
    h.init(4, 0, 0.75, 0.75, 5, 4, 2)
    h.start(0)
    for d in range(4):
        h.expect_scatter([d], 0, d, 0, 0)
    h.scatter([2], 1, 0, 0, 0)
    h.scatter([4], 2, 0, 0, 0)
    h.reduce([11], 1, 0, 0, 0, 3)
    h.reduce([10], 2, 0, 0, 0, 3)
    h.start(1)
    h.scatter([2], 1, 0, 0, 1)
    h.scatter([4], 2, 0, 0, 1)
    h.scatter([6], 3, 0, 0, 1)
    for d in range(4):
        h.expect_scatter([d + 1], 0, d, 0, 1)
    for d in range(4):
        h.expect_reduce([12], 0, d, 0, 1, 3)
    h.scatter([0], 3, 0, 0, 0)
    h.scatter([6], 3, 0, 0, 0)  # duplicate source: must not re-fire
    for d in range(4):
        h.expect_reduce([6], 0, d, 0, 0, 3)
    h.reduce([9], 3, 0, 0, 0, 3)
    h.expect_complete(0, 0)
    h.reduce([11], 1, 0, 0, 1, 3)
    h.reduce([10], 2, 0, 0, 1, 3)
    h.reduce([9], 3, 0, 0, 1, 3)
    h.expect_complete(0, 1)
    h.expect_no_msg()


def test_missed_reduce():
    """AllreduceSpec.scala:447-479 -- maxChunkSize larger than the block; complete at 3 of 4."""
    h = Harness(basic_source(4))
    h.init(4, 0, 1.0, 0.75, 5, 4, 100)
    h.start(0)
    for d in range(4):
        h.expect_scatter([d], 0, d, 0, 0)
    for s in range(4):
        h.scatter([2 * s], s, 0, 0, 0)
    for d in range(4):
        h.expect_reduce([12], 0, d, 0, 0, 4)
    h.reduce([12], 0, 0, 0, 0, 4)
    h.expect_no_msg()
    h.reduce([11], 1, 0, 0, 0, 4)
    h.expect_no_msg()
    h.reduce([10], 2, 0, 0, 0, 4)
    h.expect_complete(0, 0)


def test_delayed_future_reduce():
    """AllreduceSpec.scala:481-529 -- two rows of the ring live at once."""
    h = Harness(basic_source(4))
    h.init(4, 0, 0.75, 0.75, 5, 4, 100)
    h.start(0)
    for d in range(4):
        h.expect_scatter([d], 0, d, 0, 0)
    h.scatter([2], 1, 0, 0, 0)
    h.scatter([4], 2, 0, 0, 0)
    h.scatter([6], 3, 0, 0, 0)
    for d in range(4):
        h.expect_reduce([12], 0, d, 0, 0, 3)
    h.start(1)
    h.scatter([3], 1, 0, 0, 1)
    h.scatter([5], 2, 0, 0, 1)
    h.scatter([7], 3, 0, 0, 1)
    for d in range(4):
        h.expect_scatter([d + 1], 0, d, 0, 1)
    for d in range(4):
        h.expect_reduce([15], 0, d, 0, 1, 3)
    h.reduce([11], 1, 0, 0, 0, 3)
    h.reduce([11], 1, 0, 0, 1, 3)
    h.reduce([10], 2, 0, 0, 0, 3)
    h.reduce([10], 2, 0, 0, 1, 3)
    h.reduce([9], 3, 0, 0, 0, 3)
    h.reduce([9], 3, 0, 0, 1, 3)
    h.expect_complete(0, 0)
    h.expect_complete(0, 1)


def _expect_basic_scatter(h, i):
    for d in range(4):
        h.expect_scatter([2 * d + i, 2 * d + 1 + i], 0, d, 0, i)


def test_simple_catchup():
    """AllreduceSpec.scala:533-559,695-712 -- rounds 0-5 never complete; 6,7,8 force r-(maxLag+1) through."""
    max_lag = 5
    h = Harness(basic_source(8))
    h.init(4, 0, 1.0, 1.0, max_lag, 8, 2)
    for i in range(6):
        h.start(i)
        _expect_basic_scatter(h, i)
        h.scatter([1.0 * (i + 1)] * 2, 1, 0, 0, i)
        h.scatter([2.0 * (i + 1)] * 2, 2, 0, 0, i)
        h.scatter([4.0 * (i + 1)] * 2, 3, 0, 0, i)
        for s in (1, 2, 3):
            h.reduce([12.0, 12.0], s, 0, 0, i, 4)
    for catchup in (6, 7, 8):
        h.start(catchup)
        done = catchup - (max_lag + 1)
        for d in range(4):
            h.expect_reduce([7.0 * (done + 1)] * 2, 0, d, 0, done, 3)
        h.expect_complete(0, done)
        _expect_basic_scatter(h, catchup)
    assert h.worker.metrics()["rounds_forced"] == 3


def test_cold_catchup():
    """AllreduceSpec.scala:561-584 -- StartAllreduce(10) on a fresh worker: 5 forced empty rounds, then all scatters."""
    h = Harness(basic_source(8))
    h.init(4, 0, 1.0, 1.0, 5, 8, 2)
    h.start(10)
    for i in range(5):
        for d in range(4):
            h.expect_reduce([0, 0], 0, d, 0, i, 0)
        h.expect_complete(0, i)
    for i in range(11):
        _expect_basic_scatter(h, i)
    h.expect_no_msg()
    # forced rounds flush zeros with zero counts
    assert all((data == 0).all() and (count == 0).all() for _, data, count in h.flushed)


def test_buffer_when_uninitialized():
    """AllreduceSpec.scala:586-603 (commented out in the reference): messages before InitWorkers are deferred."""
    h = Harness(basic_source(8))
    h.start(0)
    h.expect_no_msg()
    assert not h.worker.initialized
    h.init(4, 0, 1.0, 1.0, 5, 8, 2)
    _expect_basic_scatter(h, 0)
    assert h.worker.metrics()["deferred_before_init"] == 1


def test_sanity_multi_round_v3_out_of_order_completion():
    """AllreduceSpec.scala:608-679 -- round 1 completes before round 0."""
    h = Harness(basic_source(9))
    h.init(3, 0, 0.75, 0.75, 5, 9, 2)
    h.start(0)
    h.expect_scatter([0, 1], 0, 0, 0, 0)
    h.expect_scatter([2], 0, 0, 1, 0)
    h.expect_scatter([3, 4], 0, 1, 0, 0)
    h.expect_scatter([5], 0, 1, 1, 0)
    h.expect_scatter([6, 7], 0, 2, 0, 0)
    h.expect_scatter([8], 0, 2, 1, 0)
    for s in range(3):
        h.scatter([0, 1], s, 0, 0, 0)
    for s in range(3):
        h.scatter([2], s, 0, 1, 0)
    for d in range(3):
        h.expect_reduce([0, 2], 0, d, 0, 0, 2)
    for d in range(3):
        h.expect_reduce([4], 0, d, 1, 0, 2)
    h.start(1)
    h.scatter([10, 11], 1, 0, 0, 1)
    h.scatter([12], 1, 0, 1, 1)
    h.scatter([10, 11], 2, 0, 0, 1)
    h.scatter([12], 2, 0, 1, 1)
    h.expect_scatter([1, 2], 0, 0, 0, 1)
    h.expect_scatter([3], 0, 0, 1, 1)
    h.expect_scatter([4, 5], 0, 1, 0, 1)
    h.expect_scatter([6], 0, 1, 1, 1)
    h.expect_scatter([7, 8], 0, 2, 0, 1)
    h.expect_scatter([9], 0, 2, 1, 1)
    for d in range(3):
        h.expect_reduce([20, 22], 0, d, 0, 1, 2)
    for d in range(3):
        h.expect_reduce([24], 0, d, 1, 1, 2)
    h.reduce([11, 11], 1, 0, 0, 0, 2)
    h.reduce([11], 1, 0, 1, 1, 2)
    h.reduce([11, 11], 1, 0, 0, 1, 2)
    h.reduce([11], 1, 0, 1, 0, 2)
    h.reduce([11, 11], 2, 0, 0, 0, 2)
    h.reduce([11], 2, 0, 1, 1, 2)
    h.expect_no_msg()
    h.reduce([11, 11], 2, 0, 0, 1, 2)
    h.expect_complete(0, 1)
    h.reduce([11], 2, 0, 1, 0, 2)
    h.expect_complete(0, 0)
    assert h.worker.round == 2
    assert h.worker.metrics()["rounds_out_of_order"] == 1


def test_input_size_mismatch_raises():
    """AllreduceWorker.scala:174-176 -- source size must equal the size given at init."""
    h = Harness(basic_source(5))
    h.init(2, 0, 1.0, 1.0, 1, 8, 2)
    with pytest.raises(Exception, match="different from initialization"):
        h.start(0)


def test_reduce_block_larger_than_chunk_raises():
    """AllreduceWorker.scala:135 -- assertion on the wire chunk size."""
    h = Harness(basic_source(8))
    h.init(2, 0, 1.0, 1.0, 1, 8, 2)
    h.start(0)
    with pytest.raises(Exception, match="larger than"):
        h.reduce([1, 2, 3], 1, 0, 0, 0, 2)


def test_impossible_peer_frames_are_rejected_before_they_can_make_the_worker_do_anything():
    """A well-formed frame with impossible content (ids or sizes outside the geometry, a round absurdly far ahead) raises a
    protocol error BEFORE the peer-ahead self-start can run: the data source is not asked for a single round, nothing is sent.
    (The TCP transport drops the link that delivered it and keeps serving; csrc/core/tcp_transport.cpp.)"""
    import pytest

    fetched = []

    def source(it):
        fetched.append(it)
        return np.zeros(8, dtype=np.float32)

    h = Harness(source)
    h.init(4, 1, 1.0, 1.0, 1, 8, 2)
    for bad in (lambda: h.scatter([1, 2], 0, 1, 99, 5),            # chunk id outside my block
                lambda: h.scatter([1, 2, 3], 0, 1, 0, 5),          # payload larger than the chunk's slot
                lambda: h.scatter([1, 2], 9, 1, 0, 5),             # source id outside the cluster
                lambda: h.scatter([1, 2], 0, 1, 0, 1 << 28),       # a round that would mean hours of catch-up
                lambda: h.reduce([1, 2], 0, 1, 7, 5, 4),
                lambda: h.reduce([1, 2], 0, 3, 0, 5, 4)):          # not addressed to me
        with pytest.raises(RuntimeError):
            bad()
    assert fetched == [] and not h.outbox, "a rejected frame must not have started any round"
    h.scatter([1, 2], 0, 1, 0, 3)                                   # a legitimate peer-ahead block still self-starts
    assert fetched == [0, 1, 2, 3]
