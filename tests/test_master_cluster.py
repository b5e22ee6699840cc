"""Coordinator and multi-rank tests the reference lacks ("Implication").

Master: admission quorum, round quorum, late joiner, termination (AllreduceMaster.scala:36-97).
LocalCluster: full master + N workers through the deterministic in-process router, checked against
the golden trace of SURVEY.md appendix A, plus straggler / fault-injection scenarios.
"""
import numpy as np
import pytest

from akka_allreduce_b200 import _core as core


class Recorder:
    def __init__(self):
        self.sent = []
        self.port = core.CallbackPort(lambda to, msg: self.sent.append((to, msg)))

    def of(self, cls):
        return [(to, m) for to, m in self.sent if isinstance(m, cls)]


def cfg(**kw):
    base = dict(total_workers=2, th_allreduce=1.0, th_reduce=1.0, th_complete=1.0, max_lag=1, data_size=10, max_round=3, max_chunk_size=2)
    base.update(kw)
    return core.MasterConfig(**base)


def test_admission_waits_for_quorum_then_inits_and_starts_round0():
    r = Recorder()
    m = core.Master(cfg(total_workers=3), r.port, 0)
    m.member_up(11, "worker")
    m.member_up(12, "worker")
    assert not m.started and not r.sent
    m.member_up(99, "master")  # non-worker roles never register
    assert not m.started
    m.member_up(13, "worker")
    inits = r.of(core.InitWorkers)
    assert [to for to, _ in inits] == [11, 12, 13]
    assert [i.dest_id for _, i in inits] == [0, 1, 2]  # ids in join order
    assert inits[0][1].workers == {0: 11, 1: 12, 2: 13}
    assert inits[0][1].data_size == 10 and inits[0][1].max_chunk_size == 2
    assert [(to, s.round) for to, s in r.of(core.StartAllreduce)] == [(11, 0), (12, 0), (13, 0)]
    assert m.round == 0


def test_admission_threshold_truncates():
    r = Recorder()
    m = core.Master(cfg(total_workers=4, th_allreduce=0.6), r.port, 0)  # (4 * 0.6).toInt = 2
    m.member_up(1, "worker")
    assert not m.started
    m.member_up(2, "worker")
    assert m.started


def test_round_quorum_is_float_compare_and_stops_at_max_round():
    r = Recorder()
    # 4 * 0.6 = 2.4: admission needs 2, completion needs 3 (no truncation, AllreduceMaster.scala:62)
    m = core.Master(cfg(total_workers=4, th_allreduce=0.6, max_round=1), r.port, 0)
    for a in (1, 2, 3, 4):
        m.member_up(a, "worker")
    gen = m.generation
    m.receive(core.CompleteAllreduce(0, 0, gen))
    m.receive(core.CompleteAllreduce(1, 0, gen))
    assert m.round == 0
    m.receive(core.CompleteAllreduce(1, 0, gen))  # duplicate completion is not counted twice
    assert m.round == 0
    m.receive(core.CompleteAllreduce(2, 5, gen))  # wrong round is ignored
    assert m.round == 0
    m.receive(core.CompleteAllreduce(2, 0, gen))
    assert m.round == 1
    for w in range(4):
        m.receive(core.CompleteAllreduce(w, 1, gen))
    assert m.round == 1 and m.finished  # round < maxRound guard


def test_late_joiner_reinitialises_everyone_from_round0():
    r = Recorder()
    m = core.Master(cfg(total_workers=2, th_allreduce=0.5), r.port, 0)
    m.member_up(1, "worker")
    assert m.started and m.generation == 1
    m.receive(core.CompleteAllreduce(0, 0, 1))
    assert m.round == 1
    m.member_up(2, "worker")
    assert m.generation == 2 and m.round == 0
    assert [i.dest_id for _, i in r.of(core.InitWorkers)][-2:] == [0, 1]
    m.receive(core.CompleteAllreduce(0, 0, 1))  # stale generation
    assert m.metrics()["completions_ignored"] == 1


def test_termination_removes_worker_and_ids_are_redensified_on_reinit():
    r = Recorder()
    m = core.Master(cfg(total_workers=3, th_allreduce=0.6), r.port, 0)  # admission at 1 worker
    m.member_up(1, "worker")
    m.member_up(2, "worker")
    m.member_up(3, "worker")
    m.terminated(2)
    assert m.workers == {0: 1, 2: 3}
    m.member_up(4, "worker")
    assert m.workers == {0: 1, 1: 3, 2: 4}  # dense again (the reference would leave a hole)


def test_pipelined_round_clock():
    r = Recorder()
    m = core.Master(cfg(total_workers=1, max_round=10, pipeline_depth=2, max_lag=2), r.port, 0)
    m.member_up(1, "worker")
    assert m.round == 2  # rounds 0,1,2 issued up front
    m.receive(core.CompleteAllreduce(0, 0, 1))
    assert m.round == 3
    m.receive(core.CompleteAllreduce(0, 2, 1))  # quorum for 2 arrives before 1: clock waits for 1
    assert m.round == 3
    m.receive(core.CompleteAllreduce(0, 1, 1))
    assert m.round == 5 and m.quorum_round == 2
    # the depth is capped at maxLag: further ahead, workers would force-complete rounds that nothing is wrong with
    capped = core.Master(cfg(total_workers=1, max_round=10, pipeline_depth=5, max_lag=1), Recorder().port, 0)
    capped.member_up(1, "worker")
    assert capped.round == 1


# ------------------------------------------------------------------ full in-process clusters


def make_cluster(n, sinks, **kw):
    c = cfg(total_workers=n, **kw)
    src = lambda i: (lambda it: np.arange(c.data_size, dtype=np.float32) + np.float32(it))
    snk = lambda i: (lambda data, count, it: sinks[i].append((it, np.array(data), np.array(count))))
    return core.LocalCluster(c, n, src, snk)


def test_golden_trace_config1():
    """BASELINE.json config 1 / SURVEY.md appendix A: P=2, dataSize=10, chunk=2, thresholds 1.0 -> out[i] = 2*(i+r)."""
    sinks = [[], []]
    cl = make_cluster(2, sinks, max_round=100)
    cl.join_all()
    cl.run()
    assert cl.master().finished and cl.master().round == 100
    for w in range(2):
        assert [it for it, _, _ in sinks[w]] == list(range(101))  # rounds 0..100 inclusive
        for it, data, count in sinks[w]:
            np.testing.assert_array_equal(data, 2 * (np.arange(10, dtype=np.float32) + it))
            assert list(count) == [2] * 6
    # 24 data messages + 2 starts + 2 completes per round, + 2 inits
    assert cl.delivered == 101 * 28 + 2
    assert cl.worker(0).metrics()["scatter_sent"] == 101 * 6


def test_reference_default_thresholds_give_partial_sums():
    """Appendix A: with thReduce .9 / thComplete .8 and P=2 each chunk is 'reduced' on its first arrival."""
    sinks = [[], []]
    cl = make_cluster(2, sinks, th_reduce=0.9, th_complete=0.8, max_round=2)
    cl.join_all()
    cl.run()
    for it, data, count in sinks[0]:
        assert set(count) <= {0, 1}  # never the full sum


@pytest.mark.parametrize("P,data,chunk", [(3, 9, 2), (4, 9, 2), (8, 7, 4), (5, 1003, 17), (2, 3, 2)])
def test_uneven_partitions_sum_exactly(P, data, chunk):
    sinks = [[] for _ in range(P)]
    cl = make_cluster(P, sinks, data_size=data, max_chunk_size=chunk, max_round=3)
    cl.join_all()
    cl.run()
    assert cl.master().finished
    for w in range(P):
        assert len(sinks[w]) == 4
        for it, out, count in sinks[w]:
            np.testing.assert_array_equal(out, P * (np.arange(data, dtype=np.float32) + it))
            assert (count == P).all()


def test_straggler_is_tolerated_and_catches_up():
    """One worker's mailbox is parked (a straggler); with thresholds .75 and thAllreduce .75 the other three keep
    completing rounds; when released, the straggler force-completes the rounds that fell out of the maxLag window."""
    P = 4
    sinks = [[] for _ in range(P)]
    cl = make_cluster(P, sinks, th_allreduce=0.75, th_reduce=0.75, th_complete=0.75, max_lag=2, data_size=8, max_chunk_size=2, max_round=9)
    cl.join_all()
    cl.hold(3)
    cl.run()
    assert cl.master().finished
    for w in range(3):
        assert [it for it, _, _ in sinks[w]] == list(range(10))
        for it, out, count in sinks[w]:
            # the straggler's own block (elements 6,7) never arrives -> zeros with count 0; the rest sums 3 contributors
            np.testing.assert_array_equal(out[:6], 3 * (np.arange(6, dtype=np.float32) + it))
            np.testing.assert_array_equal(out[6:], 0)
            assert list(count) == [3, 3, 3, 0]
    assert sinks[3] == []
    # released: its mailbox replays in order, so it completes every round from the parked traffic (own block summed
    # over all 4 now that its own scatter is in) without any forced completion
    cl.release(3)
    cl.run()
    assert [it for it, _, _ in sinks[3]] == list(range(10))
    assert cl.worker(3).metrics()["rounds_forced"] == 0
    assert cl.worker(3).max_scattered == 9 and cl.worker(3).round == 10


def test_worker_outside_the_lag_window_force_completes():
    """All data addressed to worker 3 is lost; it only hears the master's clock. Once maxRound - round > maxLag it pushes
    the oldest round through with whatever it has (AllreduceWorker.scala:91-97), still scattering every round once."""
    P = 4
    sinks = [[] for _ in range(P)]
    cl = make_cluster(P, sinks, th_allreduce=0.75, th_reduce=0.75, th_complete=0.75, max_lag=2, data_size=8, max_chunk_size=2, max_round=9)
    lost = core.LocalCluster.worker_addr(3)
    cl.set_drop(lambda frm, to, msg: to == lost and isinstance(msg, (core.ScatterBlock, core.ReduceBlock)))
    cl.join_all()
    cl.run()
    assert cl.master().finished
    m = cl.worker(3).metrics()
    assert m["rounds_forced"] == 7 and m["max_lag_seen"] == 3  # rounds 0..6 forced; 7, 8, 9 stay open in the window
    assert cl.worker(3).max_scattered == 9 and cl.worker(3).round == 7
    for it, out, count in sinks[3]:
        assert (out == 0).all() and (count == 0).all()
    # the other three were never held back by it, and they did receive worker 3's scatters
    for w in range(3):
        assert [it for it, _, _ in sinks[w]] == list(range(10))


def test_dropped_messages_within_threshold():
    P = 4
    sinks = [[] for _ in range(P)]
    cl = make_cluster(P, sinks, th_reduce=0.75, th_complete=0.75, data_size=8, max_chunk_size=2, max_round=4)
    # drop every ScatterBlock sent by worker 2 (address 3)
    cl.set_drop(lambda frm, to, msg: isinstance(msg, core.ScatterBlock) and frm == core.LocalCluster.worker_addr(2))
    cl.join_all()
    cl.run()
    assert cl.master().finished and cl.dropped == 5 * 4
    for it, out, count in sinks[0]:
        assert (count[count > 0] == 3).all()


def test_cpu_baseline_runs():
    sec, checksum, msgs = core.cpu_baseline_seconds(cfg(total_workers=2, max_round=100), 2)
    assert sec > 0 and msgs == 101 * 28 + 2
    assert checksum == float(sum(2 * (i + 100) for i in range(10)))


def test_start_round_continues_the_numbering_instead_of_restarting_at_zero():
    """Resume support (not in the reference, which always restarts at round 0): MasterConfig.start_round travels in InitWorkers,
    workers fetch / flush exactly the rounds start..maxRound, thresholds and the ring work on absolute round numbers."""
    sinks = [[], [], []]
    cl = make_cluster(3, sinks, data_size=9, max_chunk_size=2, max_round=12, max_lag=2, start_round=7)
    cl.join_all()
    cl.run()
    for s in sinks:
        assert [it for it, _, _ in s] == list(range(7, 13))
        for it, data, count in s:
            np.testing.assert_array_equal(data, 3 * (np.arange(9, dtype=np.float32) + it))
    # partial thresholds: same round numbering, every element is the sum of `count` contributors
    sinks = [[], [], []]
    cl = make_cluster(3, sinks, data_size=9, max_chunk_size=2, max_round=12, max_lag=2, start_round=7, th_reduce=0.7, th_complete=0.7)
    cl.join_all()
    cl.run()
    for s in sinks:
        assert sorted(it for it, _, _ in s) == list(range(7, 13))


def test_start_round_beyond_max_round_finishes_immediately():
    sinks = [[], []]
    cl = make_cluster(2, sinks, max_round=5, start_round=6)
    cl.join_all()
    cl.run()
    assert cl.master().finished and sinks == [[], []]
