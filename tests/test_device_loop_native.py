"""The device-plane round loop in C++ (csrc/core/device_loop.cpp) against fake data planes, on the CPU.

Two layers: (1) `_core.DeviceLoop` alone, its data plane a set of C function pointers (ctypes thunks here, the entry points of
csrc/cuda/aar_engine.cu on a GPU box); (2) `AllreduceWorker(native_loop=True)` in the same fake-fabric cluster as
tests/test_device_worker_logic.py, so every bookkeeping property pinned there holds for the C++ loop as well.
Reference behaviour: AllreduceWorker.scala:84-104 (StartAllreduce), :123-126 (self-start), :174-176 (size check), :253-268 (complete)."""
import ctypes as C
import threading
from types import SimpleNamespace

import numpy as np
import pytest

from akka_allreduce_b200 import AllreduceMaster, AllReduceInput
from akka_allreduce_b200 import _core as core

from test_device_worker_logic import Fabric, FakeWorker, OutOfOrderEngine, OutOfOrderFabric


class FakePlane:
    """enqueue / doorbell / flushed / synchronize as C callbacks; the test decides which rounds have flushed."""

    ENQ = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint, C.c_void_p, C.c_ulonglong)
    I2 = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)
    I1 = C.CFUNCTYPE(C.c_int, C.c_void_p)

    def __init__(self, slots=2, max_lag=1, n=8, flush_after_polls=None, fail_round=None):
        self.enqueued, self.doorbells, self.flushed, self.synced = [], [], set(), 0
        self.polls, self.flush_after_polls, self.fail_round = {}, flush_after_polls, fail_round
        self.args, self.events = {}, []
        self._cbs = (self.ENQ(self._enq), self.I2(self._db), self.I2(self._fl), self.I1(self._sy))
        a = [C.cast(f, C.c_void_p).value for f in self._cbs]
        self.plane = core.DevicePlane(engine=0x1000, enqueue=a[0], doorbell=a[1], flushed=a[2], synchronize=a[3], slots=slots, max_lag=max_lag,
                                      data_size=n, extra_flags=5)

    def _enq(self, eng, rnd, i, o, c, flags, stream, delay):
        if rnd == self.fail_round:
            return 7
        self.enqueued.append(rnd)
        self.events.append(("enq", rnd))
        self.args[rnd] = (eng, i, o, c, flags, stream, delay)
        return 0

    def _db(self, eng, newest):
        self.doorbells.append(newest)
        return 0

    def _fl(self, eng, rnd):
        if self.flush_after_polls is not None and rnd in self.enqueued:   # a device that makes progress while it is polled
            self.polls[rnd] = self.polls.get(rnd, 0) + 1
            if self.polls[rnd] >= self.flush_after_polls:
                self.flushed.add(rnd)
        return int(rnd in self.flushed)

    def _sy(self, eng):
        self.synced += 1
        return 0


def make_loop(fp, n=8, start_round=0, gen=3):
    rec = SimpleNamespace(fetched=[], sunk=[], sent=[])

    def source(r):
        rec.fetched.append(r)
        return (0x100 + r, 0x200 + r, 0x300 + r, 0x400, n)

    port = core.CallbackPort(lambda to, msg: rec.sent.append((to, msg)))
    def sink(r, slot):
        rec.sunk.append((r, slot))
        fp.events.append(("sink", r))

    loop = core.DeviceLoop(source, sink, port)
    loop.set_plane(fp.plane, 2, gen, 55, start_round)
    rec.port = port
    return loop, rec


def test_start_fetches_and_launches_every_round_once_in_order_with_the_callers_pointers():
    fp = FakePlane(slots=3, max_lag=2)
    loop, rec = make_loop(fp)
    loop.on_start(1)
    loop.on_start(1)      # a repeated StartAllreduce launches nothing twice
    loop.on_start(2)
    assert rec.fetched == [0, 1, 2] and fp.enqueued == [0, 1, 2]
    assert fp.doorbells == [1, 1, 2]
    eng, i, o, c, flags, stream, delay = fp.args[1]
    assert (eng, i, o, c, flags, stream, delay) == (0x1000, 0x101, 0x201, 0x301, 5, 0x400, 0)
    assert loop.max_round == 2 and loop.max_scattered == 2 and loop.round == 0 and loop.busy
    assert loop.deliver_flushed() == 0 and rec.sunk == []


def test_rounds_reach_the_sink_in_the_order_the_device_flushes_them():
    """AllreduceSpec.scala:675-677: round 1 completes before round 0; `round` only advances past a contiguous prefix."""
    fp = FakePlane(slots=3, max_lag=2)
    loop, rec = make_loop(fp)
    loop.on_start(2)
    fp.flushed.add(1)
    assert loop.deliver_flushed() == 1
    assert rec.sunk == [(1, 1)] and loop.round == 0 and loop.completed == [1]
    fp.flushed.update({0, 2})
    assert loop.deliver_flushed() == 2
    assert [r for r, _ in rec.sunk] == [1, 0, 2] and loop.round == 3 and loop.completed == [] and not loop.busy
    # CompleteAllreduce for every round, to the master's address, with this worker's id and generation
    assert [(to, m.src_id, m.round, m.generation) for to, m in rec.sent] == [(55, 2, 1, 3), (55, 2, 0, 3), (55, 2, 2, 3)]
    st = loop.stats()
    assert st["rounds_delivered"] == 3 and st["rounds_out_of_order"] == 1 and st["rounds_launched"] == 3 and st["rounds_forced"] == 0


def test_a_ring_slot_is_not_relaunched_before_its_previous_round_reached_the_sink():
    """Round r reuses the output and count buffers of round r - L: the loop blocks on the device (not on the master) until that
    round has been delivered."""
    fp = FakePlane(slots=2, max_lag=1, flush_after_polls=3)
    loop, rec = make_loop(fp)
    loop.on_start(5)
    assert fp.enqueued == list(range(6))
    for r in range(2, 6):   # every launch of r >= L came after the sink of r - L
        assert fp.events.index(("sink", r - 2)) < fp.events.index(("enq", r)), fp.events
    assert [r for r, _ in rec.sunk][:4] == [0, 1, 2, 3]
    st = loop.stats()
    assert st["rounds_forced"] >= 1, "rounds more than maxLag behind the newest count as forced (catch-up)"
    assert fp.doorbells[0] == 5


def test_size_mismatch_and_failed_launch_are_errors():
    fp = FakePlane(n=8)
    loop, rec = make_loop(fp, n=9)
    with pytest.raises(ValueError, match="Input data size 9 is different from initialization time 8!"):
        loop.on_start(0)
    fp2 = FakePlane(n=8, fail_round=1)
    loop2, _ = make_loop(fp2)
    with pytest.raises(RuntimeError, match="round 1 failed with code 7"):
        loop2.on_start(1)
    assert fp2.enqueued == [0] and loop2.max_scattered == 0


def test_messages_generation_filter_and_init_callback():
    fp = FakePlane()
    rec = SimpleNamespace(fetched=[], inits=[])
    loop = core.DeviceLoop(lambda r: (rec.fetched.append(r), (1, 2, 3, 4, 8))[1], lambda r, s: None, core.CallbackPort(lambda to, m: None))
    loop.receive(core.StartAllreduce(0))            # before InitWorkers: nothing to start
    assert rec.fetched == []

    def on_init(m):
        rec.inits.append(m.dest_id)
        loop.set_plane(fp.plane, m.dest_id, m.generation, m.master, m.start_round)

    loop.set_init(on_init)
    loop.receive(core.InitWorkers({0: 1, 1: 2}, 9, 1, 1.0, 1.0, 1, 8, 4, 4, 10))
    assert rec.inits == [1] and loop.id == 1 and loop.generation == 4 and loop.start_round == 10 and loop.round == 10 and loop.max_round == 9
    loop.receive(core.StartAllreduce(10, 3))        # another incarnation's message
    assert rec.fetched == []
    loop.receive(core.StartAllreduce(10, 4))
    loop.receive(core.StartAllreduce(11))           # generation 0 = unspecified
    assert rec.fetched == [10, 11] and fp.enqueued == [10, 11]


def test_peer_ahead_starts_rounds_without_the_master():
    """AllreduceWorker.scala:123-126 through the board."""
    fp = FakePlane(slots=4, max_lag=3)
    loop, rec = make_loop(fp)
    board = SimpleNamespace(posted=[], newest=-1)
    loop.set_board(lambda r: board.posted.append(r), lambda: board.newest)
    assert not loop.peer_ahead()
    board.newest = 2
    assert loop.peer_ahead() and not loop.peer_ahead()
    assert fp.enqueued == [0, 1, 2] and board.posted == [2]
    assert loop.stats()["self_started_rounds"] == 3


def test_drain_rings_the_doorbell_far_ahead_and_delivers_what_flushed():
    fp = FakePlane()
    loop, rec = make_loop(fp)
    loop.on_start(1)
    fp.flushed.add(0)
    loop.drain()
    assert fp.doorbells[-1] == 1 << 30 and fp.synced == 1 and [r for r, _ in rec.sunk] == [0]


# ---------------------------------------------------------------------------------------------- the worker on the C++ loop
def run_cluster(world, n, rounds, max_lag, pipeline_depth, fabric=None, engine_cls=None, th=1.0):
    fabric = fabric or Fabric(world, n)
    master = AllreduceMaster(world, th, 1.0, 1.0, max_lag, n, rounds - 1, 4, pipeline_depth=pipeline_depth)
    master.start(60)
    log = [SimpleNamespace(fetched=[], sunk=[], data=[]) for _ in range(world)]
    workers = []
    for i in range(world):
        def source(req, i=i):
            log[i].fetched.append(req.iteration)
            return AllReduceInput(np.arange(n, dtype=np.float32) * (i + 1) + req.iteration)

        def sink(out, i=i):
            log[i].sunk.append(out.iteration)
            log[i].data.append((np.array(out.data), int(out.count[0])))

        w = FakeWorker(fabric, source, sink, native_loop=True)
        if engine_cls is not None:
            w._make_engine = lambda m, w=w: engine_cls(fabric, m.dest_id, m.data_size, m.max_lag)
        workers.append(w)
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    return master, workers, log


@pytest.mark.parametrize("max_lag,depth", [(0, 0), (1, 1), (4, 4)])
def test_worker_on_the_native_loop_fetches_launches_and_sinks_every_round(max_lag, depth):
    world, n, rounds = 3, 16, 40
    master, workers, log = run_cluster(world, n, rounds, max_lag, depth)
    assert sorted(w.id for w in workers) == [0, 1, 2]
    for w, lg in zip(workers, log):
        assert w._loop is not None and not w._direct
        assert lg.fetched == list(range(rounds)) and lg.sunk == list(range(rounds))
        for it, (data, cnt) in enumerate(lg.data):
            np.testing.assert_array_equal(data, np.arange(n, dtype=np.float32) * 6 + 3 * it)
            assert cnt == world
        m = w.metrics()
        assert m["rounds_delivered"] == rounds and m["rounds_forced"] == 0 and m["max_scattered"] == rounds - 1
        assert sum(m["lag_hist"].values()) == rounds and max(m["lag_hist"]) <= max(max_lag, depth)
    assert master.metrics()["completions_counted"] == world * rounds


def test_worker_on_the_native_loop_delivers_out_of_order():
    """The fabric holds round 0 back until round 1 has been flushed everywhere: sinks see 1 before 0, the master still gets both."""
    world, n, rounds = 2, 8, 6
    state = SimpleNamespace(workers=None)
    fabric = OutOfOrderFabric(world, n, held=0, release=lambda: all(1 in lg.sunk for lg in state.log))
    state.log = []
    master = AllreduceMaster(world, 1.0, 1.0, 1.0, 2, n, rounds - 1, 4, pipeline_depth=2)
    master.start(60)
    workers = []
    for i in range(world):
        lg = SimpleNamespace(sunk=[])
        state.log.append(lg)
        w = FakeWorker(fabric, lambda req, i=i: AllReduceInput(np.full(n, float(i + 1 + req.iteration), dtype=np.float32)),
                       lambda out, lg=lg: lg.sunk.append(out.iteration), native_loop=True)
        w._make_engine = lambda m, w=w: OutOfOrderEngine(fabric, m.dest_id, m.data_size, m.max_lag)
        workers.append(w)
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    for w, lg in zip(workers, state.log):
        assert sorted(lg.sunk) == list(range(rounds))
        assert lg.sunk.index(1) < lg.sunk.index(0), lg.sunk
        assert w.metrics()["rounds_out_of_order"] >= 1


# ---------------------------------------------------------------------------------------------- properties
def test_random_schedules_keep_the_loop_invariants():
    """Random interleavings of StartAllreduce, board updates, device flushes and deliveries (hypothesis): every round is fetched
    once and in order, launched once, sunk once and only after the device flushed it, never relaunched into a ring slot whose
    previous round has not reached the sink, CompleteAllreduce follows every sink, and `round` is the oldest round not yet sunk."""
    from hypothesis import given, settings
    from hypothesis import strategies as st

    ops = st.lists(st.tuples(st.sampled_from(["start", "board", "flush", "deliver"]), st.integers(0, 40)), min_size=1, max_size=60)

    @settings(max_examples=150, deadline=None)
    @given(slots=st.integers(1, 4), polls=st.integers(1, 3), start_round=st.integers(0, 5), ops=ops)
    def run(slots, polls, start_round, ops):
        fp = FakePlane(slots=slots, max_lag=slots - 1, flush_after_polls=polls)
        loop, rec = make_loop(fp, start_round=start_round)
        board = SimpleNamespace(newest=start_round - 1)
        loop.set_board(lambda r: None, lambda: board.newest)
        top = start_round - 1
        for op, v in ops:
            if op == "start":
                top = max(top, min(start_round + v, top + 6))     # the master's clock moves at most a few rounds at a time
                loop.on_start(top)
            elif op == "board":
                board.newest = max(board.newest, min(start_round + v, top + 3))
                loop.peer_ahead()
                top = max(top, board.newest)
            elif op == "flush" and fp.enqueued:
                fp.flushed.add(fp.enqueued[v % len(fp.enqueued)])
            else:
                loop.deliver_flushed()
        # invariants
        launched = list(range(start_round, loop.max_scattered + 1))
        assert rec.fetched == launched and fp.enqueued == launched
        sunk = [r for r, _ in rec.sunk]
        assert len(sunk) == len(set(sunk)) and all(r in fp.flushed for r in sunk)
        assert all(slot == r % slots for r, slot in rec.sunk)
        for r in launched:
            if r - slots >= start_round:
                assert fp.events.index(("sink", r - slots)) < fp.events.index(("enq", r))
        assert [m.round for _, m in rec.sent] == sunk
        not_sunk = [r for r in launched if r not in sunk]
        assert loop.round == (min(not_sunk) if not_sunk else loop.max_scattered + 1)
        assert loop.completed == sorted(r for r in sunk if r > loop.round)
        assert loop.max_round == top and loop.max_scattered == top
        assert loop.stats()["rounds_delivered"] == len(sunk)

    run()
