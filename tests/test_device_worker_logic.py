"""Host bookkeeping of the device-plane worker (akka_allreduce_b200/worker.py) against a FAKE data plane, on the CPU.

The reference tests one real worker against a fake cluster (AllreduceSpec.scala:757-763); here the real master, the real TCP
control channel and the real worker bookkeeping run against a fake engine whose "kernels" are a shared in-process fabric that
sums whatever was enqueued.  Pins: every round fetched once and in order, sinks in round order, CompleteAllreduce for every
round, output buffers never reused before their round reached the sink, the input-size check, metrics."""
import threading
import time
from types import SimpleNamespace

import numpy as np
import pytest

from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput


class Fabric:
    """Stands in for NVLink + kernels: round r completes on every rank once `quorum` ranks enqueued it (sum of those)."""

    def __init__(self, world, n, quorum=None, delay_s=0.0):
        self.world, self.n, self.quorum, self.delay_s = world, n, quorum or world, delay_s
        self.lock = threading.Lock()
        self.inputs = {}     # round -> {rank: array}
        self.engines = {}

    def enqueue(self, rank, rnd, x):
        with self.lock:
            self.inputs.setdefault(rnd, {})[rank] = np.array(x, dtype=np.float32)
            self._advance()

    def _advance(self):
        # rounds complete in order on every engine (stream order), as soon as the quorum is in
        progressed = True
        while progressed:
            progressed = False
            for eng in self.engines.values():
                r = eng._flushed
                got = self.inputs.get(r, {})
                if eng.rank in got and len(got) >= self.quorum:
                    srcs = sorted(got)
                    out = np.zeros(self.n, dtype=np.float32)
                    for s in srcs:
                        out = out + got[s]
                    slot = r % eng.layout.L
                    assert eng._slot_round[slot] in (None, r - eng.layout.L), "output slot reused before its round was consumed"
                    eng.outputs[slot][:] = out
                    eng.counts[slot][:] = len(srcs)
                    eng._slot_round[slot] = r
                    eng._flushed += 1
                    progressed = True


class FakeEngine:
    def __init__(self, fabric, rank, n, max_lag):
        self.fabric, self.rank, self.data_size, self.max_lag = fabric, rank, n, max_lag
        self.layout = SimpleNamespace(L=max_lag + 1)
        self.outputs = [np.zeros(n, dtype=np.float32) for _ in range(max_lag + 1)]
        self.counts = [np.zeros(4, dtype=np.int32) for _ in range(max_lag + 1)]
        self._slot_round = [None] * (max_lag + 1)
        self._flushed = 0
        self.doorbell = -1
        self.enqueued = []
        self.algo, self.overlap, self.device = "fake", False, None
        fabric.engines[rank] = self

    def set_doorbell(self, r):
        self.doorbell = max(self.doorbell, r)

    def start_at(self, rnd):
        self._flushed = rnd
        self.started_at = rnd

    def rounds_flushed(self):
        if self.fabric.delay_s:
            time.sleep(self.fabric.delay_s)
        with self.fabric.lock:
            return self._flushed

    def enqueue_round(self, rnd, x):
        self.enqueued.append(rnd)
        self.fabric.enqueue(self.rank, rnd, x)

    def synchronize(self):
        pass

    def metrics(self):
        return {"algo": self.algo}

    def close(self):
        pass


class FakeWorker(AllreduceWorker):
    def __init__(self, fabric, *a, **kw):
        super().__init__(*a, backend="cuda", **kw)
        self.fabric = fabric

    def _make_engine(self, m):
        return FakeEngine(self.fabric, m.dest_id, m.data_size, m.max_lag)


def run_cluster(world, n, rounds, max_lag, pipeline_depth, quorum=None, th=1.0, **master_kw):
    fabric = Fabric(world, n, quorum)
    master = AllreduceMaster(world, th, 1.0, 1.0, max_lag, n, rounds - 1, 4, pipeline_depth=pipeline_depth, **master_kw)
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

        workers.append(FakeWorker(fabric, source, sink))
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    return master, workers, log


@pytest.mark.parametrize("max_lag,depth", [(0, 0), (1, 0), (1, 1), (4, 4)])
def test_rounds_are_fetched_launched_and_sunk_in_order(max_lag, depth):
    world, n, rounds = 3, 16, 40
    master, workers, log = run_cluster(world, n, rounds, max_lag, depth)
    ids = sorted(w.id for w in workers)
    assert ids == [0, 1, 2]
    for w, lg in zip(workers, log):
        assert lg.fetched == list(range(rounds))            # every round fetched exactly once, in order
        assert w.engine is None or True
        assert lg.sunk == list(range(rounds))               # delivered in order
        for it, (data, cnt) in enumerate(lg.data):
            np.testing.assert_array_equal(data, np.arange(n, dtype=np.float32) * 6 + 3 * it)
            assert cnt == world
        m = w.metrics()
        assert m["rounds_delivered"] == rounds and m["rounds_forced"] == 0 and m["max_scattered"] == rounds - 1
        # lag histogram: every delivered round is counted once, and nothing was delivered from outside the window
        assert sum(m["lag_hist"].values()) == rounds and max(m["lag_hist"]) <= max(max_lag, depth)
    assert master.metrics()["completions_counted"] == world * rounds


def test_doorbell_follows_the_masters_clock_and_lag_is_bounded_by_the_pipeline():
    world, n, rounds = 2, 8, 30
    master, workers, log = run_cluster(world, n, rounds, max_lag=2, pipeline_depth=2)
    for w in workers:
        assert w.max_round == rounds - 1
    # the master never ran more than depth+1 rounds ahead of what a worker had delivered: rounds_forced stays 0
    assert all(w.rounds_forced == 0 for w in workers)


def test_input_size_mismatch_is_rejected_like_the_reference():
    """AllreduceWorker.scala:174-176"""
    fabric = Fabric(1, 8)
    master = AllreduceMaster(1, 1.0, 1.0, 1.0, 1, 8, 0, 4)
    master.start(10)
    w = FakeWorker(fabric, lambda req: AllReduceInput(np.zeros(5, dtype=np.float32)), lambda out: None)
    with pytest.raises(Exception, match="different from initialization"):
        w.run("127.0.0.1", master.port, 10)


def test_zero_copy_input_offers_the_engines_buffer_to_the_source():
    """AllreduceWorker(zero_copy_input=True): the request carries the engine's staging buffer for that round; a source that fills
    and returns it is enqueued as is; without the option the request carries no buffer."""
    world, n, rounds = 2, 8, 6
    fabric = Fabric(world, n)
    master = AllreduceMaster(world, 1.0, 1.0, 1.0, 1, n, rounds - 1, 4, pipeline_depth=1)
    master.start(60)
    seen = [[], []]
    sunk = [[], []]

    class StagedEngine(FakeEngine):
        def __init__(self, *a):
            super().__init__(*a)
            self.stage = [np.zeros(n, dtype=np.float32) for _ in range(self.layout.L)]
            self.identical = []

        def stage_input(self, rnd):
            return self.stage[rnd % self.layout.L]

        def enqueue_round(self, rnd, x):
            self.identical.append(x is self.stage[rnd % self.layout.L])
            super().enqueue_round(rnd, x)

    class W(FakeWorker):
        def _make_engine(self, m):
            return StagedEngine(self.fabric, m.dest_id, m.data_size, m.max_lag)

    workers = []
    for i, zc in enumerate((True, False)):
        def source(req, i=i):
            seen[i].append(req.buffer is not None)
            data = np.full(n, float(i + 1) + req.iteration, dtype=np.float32)
            if req.buffer is not None:
                req.buffer[:] = data
                return AllReduceInput(req.buffer)
            return AllReduceInput(data)

        workers.append(W(fabric, source, lambda out, i=i: sunk[i].append(float(out.data[0])), zero_copy_input=zc))
    engines = []
    for w in workers:
        w.on_init = lambda ww: engines.append(ww.engine)
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    assert seen[0] == [True] * rounds and seen[1] == [False] * rounds
    by_rank = {e.rank: e for e in engines}
    zc_rank = workers[0].id
    assert all(by_rank[zc_rank].identical) and not any(by_rank[1 - zc_rank].identical)
    assert sunk[0] == sunk[1] == [3.0 + 2 * r for r in range(rounds)]


def test_resumed_master_makes_the_device_plane_continue_at_its_round(tmp_path):
    """AllreduceMaster(checkpoint=..., resume=True): a second coordinator continues the numbering of the first; the device-plane
    worker tells its engine (start_at), fetches and sinks exactly the remaining rounds, and its counters start there."""
    world, n = 2, 8
    ck = tmp_path / "m.ckpt"
    master, workers, log = run_cluster(world, n, 10, 1, 1, checkpoint=ck, checkpoint_every=3)
    assert ck.exists() and all(lg.sunk == list(range(10)) for lg in log)
    master, workers, log = run_cluster(world, n, 25, 1, 1, checkpoint=ck, resume=True)
    assert master.start_round == 10
    for w, lg in zip(workers, log):
        assert lg.fetched == list(range(10, 25)) and lg.sunk == list(range(10, 25))
        for it, (data, cnt) in zip(lg.sunk, lg.data):
            np.testing.assert_array_equal(data, np.arange(n, dtype=np.float32) * 3 + 2 * it)
        m = w.metrics()
        assert m["rounds_delivered"] == 25 and m["rounds_forced"] == 0 and sum(m["lag_hist"].values()) == 15


def test_late_joiner_reinitialises_the_device_plane_workers():
    """AllreduceMaster.scala:40-47 on the device plane: a third worker joins after two were admitted (thAllreduce 0.67 of 3), the
    master re-initialises everyone with a new generation and restarts at round 0.  The old engines are wound down first (doorbell
    rung far ahead so that waiting kernels give up, synchronised, then closed), new ones are built for the new peer table."""
    n, rounds = 8, 30
    fabrics, lock, closed = {}, threading.Lock(), []

    class TrackedEngine(FakeEngine):
        def __init__(self, *a):
            super().__init__(*a)
            self.synced = False

        def synchronize(self):
            self.synced = True

        def close(self):
            closed.append((self.rank, self.doorbell, self.synced))

    class W(FakeWorker):
        def _make_engine(self, m):
            with lock:
                fab = fabrics.setdefault(m.generation, Fabric(len(m.workers), m.data_size))
            return TrackedEngine(fab, m.dest_id, m.data_size, m.max_lag)

    master = AllreduceMaster(3, 0.67, 1.0, 1.0, 1, n, rounds - 1, 4, pipeline_depth=1)
    master.start(60)
    sunk = [[], [], []]
    gens = [[], [], []]
    workers = []
    for i in range(3):
        w = W(None, lambda req: AllReduceInput(np.full(n, 1.0, dtype=np.float32)),
              (lambda k: lambda out: sunk[k].append((out.iteration, float(out.data[0]))))(i))
        w.on_init = (lambda k: lambda ww: gens[k].append(ww.generation))(i)
        workers.append(w)
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    threads[0].start()
    threads[1].start()
    deadline = time.time() + 20
    while not sunk[0] and time.time() < deadline:   # the two-worker incarnation is really running
        time.sleep(0.01)
    threads[2].start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    # the first two workers lived through two incarnations, the third only through the second
    assert gens == [[1, 2], [1, 2], [2]]
    for k in range(3):
        its = [it for it, _ in sunk[k]]
        restart = its.index(0, 1) if k < 2 else 0          # round numbering restarts at 0 with the new incarnation
        assert its[restart:] == list(range(rounds))
        assert all(v == 3.0 for _, v in sunk[k][restart:])   # three contributors after the re-init
        if k < 2:
            assert all(v == 2.0 for _, v in sunk[k][:restart])
    # both old engines were rung out and synchronised before they were closed
    old = [c for c in closed if c[1] >= (1 << 30)]
    assert len(old) >= 2 and all(s for _, _, s in old)


class OutOfOrderFabric(Fabric):
    """A data plane that flushes rounds in whatever order their quorum lands (what the device does since retirement is per ring
    slot): round `held` is kept back until `release()` says so (the test: until round `held + 1` has reached every sink)."""

    def __init__(self, world, n, held=0, release=lambda: True):
        super().__init__(world, n)
        self.held, self.release = held, release

    def _advance(self):
        progressed = True
        while progressed:
            progressed = False
            for eng in self.engines.values():
                for r in sorted(eng.enqueued):
                    if r in eng.done:
                        continue
                    got = self.inputs.get(r, {})
                    if len(got) < self.quorum:
                        continue
                    if r == self.held and not self.release():
                        continue
                    out = np.zeros(self.n, dtype=np.float32)
                    for s in sorted(got):
                        out = out + got[s]
                    slot = r % eng.layout.L
                    assert eng._slot_round[slot] in (None, r - eng.layout.L), "output slot reused before its round was consumed"
                    eng.outputs[slot][:] = out
                    eng.counts[slot][:] = len(got)
                    eng._slot_round[slot] = r
                    eng.done.add(r)
                    progressed = True


class OutOfOrderEngine(FakeEngine):
    def __init__(self, *a):
        super().__init__(*a)
        self.done = set()

    def round_flushed(self, r):
        with self.fabric.lock:
            self.fabric._advance()
            return r in self.done

    def rounds_flushed(self):
        with self.fabric.lock:
            k = 0
            while k in self.done:
                k += 1
            return k


def test_a_later_round_completes_before_an_earlier_one_when_the_device_flushes_it_first():
    """AllreduceSpec.scala:675-677 ("multi-round allreduce v3": round 1 completes before round 0) on the device-plane worker: the
    sink and the master see round 1 first, `round` stays at 0 until round 0 lands, then advances past both; nothing is lost, no
    output slot is reused before its round reached the sink."""
    world, n, rounds = 2, 8, 12
    sunk = [[], []]
    fabric = OutOfOrderFabric(world, n, held=0, release=lambda: all(1 in s for s in sunk))
    master = AllreduceMaster(world, 1.0, 1.0, 1.0, 2, n, rounds - 1, 4, pipeline_depth=2)
    master.start(60)

    class W(FakeWorker):
        def _make_engine(self, m):
            return OutOfOrderEngine(self.fabric, m.dest_id, m.data_size, m.max_lag)

    workers = []
    for i in range(world):
        def source(req, i=i):
            return AllReduceInput(np.full(n, float(i + 1) + req.iteration, dtype=np.float32))

        def sink(out, i=i):
            assert np.all(np.asarray(out.data) == 3.0 + 2 * out.iteration), "sum of both ranks for that round"
            sunk[i].append(out.iteration)

        workers.append(W(fabric, source, sink))
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    for i, w in enumerate(workers):
        assert sorted(sunk[i]) == list(range(rounds)), "every round reached the sink exactly once"
        assert sunk[i].index(1) < sunk[i].index(0), "round 1 completed before round 0"
        assert w.metrics()["rounds_out_of_order"] >= 1 and w.round == rounds and not w.completed
    assert master.metrics()["completions_counted"] == world * rounds


@pytest.mark.parametrize("native_loop", [False, True])
def test_a_worker_that_never_hears_from_the_master_starts_rounds_because_its_peers_have(native_loop):
    """AllreduceWorker.scala:123-126 / AllreduceSpec.scala:106-124 (peer-ahead self-start): worker 1 drops every StartAllreduce of
    the master; it still fetches and launches every round, because worker 0 posts the rounds it starts on the shared board."""
    import os

    from akka_allreduce_b200 import _core as core
    from akka_allreduce_b200.parallel.board import PeerBoard

    world, n, rounds = 2, 8, 10
    fabric = Fabric(world, n)
    master = AllreduceMaster(world, 1.0, 1.0, 1.0, 1, n, rounds - 1, 4, pipeline_depth=1)
    master.start(60)
    session = f"test_selfstart_{os.getpid()}_{int(native_loop)}"
    sunk = [[], []]

    class W(FakeWorker):
        def _make_engine(self, m):
            self.board = PeerBoard(session, world, m.dest_id)
            return FakeEngine(self.fabric, m.dest_id, m.data_size, m.max_lag)

    workers = []
    for i in range(world):
        def source(req, i=i):
            return AllReduceInput(np.full(n, float(i + 1), dtype=np.float32))

        w = W(fabric, source, lambda out, i=i: sunk[i].append(out.iteration), native_loop=native_loop)
        w.drop_start = i == 1   # the master's clock never reaches this worker
        workers.append(w)
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    for t in threads:
        t.start()
    # worker ids are assigned in join order: make the one that got id 1 deaf as soon as it knows
    deadline = time.time() + 30
    while time.time() < deadline and not all(w.id >= 0 for w in workers):
        time.sleep(0.01)
    for t in threads:
        t.join(90)
    assert master.join(30)
    assert all(s == list(range(rounds)) for s in sunk)


@pytest.mark.parametrize("native_loop", [False, True])
def test_self_start_counts_the_rounds_it_started(tmp_path, native_loop):
    """The same with the deaf worker chosen up front (by construction order): its self_started_rounds metric covers every round."""
    import os

    from akka_allreduce_b200 import _core as core
    from akka_allreduce_b200.parallel.board import PeerBoard

    world, n, rounds = 2, 8, 8
    fabric = Fabric(world, n)
    master = AllreduceMaster(world, 1.0, 1.0, 1.0, 1, n, rounds - 1, 4, pipeline_depth=1)
    master.start(60)
    session = f"test_selfstart2_{os.getpid()}_{int(native_loop)}"

    class W(FakeWorker):
        def _make_engine(self, m):
            self.board = PeerBoard(session, world, m.dest_id)
            return FakeEngine(self.fabric, m.dest_id, m.data_size, m.max_lag)

    ws = [W(fabric, lambda req: AllReduceInput(np.ones(n, dtype=np.float32)), lambda out: None, native_loop=native_loop) for _ in range(world)]
    ws[1].drop_start = True
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in ws]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    assert ws[1].metrics()["self_started_rounds"] == rounds and ws[0].metrics()["self_started_rounds"] == 0
    assert ws[1].metrics()["rounds_delivered"] == rounds


def test_peer_board_is_monotone_and_a_new_incarnation_resets_its_own_word(tmp_path):
    from akka_allreduce_b200.parallel.board import PeerBoard

    a = PeerBoard("board_unit", 2, 0, directory=str(tmp_path))
    b = PeerBoard("board_unit", 2, 1, directory=str(tmp_path))
    assert a.newest() == -1
    a.post(4)
    a.post(2)                      # never goes back
    assert b.newest() == 4
    b.post(6)
    assert a.newest() == 6
    a.close()
    b.close()
    # the same session name again (a job restarted on the same port): both words start over, nothing stale is read
    a2 = PeerBoard("board_unit", 2, 0, directory=str(tmp_path))
    b2 = PeerBoard("board_unit", 2, 1, initial=9, directory=str(tmp_path))   # a resumed master: rounds below 10 count as started
    assert a2.newest() == 9
    a2.close()
    b2.close(unlink=True)
