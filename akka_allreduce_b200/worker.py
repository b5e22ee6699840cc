"""AllreduceWorker: one per process (one per GPU on the device data plane).

    worker = AllreduceWorker(data_source, data_sink, backend="cuda")
    worker.run(master_host, master_port)

Same contract as the reference actor (AllreduceWorker.scala:9-10): the worker pulls a `dataSize` buffer from
`data_source(AllReduceInputRequest(round))` for every round, in order, exactly once, and pushes the reduced
result to `data_sink(AllReduceOutput(data, count, round))`.

backend="cpu"  : the C++ message-passing engine (csrc/core/worker.cpp) over TCP -- bit-compatible restatement of
                 the reference, used for plumbing runs and as the oracle.
backend="cuda" : control messages (InitWorkers / StartAllreduce / CompleteAllreduce) over the same TCP channel, data
                 over NVLink by the sm_100a kernels.  Up to maxLag + 1 rounds of one worker are in flight; a round reaches the
                 sink (and CompleteAllreduce the master) as soon as the device has flushed it, so round r+1 can complete before
                 round r (AllreduceWorker.scala:253-268, AllreduceSpec.scala:675-677).  A worker that falls more than maxLag
                 rounds behind the newest round is forced through by the device doorbell (the catch-up loop of
                 AllreduceWorker.scala:91-97 without a host round-trip), and a worker whose peers have started a round it has not
                 heard of from the master starts that round itself (peer-ahead self-start, AllreduceWorker.scala:123-126): the
                 peers' progress is read from a board in shared host memory, next to the data that is already in its ring.
"""
from __future__ import annotations

import os
import time
from typing import Callable

import numpy as np

from . import _core as core
from .api import AllReduceInput, AllReduceInputRequest, AllReduceOutput, DataSink, DataSource


class AllreduceWorker:
    def __init__(self, data_source: DataSource, data_sink: DataSink, backend: str = "cuda", *, device: int | None = None,
                 dtype: str = "fp32", algo: str = "auto", normalize: bool = False, exact_quorum: bool = False,
                 spin_timeout_ms: float = 0.0, grid: int = 0, session: str | None = None, want_multicast: bool = True,
                 zero_copy_input: bool = False, on_init: Callable[["AllreduceWorker"], None] | None = None,
                 native_loop: bool | None = None):
        if backend not in ("cpu", "cuda"):
            raise ValueError("backend must be 'cpu' or 'cuda'")
        self.data_source, self.data_sink, self.backend = data_source, data_sink, backend
        self.device, self.dtype, self.algo = device, dtype, algo
        self.normalize, self.exact_quorum, self.spin_timeout_ms, self.grid = normalize, exact_quorum, spin_timeout_ms, grid
        self.session, self.want_multicast, self.on_init = session, want_multicast, on_init
        self.zero_copy_input = zero_copy_input
        # the round loop in C++ (csrc/core/device_loop.cpp): Python is called for the source, the sink and InitWorkers only
        self.native_loop = (os.environ.get("AAR_NATIVE_LOOP", "1") not in ("", "0")) if native_loop is None else bool(native_loop)
        self.drop_start = False     # fault injection: the master's StartAllreduce never reaches this worker (tests of the self-start)
        self._loop = None
        self._loop_error = None
        self._direct = False
        self._plane_callbacks = None   # ctypes thunks of a Python data plane (tests): must outlive the plane
        self.node: core.WorkerNode | None = None
        self.id = -1
        self.generation = 0
        # device data plane state
        self.engine = None
        self.symm = None
        self.rdv = None
        self.max_round = -1
        self.max_scattered = -1
        self.delivered = 0          # rounds handed to the sink, counted from round 0 (start_round + the ones of this incarnation)
        self.round = 0              # oldest round that has not reached the sink yet (the reference's `round`)
        self.completed: set[int] = set()   # rounds >= self.round that have (the reference's `completed`, AllreduceWorker.scala:253-268)
        self.out_of_order = 0       # rounds that reached the sink while an older round was still in flight
        self.self_starts = 0        # rounds started because a peer was ahead of the master's message to this worker
        self.board = None           # shared host memory: newest round every peer has started
        self.rounds_forced = 0
        self.lag_hist: dict[int, int] = {}   # (master's newest round - round) at the moment a round reached the sink
        self.master_lost = False    # run() ended because the coordinator's link closed without a Shutdown
        self.start_round = 0        # first round of this incarnation (> 0 when the master resumed from a checkpoint)
        self._t_first: float | None = None   # wall clock of the first / latest delivery (rounds_per_sec in metrics())
        self._t_last = 0.0
        self._stream = None
        self._inputs = {}           # round -> input tensor kept alive until the round is flushed
        self._outs = {}             # round -> caller-provided output tensor (AllReduceInput.out), else the ring slot's buffer
        self._cpu_worker = None
        self.stream_ordered = False  # local mode: the sink runs right after the launch, ordered on the CUDA stream

    # ------------------------------------------------------------------ lifecycle
    def connect(self, master_host: str = "127.0.0.1", master_port: int = 2551, listen_port: int = 0, heartbeat_ms: float = 1000.0):
        self.master_port = master_port
        self.node = core.WorkerNode(master_host, master_port, "127.0.0.1", listen_port, heartbeat_ms)
        if self.backend == "cpu":
            def src(it):
                d = self.data_source(AllReduceInputRequest(it)).data
                return np.ascontiguousarray(np.asarray(d, dtype=np.float32))

            def snk(data, count, it):
                self.delivered += 1
                self.data_sink(AllReduceOutput(data, count, it))

            self._cpu_worker = core.Worker(src, snk, self.node)
            self.node.attach(self._cpu_worker)
        elif self.native_loop:
            self._loop = core.DeviceLoop(self._loop_source, self._loop_sink, self.node)
            self._loop.set_init(self._loop_init)
        else:
            self.node.set_handler(self._on_message)
        return self

    def run(self, master_host: str = "127.0.0.1", master_port: int = 2551, timeout_sec: float = 3600.0) -> bool:
        """Join the master and serve until it says Shutdown (returns True) or the timeout expires."""
        if self.node is None:
            self.connect(master_host, master_port)
        if self.backend == "cpu":
            ok = self.node.run(timeout_sec)
            self.id = self._cpu_worker.id
            self.master_lost = ok and not self.node.shutdown_by_master
            return ok
        if self._loop is not None:
            self._loop_error = None
            try:
                ok = self._loop.run(self.node, timeout_sec)   # poll -> peer-ahead -> deliver, without the interpreter lock
            except RuntimeError:
                if self._loop_error is not None:   # a Python data plane failed inside a C thunk: surface its own exception
                    raise self._loop_error
                raise
            finally:
                self._sync_from_loop()
            self.master_lost = ok and not self.node.shutdown_by_master
            return ok
        deadline = time.time() + timeout_sec
        while not self.node.shutdown_requested and time.time() < deadline:
            busy = self.engine is not None and self.delivered < self.max_scattered + 1
            self.node.poll(0 if busy else 1)
            self._peer_ahead()
            self._deliver_flushed()
        self._drain()
        # True also when the master's link closed without a Shutdown message: `master_lost` tells the two apart
        self.master_lost = self.node.shutdown_requested and not self.node.shutdown_by_master
        return self.node.shutdown_requested

    def step(self, timeout_ms: int = 0) -> None:
        """One iteration of the event loop (for callers that own the loop)."""
        if self._loop is not None:
            self.node.set_handler(self._loop.receive)
            self.node.poll(timeout_ms)
            self._loop.peer_ahead()
            self._loop.deliver_flushed()
            self._sync_from_loop()
            return
        self.node.poll(timeout_ms)
        if self.backend == "cuda":
            self._peer_ahead()
            self._deliver_flushed()

    def _peer_ahead(self) -> None:
        """AllreduceWorker.scala:123-126: a ScatterBlock / ReduceBlock for a round beyond maxRound makes the worker start that
        round itself instead of waiting for the master's StartAllreduce.  On the device plane peers' blocks land in this
        worker's ring without any host involvement, so the trigger is the board: every worker posts the newest round it has
        started, and a worker that sees a peer ahead of its own maxRound fetches and launches up to that round."""
        if self.board is None or self.engine is None:
            return
        newest = self.board.newest()
        if newest > self.max_round:
            self.self_starts += newest - self.max_round
            self._on_start(newest)

    def metrics(self) -> dict:
        """rounds/s, forced completions, lag and the data-plane counters (CPU backend: the C++ worker's counters)."""
        if self.backend == "cpu":
            return dict(self._cpu_worker.metrics()) if self._cpu_worker else {}
        if self._loop is not None:
            self._sync_from_loop()
        m = {"id": self.id, "generation": self.generation, "rounds_delivered": self.delivered, "rounds_forced": self.rounds_forced,
             "max_round": self.max_round, "max_scattered": self.max_scattered, "lag": self.max_round - self.delivered + 1,
             "rounds_out_of_order": self.out_of_order, "self_started_rounds": self.self_starts}
        if self.engine is not None:
            m.update(self.engine.metrics())
        m["lag_hist"] = dict(sorted(self.lag_hist.items()))
        done = self.delivered - self.start_round
        if self._t_first is not None and done > 1:
            m["rounds_per_sec"] = rps = (done - 1) / max(1e-9, self._t_last - self._t_first)
            lay = getattr(self.engine, "layout", None)
            if lay is not None and getattr(lay, "P", 0) > 1:  # NCCL's bus-bandwidth convention: algbw * 2(P-1)/P
                m["bus_gbps"] = rps * lay.dataSize * lay.elemSize * 2 * (lay.P - 1) / lay.P / 1e9
        return m

    def close(self) -> None:
        if self._loop is not None:
            self._loop.clear_plane()   # the handle below is about to be destroyed
        if self.engine is not None:
            # Kernels of this incarnation may still be waiting for peers that are being re-initialised themselves (a late joiner
            # makes the master re-init everyone).  Their region, doorbell and completion words are about to be unmapped: ring the
            # doorbell far ahead so that every device-side wait gives up (the catch-up predicate), let them finish, then free.
            try:
                self.engine.set_doorbell(1 << 30)
                self.engine.synchronize()
            except Exception:   # teardown is best effort: a dead context must not mask the error that brought us here
                pass
            self.engine.close()
            self.engine = None
        if self.symm is not None:
            self.symm.close()
            self.symm = None
        if self.rdv is not None:
            self.rdv.close()
            self.rdv = None
        if self.board is not None:
            self.board.close(unlink=self.id == 0)
            self.board = None

    # ------------------------------------------------------------------ C++ round loop: what it calls back
    def _loop_init(self, m) -> None:
        """InitWorkers under the C++ loop: build the data plane as always, then hand the loop its handle."""
        self._loop.clear_plane()
        self._on_init(m)
        self._loop.set_drop_start(bool(self.drop_start))   # after the caller's on_init hook, which may decide it from the id
        eng = self.engine
        plane = eng.native_plane() if hasattr(eng, "native_plane") else None
        self._direct = plane is not None   # launches go C++ -> aar_engine_enqueue_ex; otherwise through the thunks below
        if plane is None:
            plane = self._python_plane(eng)
        self._loop.set_plane(plane, self.id, self.generation, m.master, self.start_round)
        if self.board is not None and len(m.workers) > 1:   # nobody to be ahead of in a world of one
            self._loop.set_board(self.board.post, self.board.newest)
        else:
            self._loop.clear_board()

    def _python_plane(self, eng):
        """A data plane that only exists in Python (the launch path with AAR_NATIVE_ENGINE=0, fake engines in tests), wrapped in C
        thunks so the same loop drives it."""
        import ctypes as C

        ENQ = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint, C.c_void_p, C.c_ulonglong)
        I2 = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)
        I1 = C.CFUNCTYPE(C.c_int, C.c_void_p)
        per_round = getattr(eng, "round_flushed", None)

        def enqueue(_e, rnd, _i, _o, _c, _f, _s, _d):
            x, out = self._inputs[rnd], self._outs.get(rnd)
            try:
                eng.enqueue_round(rnd, x, out=out) if out is not None else eng.enqueue_round(rnd, x)
            except Exception as exc:   # a C caller cannot unwind a Python exception: report it as a failed launch
                self._loop_error = exc
                return 1
            return 0

        def doorbell(_e, newest):
            eng.set_doorbell(newest)
            return 0

        scan = {"last": -1, "in_order": 0}

        def flushed(_e, rnd):
            if per_round is not None:
                return int(per_round(rnd))
            # a data plane that only reports an in-order count completes in order: one snapshot per ascending scan of the loop
            if scan["last"] < 0 or rnd <= scan["last"]:
                scan["in_order"] = eng.rounds_flushed()
            scan["last"] = rnd
            return int(rnd < scan["in_order"])

        def synchronize(_e):
            eng.synchronize()
            return 0

        self._plane_callbacks = (ENQ(enqueue), I2(doorbell), I2(flushed), I1(synchronize))
        a = [C.cast(f, C.c_void_p).value for f in self._plane_callbacks]
        return core.DevicePlane(engine=1, enqueue=a[0], doorbell=a[1], flushed=a[2], synchronize=a[3], slots=eng.layout.L,
                                max_lag=eng.max_lag, data_size=eng.data_size)

    def _loop_source(self, rnd: int):
        """source(round) for the C++ loop: the caller's data_source plus everything that needs the tensors (size check, output
        choice, alignment bounce); returns raw pointers."""
        eng = self.engine
        req = AllReduceInputRequest(rnd)
        if self.zero_copy_input and hasattr(eng, "stage_input"):
            req.buffer = eng.stage_input(rnd)
        inp = self.data_source(req)
        x = inp.data
        n = x.numel() if hasattr(x, "numel") else len(x)
        out = getattr(inp, "out", None)
        if n != eng.data_size:   # the loop raises the reference's message (AllreduceWorker.scala:174-176)
            return (0, 0, 0, 0, n)
        if self._direct:
            import torch

            x, output = eng.prepare_round(rnd, x, out)
            eng.note_launched(rnd)
            slot = rnd % eng.layout.L
            ptrs = (x.data_ptr(), output.data_ptr(), eng.counts[slot].data_ptr(), torch.cuda.current_stream(eng.device).cuda_stream, n)
        else:
            ptrs = (0, 0, 0, 0, n)
        self._inputs[rnd] = x
        if out is not None:
            self._outs[rnd] = out
        return ptrs

    def _loop_sink(self, rnd: int, slot: int) -> None:
        eng = self.engine
        self._inputs.pop(rnd, None)
        self.data_sink(AllReduceOutput(self._outs.pop(rnd, eng.outputs[slot]), eng.counts[slot], rnd))
        self._t_last = time.time()
        if self._t_first is None:
            self._t_first = self._t_last

    def _sync_from_loop(self) -> None:
        """Mirror the loop's state into the attributes metrics() and the tests read."""
        st = self._loop.stats()   # one consistent snapshot (metrics() may be called from another thread while the loop runs)
        self.round, self.max_round, self.max_scattered = st["round"], st["max_round"], st["max_scattered"]
        self.completed = set(st["completed"])
        self.delivered = self.start_round + st["rounds_delivered"]
        self.out_of_order, self.self_starts, self.rounds_forced = st["rounds_out_of_order"], st["self_started_rounds"], st["rounds_forced"]
        self.lag_hist = dict(st["lag_hist"])

    # ------------------------------------------------------------------ masterless mode
    @classmethod
    def local(cls, data_source: DataSource, data_sink: DataSink, engine, *, worker_id: int = 0, stream_ordered: bool = True) -> "AllreduceWorker":
        """A worker whose round clock is its caller instead of an AllreduceMaster: `start_allreduce(r)` is the StartAllreduce message,
        completions are counted locally.  For callers that are already in lock-step -- a data-parallel training step is one
        round per gradient bucket (parallel/ddp.py) -- and must not pay a TCP round trip per round.  Same contract otherwise:
        every round is fetched from the source exactly once, in order, and its result reaches the sink.  stream_ordered: the sink
        is called as soon as the round is launched, with tensors that are valid in CUDA stream order (the current stream has
        been made to wait for the round) -- the host never blocks; False: the sink is called when the host sees the round flushed."""
        w = cls(data_source, data_sink, backend="cuda")
        w.engine = engine
        w.id = worker_id
        w.node = _LocalClock()
        w.stream_ordered = stream_ordered
        w.max_round = w.max_scattered = -1
        return w

    def start_allreduce(self, rnd: int, *, deliver: bool = True) -> None:
        """Local mode: StartAllreduce(rnd) from the caller's clock.  deliver=False only launches: the caller's stream is NOT made
        to wait for the round here, so whatever it enqueues next (the rest of a backward pass) runs beside the round; the
        caller joins later with deliver()."""
        self._on_start(rnd)
        if deliver:
            self.deliver()

    def deliver(self) -> None:
        """Local mode: hand every launched round to the sink (stream-ordered: the current stream waits for each of them)."""
        if self.stream_ordered:
            self._deliver_stream_ordered()
        else:
            self._deliver_flushed()

    def _deliver_stream_ordered(self) -> None:
        eng = self.engine
        while self.delivered <= self.max_scattered:
            r = self.delivered
            slot = r % eng.layout.L
            eng.wait_round(r)           # the CURRENT stream waits for the round; nothing blocks on the host
            self._inputs.pop(r, None)
            self.data_sink(AllReduceOutput(self._outs.pop(r, eng.outputs[slot]), eng.counts[slot], r))
            self.node.send_complete(self.id, r, self.generation)
            self.delivered += 1
            self.round = self.delivered

    # ------------------------------------------------------------------ device data plane
    def _on_message(self, msg) -> None:
        if isinstance(msg, core.InitWorkers):
            self._on_init(msg)
        elif isinstance(msg, core.StartAllreduce):
            if not self.drop_start and msg.generation in (0, self.generation):
                self._on_start(msg.round)

    def _on_init(self, m) -> None:
        """AllreduceWorker.scala:37-82: (re)initialisation wipes all round state and allocates the rings -- here the
        symmetric HBM region, mapped into every peer."""
        self.close()
        self.id, self.generation = m.dest_id, m.generation
        # a resumed master continues its round numbering (InitWorkers.start_round; 0 in the reference): earlier rounds count as
        # retired everywhere, which the device-side flow control is told inside _make_engine, BEFORE the barrier that lets peers
        # launch (a peer's first retirement must not be overwritten by this rank's initialisation)
        start = getattr(m, "start_round", 0)
        self.engine = self._make_engine(m)
        if start > 0 and not getattr(self.engine, "_start_applied", False):   # data planes built elsewhere (tests)
            self.engine.start_at(start)
        self.max_round = self.max_scattered = start - 1
        self.delivered = self.round = start
        self.completed = set()
        self.out_of_order = self.self_starts = 0
        self.start_round = start
        self.lag_hist = {}
        self._t_first = None
        self._inputs.clear()
        self._outs.clear()
        if self.on_init:
            self.on_init(self)

    def _make_engine(self, m):
        """Build this rank's data plane for the geometry in InitWorkers: symmetric region mapped into every peer + round
        engine.  Kept separate from the bookkeeping so the host state machine can be tested against a fake data plane
        (tests/test_device_worker_logic.py), the way the reference tests one worker against a fake cluster."""
        import torch

        from .ops import lib as L
        from .parallel.engine import ONESHOT_MAX_BYTES, RoundEngine
        from .parallel.symmetric import SymmetricMemory, UnixRendezvous

        P = len(m.workers)
        dev = self.device if self.device is not None else self.id % torch.cuda.device_count()
        torch.cuda.set_device(dev)
        dt = L.AAR_F32 if self.dtype == "fp32" else L.AAR_BF16
        es = 4 if dt == L.AAR_F32 else 2
        layout = L.compute_layout(P, m.data_size, m.max_chunk_size, m.max_lag, dt, ONESHOT_MAX_BYTES // es, self.want_multicast)
        session = (self.session or f"w{self.master_port}") + f"_g{m.generation}"
        self.rdv = UnixRendezvous(self.id, P, session)
        self.symm = SymmetricMemory(self.rdv, layout.regionBytes, dev, want_multicast=self.want_multicast and P > 1)
        engine = RoundEngine.from_symmetric(self.symm, layout, th_reduce=m.th_reduce, th_complete=m.th_complete,
                                            normalize=self.normalize, exact_quorum=self.exact_quorum,
                                            timeout_ms=self.spin_timeout_ms, grid=self.grid, algo=self.algo)
        start = getattr(m, "start_round", 0)
        if start > 0:
            engine.start_at(start)
            engine._start_applied = True
        from .parallel.board import PeerBoard

        self.board = PeerBoard(session, P, self.id, start - 1)
        self.rdv.barrier()  # nobody launches before every peer has mapped every region (and set up its words)
        return engine

    def _on_start(self, rnd: int) -> None:
        """AllreduceWorker.scala:84-104.  The catch-up loop is the doorbell: kernels of rounds that are now more than
        maxLag behind stop waiting and complete with what has arrived.  Then every not-yet-scattered round is fetched and
        launched, in order, exactly once."""
        eng = self.engine
        if eng is None:
            return
        self.max_round = max(self.max_round, rnd)
        eng.set_doorbell(self.max_round)
        if self.board is not None:
            self.board.post(self.max_round)
        L = eng.layout.L
        while self.max_scattered < self.max_round:
            r = self.max_scattered + 1
            # the output / count buffers of round r - L must have reached the sink before round r reuses them
            while r - L >= self.round and (r - L) not in self.completed:
                if self.stream_ordered:
                    self._deliver_stream_ordered()
                else:
                    self._deliver_flushed(block=True)
            req = AllReduceInputRequest(r)
            if self.zero_copy_input and hasattr(eng, "stage_input"):
                req.buffer = eng.stage_input(r)   # None unless the in-switch path is active
            inp = self.data_source(req)
            x = inp.data
            n = x.numel() if hasattr(x, "numel") else len(x)
            if n != eng.data_size:
                raise ValueError(f"Input data size {n} is different from initialization time {eng.data_size}!")
            # The engine orders its launch after the current stream (where the source produced x) and, when the ring has
            # two slots, alternates rounds between two internal streams so consecutive rounds overlap (the maxLag window).
            out = getattr(inp, "out", None)
            if out is not None:
                eng.enqueue_round(r, x, out=out)
                self._outs[r] = out
            else:
                eng.enqueue_round(r, x)
            self._inputs[r] = x
            self.max_scattered = r

    def _deliver_flushed(self, block: bool = False) -> None:
        """AllreduceWorker.scala:253-268 (complete): sink, then CompleteAllreduce to the master, for every round the device has
        flushed -- in whatever order they flush.  `round` (the oldest round still in flight) advances past completed rounds."""
        eng = self.engine
        if eng is None:
            return
        per_round = getattr(eng, "round_flushed", None)   # data planes that only report an in-order count complete in order
        while True:
            progressed = False
            in_order = eng.rounds_flushed() if per_round is None else None
            for r in range(self.round, self.max_scattered + 1):
                if r < self.round or r in self.completed:   # self.round moves inside this loop
                    continue
                if not (per_round(r) if per_round is not None else r < in_order):
                    continue
                slot = r % eng.layout.L
                self._inputs.pop(r, None)
                lag = self.max_round - r
                self.lag_hist[lag] = self.lag_hist.get(lag, 0) + 1
                if lag > eng.max_lag:
                    self.rounds_forced += 1
                if r != self.round:
                    self.out_of_order += 1
                self.data_sink(AllReduceOutput(self._outs.pop(r, eng.outputs[slot]), eng.counts[slot], r))
                self.node.send_complete(self.id, r, self.generation)
                self.completed.add(r)
                self.delivered += 1
                self._t_last = time.time()
                if self._t_first is None:
                    self._t_first = self._t_last
                progressed = True
                while self.round in self.completed:   # AllreduceWorker.scala:261-267: past every already-completed round
                    self.completed.discard(self.round)
                    self.round += 1
            if not block or progressed or self.delivered >= self.max_scattered + 1:
                return
            time.sleep(0)  # blocking mode: yield until the next round flushes

    def _drain(self) -> None:
        """The job is over (Shutdown, lost master or timeout).  Rounds still in flight here can only be ones the master no longer
        waits for (thAllreduce < 1) or ones whose peers are gone: ring the doorbell far ahead so that their device-side waits give
        up instead of spinning on peers that have exited, then deliver what they flush."""
        if self.engine is not None:
            self.engine.set_doorbell(1 << 30)
            self.engine.synchronize()
            self._deliver_flushed()


class _LocalClock:
    """Stands in for the control channel of a masterless worker: completions are counted, nothing is sent."""

    shutdown_requested = False
    shutdown_by_master = False

    def __init__(self):
        self.completed: list[int] = []

    def send_complete(self, worker_id: int, rnd: int, generation: int) -> None:
        self.completed.append(rnd)

    def poll(self, timeout_ms: int = 0) -> None:
        pass
