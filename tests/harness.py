"""Fake-cluster harness: one real worker, every peer and the master impersonated by the test.

Same idea as the reference's TestKit set-up (AllreduceSpec.scala:746-763): all peer ids map to the
test's own address, so everything the worker sends is recorded and inbound traffic is scripted.
"""
from collections import deque

import numpy as np

from akka_allreduce_b200 import _core as core

TEST = 100   # address of the test actor (all fake peers + master)
REAL = 7     # address used when the worker should receive its own sends


def basic_source(size):
    return lambda it: np.arange(size, dtype=np.float32) + np.float32(it)


def custom_source(size, fn):
    return lambda it: np.array([fn(i, it) for i in range(size)], dtype=np.float32)


class Harness:
    def __init__(self, source, sink=None):
        self.outbox = deque()
        self.self_inbox = deque()
        self.flushed = []
        self.port = core.CallbackPort(self._send)
        self._user_sink = sink
        self.worker = core.Worker(source, self._sink, self.port)

    def _sink(self, data, count, iteration):
        self.flushed.append((iteration, np.array(data), np.array(count)))
        if self._user_sink:
            self._user_sink(data, count, iteration)

    def _send(self, to, msg):
        (self.self_inbox if to == REAL else self.outbox).append(msg)

    def init(self, num, idx, th_reduce, th_complete, max_lag, data_size, max_chunk, real_self=False):
        workers = {i: TEST for i in range(num)}
        if real_self:
            workers[idx] = REAL
        self.tell(core.InitWorkers(workers, TEST, idx, th_reduce, th_complete, max_lag, data_size, max_chunk))

    def tell(self, msg):
        self.worker.receive(msg)
        while self.self_inbox:
            self.worker.receive(self.self_inbox.popleft())

    def start(self, r):
        self.tell(core.StartAllreduce(r))

    def scatter(self, value, src, dest, chunk, rnd):
        self.tell(core.ScatterBlock([float(v) for v in value], src, dest, chunk, rnd))

    def reduce(self, value, src, dest, chunk, rnd, count):
        self.tell(core.ReduceBlock([float(v) for v in value], src, dest, chunk, rnd, count))

    # --- expectations -------------------------------------------------------------------------
    def expect_scatter(self, value, src, dest, chunk, rnd):
        assert self.outbox, "expected a ScatterBlock, outbox is empty"
        m = self.outbox.popleft()
        assert isinstance(m, core.ScatterBlock), f"expected ScatterBlock, got {type(m).__name__}"
        assert (m.src_id, m.dest_id, m.round, m.chunk_id) == (src, dest, rnd, chunk)
        assert list(m.value) == [float(v) for v in value]

    def expect_reduce(self, value, src, dest, chunk, rnd, count):
        assert self.outbox, "expected a ReduceBlock, outbox is empty"
        m = self.outbox.popleft()
        assert isinstance(m, core.ReduceBlock), f"expected ReduceBlock, got {type(m).__name__}"
        assert (m.src_id, m.dest_id, m.round, m.chunk_id, m.count) == (src, dest, rnd, chunk, count)
        assert list(m.value) == [float(v) for v in value]

    def expect_complete(self, src, rnd):
        assert self.outbox, "expected CompleteAllreduce, outbox is empty"
        m = self.outbox.popleft()
        assert isinstance(m, core.CompleteAllreduce), f"expected CompleteAllreduce, got {type(m).__name__}"
        assert (m.src_id, m.round) == (src, rnd)

    def fish_complete(self, src, rnd):
        while self.outbox:
            m = self.outbox.popleft()
            if isinstance(m, core.CompleteAllreduce) and (m.src_id, m.round) == (src, rnd):
                return
        raise AssertionError(f"CompleteAllreduce({src}, {rnd}) never sent")

    def expect_no_msg(self):
        assert not self.outbox, f"unexpected message {type(self.outbox[0]).__name__}"
