"""Data-parallel gradient averaging through the worker source/sink contract, overlapped with backward.

dataSize is fixed when a worker is initialised (AllreduceWorker.scala:174-176), so the flat gradient is cut into equal
fixed-size buckets and every bucket gets its own AllreduceWorker ("one worker per bucket") in
masterless mode: the training step is the round clock (AllreduceWorker.local / start_allreduce), so no control message crosses a
socket per round, but the data still flows through the worker contract -- the worker calls the bucket's data_source for the
round's input and its data_sink with the result.  Parameters' `.grad` tensors are views into the bucket buffers, so autograd
accumulates straight into the memory the kernels read.  A post-accumulate-grad hook counts down each bucket; the moment a
bucket is complete its worker is started on the step's round: the source hands over the buffer (AllReduceInput.out = the same
buffer: in place), the engine launches the round on its own streams while autograd keeps running.  The sink sees the averaged
gradients (normalisation by the ACTUAL contributor count is fused into the reduce kernel), stream-ordered.

`zero_grad()` of THIS class must be used (it zeroes the buckets); `optimizer.zero_grad(set_to_none=True)` -- the torch default --
would detach the gradients from the buckets.  The hook notices a detached or re-bound `.grad` and copies it back in (correct,
one extra copy), and a second backward before finish() -- gradient accumulation -- defers the launch to finish().
"""
from __future__ import annotations

from typing import Callable

import torch

from ..api import AllReduceInput, AllReduceInputRequest, AllReduceOutput
from ..ops import lib as L
from .engine import RoundEngine
from .symmetric import SymmetricMemory, UnixRendezvous


def plan_buckets(sizes: list[int], per_bucket: int, align: int) -> list[tuple[list[int], int]]:
    """Greedy packing of parameters (given in the order their gradients become ready) into buckets of at most `per_bucket`
    elements -- a single larger parameter gets a bucket of its own -- each padded to a multiple of `align` elements.
    Returns [(indices into sizes, padded bucket length)].  Pure function: the fixed-dataSize contract of the worker
    (AllreduceWorker.scala:174-176) makes the plan part of the initialisation, so it is tested on its own."""
    groups: list[list[int]] = []
    cur: list[int] = []
    size = 0
    for i, n in enumerate(sizes):
        if size and size + n > per_bucket:
            groups.append(cur)
            cur, size = [], 0
        cur.append(i)
        size += n
    if cur:
        groups.append(cur)
    out = []
    for g in groups:
        n = sum(sizes[i] for i in g)
        out.append((g, (n + align - 1) // align * align))
    return out


class GradientBucket:
    def __init__(self, index: int, numel: int, dtype: torch.dtype, device: torch.device):
        self.index = index
        self.buffer = torch.zeros(numel, dtype=dtype, device=device)
        self.params: list[torch.nn.Parameter] = []
        self.pending = 0
        self.launched_step = -1     # last training step whose round was started
        self.engine: RoundEngine | None = None
        self.symm: SymmetricMemory | None = None
        self.worker = None          # AllreduceWorker in masterless mode
        self.views: dict[int, torch.Tensor] = {}   # id(param) -> its view of the buffer
        self.last_out: AllReduceOutput | None = None

    # the worker contract (DataWrapper.scala:3-7)
    def data_source(self, req: AllReduceInputRequest) -> AllReduceInput:
        return AllReduceInput(self.buffer, out=self.buffer)   # in place

    def data_sink(self, out: AllReduceOutput) -> None:
        self.last_out = out
        self.last_counts = out.count


class ThresholdDDP:
    def __init__(self, module: torch.nn.Module, rank: int, world: int, device: int, *, bucket_mb: float = 32.0, max_chunk: int = 1 << 20,
                 max_lag: int = 1, th_reduce: float = 1.0, th_complete: float = 1.0, session: str = "ddp", algo: str = "auto",
                 sink: Callable[[int, AllReduceOutput], None] | None = None, comm_ctas: int = 0, comm_split: tuple[int, int] | None = None,
                 engine_factory=None):
        """comm_ctas: SM budget of the gradient allreduce (CTAs per launch, one CTA occupies one SM); 0 = all SMs.  A bucket's
        round runs while backward is still producing the next bucket, so every SM it holds is taken from the GEMMs; the last
        bucket (ready when backward ends) always gets all SMs.  comm_split = (push CTAs, reduce CTAs) of a budgeted round on the
        bulk kernels, the rest gathers (the library's default split is tuned for a full grid).
        engine_factory(bucket_index, numel, chunk) -> engine: replaces the data plane (the hook / bucket bookkeeping is then
        testable without a GPU, tests/test_models_ddp_cpu.py); `device` may then be a torch.device such as cpu."""
        self.module, self.rank, self.world = module, rank, world
        self.device = device if isinstance(device, torch.device) else torch.device("cuda", device)
        self.step = 0
        self.rebinds = 0      # how often a hook found .grad detached from its bucket and copied it back in
        self.user_sink = sink
        params = [p for p in module.parameters() if p.requires_grad]
        dtype = params[0].dtype
        assert dtype in (torch.float32, torch.bfloat16) and all(p.dtype == dtype for p in params), "one gradient dtype (fp32 or bf16)"
        es = params[0].element_size()
        per_bucket = max(1, int(bucket_mb * (1 << 20)) // es)
        # Reverse registration order ~ the order gradients become ready; pack greedily, pad every bucket to a size the
        # 16-byte vector path likes (multiple of world * 64 elements).
        self.buckets: list[GradientBucket] = []
        ordered = list(reversed(params))
        code = L.AAR_F32 if dtype == torch.float32 else L.AAR_BF16
        plan = plan_buckets([p.numel() for p in ordered], per_bucket, world * 64)
        for i, (idx, n) in enumerate(plan):
            # the last bucket becomes ready when backward ends: nothing is left to share the SMs with, its round is the exposed tail
            grid = 0 if i == len(plan) - 1 else comm_ctas
            g = [ordered[k] for k in idx]
            b = GradientBucket(i, n, dtype, self.device)
            off = 0
            for p in g:
                p.grad = b.views[id(p)] = b.buffer[off:off + p.numel()].view_as(p)   # autograd accumulates in place from now on
                off += p.numel()
                b.params.append(p)
                p.register_post_accumulate_grad_hook(self._make_hook(b))
            chunk = max(1, min(max_chunk, (n + world - 1) // world))
            if engine_factory is not None:
                b.engine = engine_factory(i, n, chunk)
            elif world > 1:
                lay = L.compute_layout(world, n, chunk, max_lag, code, 0, True)
                b.symm = SymmetricMemory(UnixRendezvous(rank, world, f"{session}_b{i}"), lay.regionBytes, device, want_multicast=True)
                b.engine = RoundEngine.from_symmetric(b.symm, lay, th_reduce=th_reduce, th_complete=th_complete, normalize=True, algo=algo,
                                                      grid=grid)
            else:
                b.engine = RoundEngine.single_process(n, chunk, max_lag, dtype=dtype, device=device, normalize=True, grid=grid)
            if comm_split and grid and hasattr(b.engine, "bulk_split"):
                b.engine.bulk_split = int(comm_split[0]) | (int(comm_split[1]) << 16)
            from ..worker import AllreduceWorker

            b.worker = AllreduceWorker.local(b.data_source, b.data_sink, b.engine, worker_id=rank)
            self.buckets.append(b)
        self._arm()

    def _arm(self):
        for b in self.buckets:
            b.pending = len(b.params)

    def _make_hook(self, b: GradientBucket):
        def hook(param):
            view = b.views[id(param)]
            g = param.grad
            if g is not None and g.data_ptr() != view.data_ptr():
                # .grad was re-bound (optimizer.zero_grad(set_to_none=True), a hook, manual assignment): the bucket would
                # reduce stale memory while the optimizer steps on un-reduced gradients.  Copy in and re-attach.
                view.copy_(g)
                param.grad = view
                self.rebinds += 1
            b.pending -= 1
            if b.pending == 0 and b.launched_step < self.step:
                self._launch(b)
            # pending < 0: a second backward before finish() (gradient accumulation).  The round of this step is already in
            # flight on memory autograd is accumulating into again; finish() notices and refuses.
        return hook

    def _launch(self, b: GradientBucket):
        # in place: every element range is pushed (read) before the chunk that overwrites it can have been reduced
        b.launched_step = self.step
        # -> data_source -> kernels.  The sink is NOT reached here: delivering makes the current stream -- the one backward is
        # running on -- wait for the round, which would serialise every bucket's round with the rest of backward.  finish() joins.
        b.worker.start_allreduce(self.step, deliver=False)

    def finish(self) -> None:
        """Call after backward(): the current stream waits for every bucket's round, sinks have been invoked, hooks re-armed."""
        for b in self.buckets:
            if b.pending < 0:
                raise RuntimeError("ThresholdDDP: a second backward() ran before finish(): this step's gradient rounds were launched after "
                                   "the first one and are reducing memory that autograd has written again.  For gradient accumulation call "
                                   "backward() under ddp.no_sync() for all but the last micro-batch.")
            if b.launched_step < self.step:   # a parameter without gradient this step, or no_sync(): launch what we have
                self._launch(b)
        for b in self.buckets:
            b.worker.deliver()                # the current stream waits for the bucket's round; data_sink runs
            if self.user_sink and b.last_out is not None:
                self.user_sink(b.index, b.last_out)
        self.step += 1
        self._arm()

    def no_sync(self):
        """Context manager: backward() inside it accumulates into the buckets without starting their rounds (gradient
        accumulation); the last micro-batch runs outside it, or finish() launches whatever was deferred."""
        ddp = self

        class _NoSync:
            def __enter__(self_inner):
                for b in ddp.buckets:
                    b.pending = 1 << 30   # never reaches zero inside the block
                return ddp

            def __exit__(self_inner, *exc):
                ddp._arm()
                return False

        return _NoSync()

    def zero_grad(self) -> None:
        for b in self.buckets:
            b.buffer.zero_()

    def launches(self) -> int:
        return sum(b.engine.launches for b in self.buckets)

    def close(self) -> None:
        for b in self.buckets:
            b.engine.close()
            if b.symm:
                b.symm.close()
