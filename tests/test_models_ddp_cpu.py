"""CPU checks of the data-parallel client pieces: the GPT-2 small definition (BASELINE.json config 4) and the bucket plan.
The reference has no model code; these only pin what bench/gpt2_dp.py relies on."""
import torch

from akka_allreduce_b200.models.gpt2 import GPT2, GPT2Config
from akka_allreduce_b200.parallel.ddp import plan_buckets


def test_gpt2_small_has_124m_parameters():
    with torch.device("meta"):
        m = GPT2(GPT2Config())
    n = sum(p.numel() for p in m.parameters())
    # 12 layers x 7.09 M + (50304 + 1024) x 768 embeddings (tied head) + final LayerNorm
    assert n == 12 * (4 * 768 * 768 + 4 * 768 + 2 * 4 * 768 * 768 + 5 * 768 + 4 * 768) + (50304 + 1024) * 768 + 2 * 768
    assert 123e6 < n < 125e6


def test_gpt2_tiny_trains_on_cpu():
    torch.manual_seed(0)
    cfg = GPT2Config(vocab_size=64, n_layer=2, n_head=2, n_embd=32, block_size=16)
    m = GPT2(cfg)
    x = torch.randint(0, 64, (4, 16))
    y = torch.roll(x, -1, dims=1)
    opt = torch.optim.SGD(m.parameters(), lr=0.5)
    losses = []
    for _ in range(8):
        out = m(x, y)
        loss = out[1] if isinstance(out, tuple) else out
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(loss.item())
    assert abs(losses[0] - torch.log(torch.tensor(64.0)).item()) < 0.5     # random init: ~ln(vocab)
    assert losses[-1] < losses[0] - 0.2
    assert all(p.grad is not None for p in m.parameters())


def test_bucket_plan_is_greedy_padded_and_complete():
    sizes = [10, 20, 70, 5, 200, 1, 1]
    plan = plan_buckets(sizes, per_bucket=100, align=16)
    assert [g for g, _ in plan] == [[0, 1, 2], [3], [4], [5, 6]]          # 200 > per_bucket gets its own bucket
    assert [n for _, n in plan] == [112, 16, 208, 16]
    assert sorted(i for g, _ in plan for i in g) == list(range(len(sizes)))
    for g, n in plan:
        assert n % 16 == 0 and n >= sum(sizes[i] for i in g)
    assert plan_buckets([], 100, 16) == []


def test_bucket_plan_for_gpt2_small_matches_the_benchmark_shape():
    """bench/gpt2_dp.py: bf16, 32 MiB buckets, 4 ranks -> 8 buckets of at most 16 Mi elements (profiles/README.md quotes 7 of
    16.5 M with a larger bucket size); every bucket length is a multiple of world * 64 so the 16-byte vector path applies."""
    with torch.device("meta"):
        m = GPT2(GPT2Config())
    sizes = [p.numel() for p in reversed(list(m.parameters()))]
    plan = plan_buckets(sizes, per_bucket=(32 << 20) // 2, align=4 * 64)
    assert sum(n for _, n in plan) >= sum(sizes)
    assert all(n % 256 == 0 for _, n in plan)
    assert max(n for _, n in plan) <= 50304 * 768 + 256    # the embedding is the largest single tensor


class _FakeEngine:
    """Stands in for RoundEngine: 'averages' in place over `world` identical contributions and records what was asked of it."""

    def __init__(self, index, numel, chunk, log):
        from types import SimpleNamespace

        self.index, self.numel, self.chunk, self.log = index, numel, chunk, log
        self.launches = 0
        self.closed = False
        # what the masterless AllreduceWorker asks of its data plane
        self.data_size, self.max_lag, self.layout = numel, 1, SimpleNamespace(L=2)
        self.outputs = [None, None]
        self.counts = [torch.full((4,), 2, dtype=torch.int32) for _ in range(2)]
        self.doorbell = -1

    def set_doorbell(self, r):
        self.doorbell = max(self.doorbell, r)

    def enqueue_round(self, rnd, x, out=None):
        assert x.numel() == self.numel and out is x, "the bucket buffer is reduced in place"
        self.log.append(("enqueue", self.index, rnd))
        self.launches += 1
        return x, torch.full((4,), 2, dtype=torch.int32)

    def wait_round(self, rnd):
        self.log.append(("wait", self.index, rnd))

    def close(self):
        self.closed = True


def test_threshold_ddp_bookkeeping_with_a_fake_data_plane():
    """ThresholdDDP without a GPU (engine_factory): .grad tensors are views into the bucket buffers, a bucket's AllreduceWorker is
    started by the hook of its LAST gradient (so early buckets go out while backward is still running) and goes through the
    worker contract (data_source -> round -> data_sink, CompleteAllreduce counted), finish() launches what a parameter without
    gradient left behind, hands the counts to the sink and re-arms the hooks."""
    from akka_allreduce_b200.parallel.ddp import ThresholdDDP

    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16), torch.nn.ReLU(), torch.nn.Linear(16, 4))
    unused = torch.nn.Parameter(torch.zeros(5))          # never takes part in the loss: no gradient, no hook call
    model.register_parameter("unused", unused)
    log, engines, sunk = [], [], []

    def factory(i, n, chunk):
        engines.append(_FakeEngine(i, n, chunk, log))
        return engines[-1]

    ddp = ThresholdDDP(model, 0, 2, torch.device("cpu"), bucket_mb=300 * 4 / (1 << 20), engine_factory=factory,
                       sink=lambda idx, out: sunk.append((idx, out.iteration, int(out.count[0]))))
    assert len(ddp.buckets) >= 3 and all(b.buffer.numel() % (2 * 64) == 0 for b in ddp.buckets)
    for b in ddp.buckets:
        for p in b.params:
            assert p.grad.data_ptr() >= b.buffer.data_ptr() and p.grad.data_ptr() < b.buffer.data_ptr() + b.buffer.numel() * 4
    for step in range(2):
        ddp.zero_grad()
        log.clear()
        loss = model(torch.randn(3, 8)).square().mean()
        loss.backward()
        launched_in_backward = [e for e in log if e[0] == "enqueue"]
        assert launched_in_backward and all(r == step for _, _, r in launched_in_backward)
        # backward's stream must not have been made to wait for any round yet: that would serialise the round with the rest
        # of backward (measured on B200: the whole gradient allreduce time added to the step).  finish() is the join.
        assert not [e for e in log if e[0] == "wait"], "a hook made the stream wait for its bucket's round"
        ddp.finish()
        enq = [i for k, i, _ in log if k == "enqueue"]
        assert sorted(enq) == list(range(len(ddp.buckets))), "every bucket exactly once per step (the one holding `unused` from finish())"
        assert sorted(i for k, i, _ in log if k == "wait") == list(range(len(ddp.buckets))), "the stream waits for every bucket's round"
        assert all(b.worker.node.completed == list(range(step + 1)) for b in ddp.buckets), "each bucket's worker completed every round, in order"
        assert ddp.step == step + 1
    assert len(sunk) == 2 * len(ddp.buckets) and all(c == 2 for _, _, c in sunk)
    # gradients really live in the buckets: what autograd accumulated is what the engine was handed
    w = model[4].weight
    assert torch.count_nonzero(w.grad) > 0
    owner = next(b for b in ddp.buckets if any(p is w for p in b.params))
    assert torch.count_nonzero(owner.buffer) > 0
    assert ddp.launches() == 2 * len(ddp.buckets)
    ddp.close()
    assert all(e.closed for e in engines)


def test_threshold_ddp_survives_a_rebound_grad_and_refuses_a_second_backward():
    """optimizer.zero_grad(set_to_none=True) -- the torch default -- detaches .grad from the bucket: the hook copies the fresh gradient
    back into the bucket and re-attaches it (the round reduces the real gradient, not stale memory); a second backward() before
    finish() is refused; under no_sync() it accumulates and finish() launches once."""
    import pytest

    from akka_allreduce_b200.parallel.ddp import ThresholdDDP

    torch.manual_seed(1)
    model = torch.nn.Sequential(torch.nn.Linear(8, 8), torch.nn.Tanh(), torch.nn.Linear(8, 2))
    log = []
    ddp = ThresholdDDP(model, 0, 2, torch.device("cpu"), bucket_mb=1.0, engine_factory=lambda i, n, c: _FakeEngine(i, n, c, log))
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    opt.zero_grad(set_to_none=True)                 # detaches every .grad from its bucket
    assert all(p.grad is None for p in model.parameters())
    x = torch.randn(5, 8)
    model(x).square().mean().backward()
    ddp.finish()
    assert ddp.rebinds == len(list(model.parameters()))
    for b in ddp.buckets:
        for p in b.params:
            assert p.grad.data_ptr() == b.views[id(p)].data_ptr() and torch.count_nonzero(p.grad) > 0
        assert torch.count_nonzero(b.buffer) > 0, "the bucket holds the real gradient"
    ddp.zero_grad()
    model(x).square().mean().backward()
    model(x).square().mean().backward()             # second backward before finish()
    with pytest.raises(RuntimeError, match="second backward"):
        ddp.finish()
    ddp.close()
    # gradient accumulation done right: all but the last micro-batch under no_sync()
    model = torch.nn.Sequential(torch.nn.Linear(8, 8), torch.nn.Tanh(), torch.nn.Linear(8, 2))
    ddp = ThresholdDDP(model, 0, 2, torch.device("cpu"), bucket_mb=1.0, engine_factory=lambda i, n, c: _FakeEngine(i, n, c, log))
    ddp.zero_grad()
    with ddp.no_sync():
        model(x).square().mean().backward()
        assert ddp.launches() == 0, "no round inside no_sync()"
    g1 = model[0].weight.grad.clone()
    model(x).square().mean().backward()
    ddp.finish()
    assert ddp.launches() == len(ddp.buckets)
    assert torch.allclose(model[0].weight.grad, 2 * g1), "both micro-batches accumulated into the bucket before its round"
    ddp.close()
