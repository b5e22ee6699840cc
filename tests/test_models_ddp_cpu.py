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
