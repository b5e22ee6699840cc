"""Property tests of the host protocol engine over random geometries and schedules (hypothesis).

The reference pins behaviour with 15 hand-written scenarios (AllreduceSpec.scala); these check the invariants behind them
for arbitrary P / dataSize / maxChunkSize / thresholds / maxLag and for arbitrary straggler schedules."""
import numpy as np
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from akka_allreduce_b200 import _core as core


def run_cluster(P, n, chunk, th_all, th_red, th_com, max_lag, rounds, holds):
    cfg = core.MasterConfig(total_workers=P, th_allreduce=th_all, th_reduce=th_red, th_complete=th_com, max_lag=max_lag, data_size=n,
                            max_round=rounds - 1, max_chunk_size=chunk)
    sinks = [[] for _ in range(P)]
    fetched = [[] for _ in range(P)]

    def src(i):
        def f(it):
            fetched[i].append(it)
            return (np.arange(n, dtype=np.float32) + 1) * (i + 1) + it
        return f

    cl = core.LocalCluster(cfg, P, src, lambda i: (lambda d, c, it: sinks[i].append((it, np.array(d), np.array(c)))))
    cl.join_all()
    for w, budget in holds:
        cl.hold(w)
        cl.run(budget)
        cl.release(w)
    cl.run()
    return cl, sinks, fetched


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(P=st.integers(1, 9), n=st.integers(1, 200), chunk=st.integers(1, 64), max_lag=st.integers(0, 4), rounds=st.integers(1, 6))
def test_full_thresholds_always_give_the_exact_sum(P, n, chunk, max_lag, rounds):
    cl, sinks, fetched = run_cluster(P, n, chunk, 1.0, 1.0, 1.0, max_lag, rounds, [])
    assert cl.master().finished
    want0 = (np.arange(n, dtype=np.float32) + 1) * sum(range(1, P + 1))
    geo = core.Geometry(P, n, chunk)
    for w in range(P):
        assert [it for it, _, _ in sinks[w]] == list(range(rounds))
        assert fetched[w] == list(range(rounds))                      # every round fetched exactly once, in order
        for it, data, count in sinks[w]:
            np.testing.assert_array_equal(data, want0 + P * it)
            assert len(count) == geo.total_chunks() and (count == P).all()


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(P=st.integers(2, 8), n=st.integers(1, 120), chunk=st.integers(1, 32), max_lag=st.integers(0, 3), rounds=st.integers(1, 8),
       th=st.sampled_from([0.5, 0.6, 0.75, 0.9]), data=st.data())
def test_thresholds_and_stragglers_never_break_the_invariants(P, n, chunk, max_lag, rounds, th, data):
    holds = data.draw(st.lists(st.tuples(st.integers(0, P - 1), st.integers(0, 400)), max_size=4))
    # thAllreduce stays 1.0: below that a late joiner re-initialises everyone from round 0 (AllreduceMaster.scala:40-47), which
    # legitimately repeats fetches
    cl, sinks, fetched = run_cluster(P, n, chunk, 1.0, th, th, max_lag, rounds, holds)
    geo = core.Geometry(P, n, chunk)
    min_req = core.Geometry.min_required(th, P)
    for w in range(P):
        its = [it for it, _, _ in sinks[w]]
        assert len(its) == len(set(its)), "a round was flushed twice"
        assert fetched[w] == list(range(len(fetched[w]))), "rounds must be fetched in order, each once"
        m = cl.worker(w).metrics()
        for it, d, count in sinks[w]:
            assert len(count) == geo.total_chunks() and ((count >= 0) & (count <= P)).all()
            # a chunk that was reduced normally has at least minRequired contributors; forced rounds may have fewer
            if m["rounds_forced"] == 0:
                assert ((count == 0) | (count >= min_req)).all()
            # every flushed element is a sum of `count` distinct workers' values: bounded by the extreme subsets
            ci = 0
            for p in range(P):
                for c in range(geo.num_chunks(p)):
                    s = geo.block_start(p) + c * chunk
                    e = s + geo.chunk_len(p, c)
                    k = int(count[ci])
                    base = np.arange(s, e, dtype=np.float64) + 1
                    lo = base * sum(range(1, k + 1)) + k * it
                    hi = base * sum(range(P - k + 1, P + 1)) + k * it
                    assert (d[s:e] >= lo - 1e-3).all() and (d[s:e] <= hi + 1e-3).all()
                    ci += 1
