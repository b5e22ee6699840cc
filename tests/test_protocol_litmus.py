"""Device-protocol litmus under ThreadSanitizer (csrc/tests/protocol_litmus.cpp): the kernels' flag / decision / election /
flow-control logic re-stated on host threads with a NON-atomic ring payload, so any missing happens-before edge is a
reported data race.  Replaces the compute-sanitizer runs that are not possible on the build pool.

What is pinned:
  * thresholds = 1.0 -- the protocol exactly as the round-1 kernels implement it (no flow control on the broadcast) is
    race-free, bit-exact and deadlock-free, with one or two rounds in flight, for several ring depths;
  * thresholds < 1.0 -- race-free only when the reduce-broadcast obeys the same slot flow control as the scatter.  The
    round-1 kernels do not do that yet: an owner whose quorum excludes a peer lagging by a whole ring can overwrite the slot
    that peer is still gathering from (found by this model, see docs/PROTOCOL.md "Known limits").
  * degraded mode -- one rank slow in every round plus tiny wait budgets, so pushes are dropped, chunk and round decisions are
    forced over incomplete sets and the gather gives up on chunks: race-free, bit-exact over the decided sets, and every chunk
    flushed whole or as zeros, PROVIDED the broadcast adopts the scatter's per-destination decision (AAR_FLAG_BCAST_FC, as
    written in aar_kernels.cu) and the eager gather decides once per chunk (AAR_FLAG_EAGER_DEC; second hole found by this model,
    `eagerDecision` = 0 reproduces half-flushed chunks at thresholds = 1 when a wait gives up).  Both fixes are opt-in kernel
    variants until they have run on hardware (AAR_STRICT=1).
"""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def litmus(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("needs g++")
    exe = tmp_path_factory.mktemp("litmus") / "protocol_litmus"
    b = subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-fno-omit-frame-pointer",
                        str(ROOT / "csrc/tests/protocol_litmus.cpp"), "-o", str(exe), "-lpthread"], capture_output=True, text=True, timeout=300)
    if b.returncode != 0:
        pytest.skip("ThreadSanitizer build not available: " + b.stderr[-300:])
    return exe


def run(exe, *args):
    r = subprocess.run([str(exe), *map(str, args)], capture_output=True, text=True, timeout=280, env={"TSAN_OPTIONS": "exitcode=66"})
    if "unexpected memory mapping" in r.stderr:
        pytest.skip("ThreadSanitizer cannot run in this container")
    return r


# seed, P, rounds, threshold %, maxLag, overlap, CTAs per kernel
FULL = [(1, 4, 10, 100, 1, 1, 3), (2, 4, 10, 100, 1, 0, 3), (4, 3, 8, 100, 2, 1, 2), (6, 2, 12, 100, 0, 0, 3), (9, 5, 6, 100, 3, 1, 2)]


@pytest.mark.parametrize("cfg", FULL)
def test_round1_protocol_is_race_free_at_full_thresholds(litmus, cfg):
    r = run(litmus, *cfg, 0, 0, 0)   # no flow control on the broadcast, no lagger, no eager decision word: the shipped kernels
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]


@pytest.mark.parametrize("cfg", [(3, 4, 12, 75, 3, 1, 2), (5, 4, 12, 75, 1, 1, 2), (7, 6, 6, 50, 2, 1, 2)])
def test_partial_thresholds_are_race_free_with_broadcast_flow_control(litmus, cfg):
    r = run(litmus, *cfg, 1)
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]


# seed, P, rounds, threshold %, maxLag, overlap, CTAs per kernel, bcastFlowControl, lagger, eagerDecision, spinBudget
DEGRADED = [(3, 4, 14, 75, 1, 1, 4, 1, 1, 1, 64), (4, 4, 14, 100, 1, 1, 4, 1, 1, 1, 64), (6, 3, 16, 60, 2, 1, 3, 1, 1, 1, 128),
            (12, 4, 14, 75, 1, 0, 4, 1, 1, 1, 32),
            # ... and with late-work elision (AAR_FLAG_ELIDE_LATE): a rank a whole round late skips what nobody will read
            (13, 5, 12, 80, 1, 1, 4, 1, 1, 1, 64, 1), (6, 3, 16, 60, 2, 1, 3, 1, 1, 1, 128, 1)]


@pytest.mark.parametrize("cfg", DEGRADED)
def test_degraded_mode_is_race_free_with_both_fixes(litmus, cfg):
    r = run(litmus, *cfg)
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]
    # the run really was degraded: pushes were dropped and decisions were forced
    import re

    m = re.search(r"drops=(\d+) forcedChunks=(\d+) forcedRounds=(\d+) abandoned=(\d+)", r.stdout)
    assert m and int(m.group(1)) > 0 and int(m.group(2)) > 0, r.stdout


def test_late_work_elision_is_race_free_and_exact(litmus):
    """No wait budgets here: the lagger is tolerated purely by the thresholds, falls a round behind, and elides pushes and
    broadcasts (the counters prove it); every flushed vector still equals the sum over the decided sets."""
    import re

    r = run(litmus, 5, 4, 14, 75, 3, 1, 4, 1, 1, 1, 1 << 30, 1)
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]
    m = re.search(r"elidedPushes=(\d+) elidedBcasts=(\d+)", r.stdout)
    assert m and int(m.group(1)) > 0 and int(m.group(2)) > 0, r.stdout


# seed, P, rounds, threshold %, maxLag, overlap, CTAs, fc, lagger, eagerDecision, spinBudget, elide, nvls
NVLS_SHIPPED = [(1, 4, 12, 100, 1, 1, 4, 0, 0, 0, 1 << 30, 0, 1), (2, 4, 12, 100, 1, 0, 4, 0, 0, 0, 1 << 30, 0, 1),
                (3, 3, 12, 100, 2, 1, 3, 0, 0, 0, 1 << 30, 0, 1), (4, 5, 8, 100, 3, 1, 4, 0, 0, 0, 1 << 30, 0, 1)]
NVLS_STRICT_DEGRADED = [(5, 4, 12, 100, 1, 1, 4, 0, 1, 1, 64, 0, 2), (6, 3, 14, 100, 2, 1, 3, 0, 1, 1, 128, 0, 2), (8, 4, 12, 100, 1, 0, 4, 0, 1, 1, 32, 0, 2)]


@pytest.mark.parametrize("cfg", NVLS_SHIPPED)
def test_nvls_path_as_shipped_is_race_free_at_full_thresholds(litmus, cfg):
    """The in-switch path of the headline number: staging slots WITHOUT flow control, the switch sums every rank's slot, multicast
    store into every reduce ring, eager gather, two rounds in flight -- race-free and exact as long as no wait gives up."""
    r = run(litmus, *cfg)
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]


@pytest.mark.parametrize("cfg", NVLS_STRICT_DEGRADED)
def test_strict_nvls_survives_degraded_mode(litmus, cfg):
    """With a persistent lagger and waits that give up, the shipped NVLS path races (staging slot rewritten under a late reduce)
    and publishes sums that do not match their counts; the strict variant -- staging obeys the slot flow control, a decision
    forced over an incomplete set is not carried out, the eager gather decides once -- stays race-free and exact."""
    r = run(litmus, *cfg)
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]


# seed, P, rounds, threshold %, maxLag, overlap, CTAs, fc, lagger, eagerDecision, spinBudget, elide, nvls, oneshot
ONESHOT = [(1, 4, 12, 100, 1, 1, 4, 0, 0, 0, 1 << 30, 0, 0, 1), (2, 4, 12, 75, 1, 1, 4, 0, 0, 0, 1 << 30, 0, 0, 1),
           (4, 5, 10, 80, 1, 0, 4, 0, 1, 0, 64, 0, 0, 1), (5, 4, 12, 100, 3, 1, 4, 0, 1, 0, 64, 0, 0, 1)]


@pytest.mark.parametrize("cfg", ONESHOT)
def test_oneshot_path_as_shipped_is_race_free_also_in_degraded_mode(litmus, cfg):
    """K5 (small messages): whole-vector pushes under the per-destination flow control, one arrival flag per source, one
    contributor decision per round, local sum -- clean at full and partial thresholds, with and without a lagger and give-ups."""
    r = run(litmus, *cfg)
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]


# seed, P, rounds, threshold %, maxLag, overlap, CTAs, fc, lagger, eagerDecision, spinBudget, elide, nvls (3 = one hop)
ONE_HOP = [(1, 4, 14, 100, 1, 1, 4, 0, 0, 0, 1 << 30, 0, 3), (2, 2, 16, 100, 1, 1, 4, 0, 0, 0, 1 << 30, 0, 3),
           (3, 3, 12, 100, 2, 1, 5, 0, 0, 0, 1 << 30, 0, 3), (4, 5, 10, 100, 1, 0, 4, 0, 0, 0, 1 << 30, 0, 3),
           (5, 4, 12, 100, 3, 1, 6, 0, 0, 0, 1 << 30, 0, 3)]


@pytest.mark.parametrize("cfg", ONE_HOP)
def test_one_hop_in_switch_round_is_race_free_in_normal_mode(litmus, cfg):
    """nvlsOneHopKernel: every rank reads every rank's staging slot, so a slot may only be restaged once ALL ranks retired its
    previous tenant; one flag per source after the last staged tile; whole-vector sum into the local output.  Two rounds in
    flight, rotating stragglers.  (A forced round is outside the model: the kernel's give-up is decided per CTA.)"""
    r = run(litmus, *cfg)
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "failures=0" in r.stdout, r.stdout + r.stderr[-1000:]
