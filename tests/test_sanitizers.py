"""AddressSanitizer + UndefinedBehaviorSanitizer runs of the host runtime (race detection / sanitizers).

  * the coordinator + TCP transport + CPU workers cluster that tests/test_tsan.py runs under ThreadSanitizer;
  * csrc/tests/fuzz_protocol.cpp: random, partly hostile message sequences against one worker and one master (ids, chunk numbers
    and rounds outside every range, empty payloads, re-initialisation in the middle of a round).  Contract: a bad message is
    rejected with an exception or ignored, never undefined behaviour.  (First run found memcpy from a null pointer for an empty
    payload and an unbounded catch-up on an absurd round number.)
"""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
FLAGS = ["-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer", f"-I{ROOT / 'include'}"]
CORE = ["worker.cpp", "master.cpp", "config.cpp", "checkpoint.cpp", "local_cluster.cpp", "tcp_transport.cpp"]


TARGETS = {
    "fuzz_protocol": ("fuzz_protocol.cpp", ["worker.cpp", "master.cpp"]),
    "fuzz_wire": ("fuzz_wire.cpp", CORE),
    "fuzz_checkpoint": ("fuzz_checkpoint.cpp", ["worker.cpp", "master.cpp", "checkpoint.cpp"]),
    "fuzz_cluster": ("fuzz_cluster.cpp", ["worker.cpp", "master.cpp", "local_cluster.cpp"]),
    "asan_cluster": ("tsan_cluster.cpp", CORE),
    "asan_device_loop": ("tsan_device_loop.cpp", CORE + ["device_loop.cpp"]),
}


@pytest.fixture(scope="module")
def exes(tmp_path_factory):
    """All sanitizer binaries, compiled side by side (the builds, not the runs, dominate the time of this file)."""
    from concurrent.futures import ThreadPoolExecutor

    if shutil.which("g++") is None:
        pytest.skip("needs g++")
    out = tmp_path_factory.mktemp("sanitizers")

    def build(item):
        name, (main, srcs) = item
        exe = out / name
        b = subprocess.run(["g++", *FLAGS, str(ROOT / "csrc/tests" / main), *[str(ROOT / "csrc/core" / s) for s in srcs], "-o", str(exe), "-lpthread"],
                           capture_output=True, text=True, timeout=900)
        return name, exe, b

    with ThreadPoolExecutor(max_workers=len(TARGETS)) as ex:
        built = list(ex.map(build, TARGETS.items()))
    res = {}
    for name, exe, b in built:
        if b.returncode != 0 and ("asan" in b.stderr.lower() or "ubsan" in b.stderr.lower()):
            pytest.skip("sanitizer runtimes are not installed in this image")
        assert b.returncode == 0, f"{name}: {b.stderr[-2000:]}"
        res[name] = exe
    return res


def test_protocol_engine_survives_hostile_messages_under_asan_ubsan(exes):
    exe = exes["fuzz_protocol"]
    for seed in (1, 2):
        r = subprocess.run([str(exe), str(seed), "500"], capture_output=True, text=True, timeout=280)
        assert "runtime error" not in r.stderr and "AddressSanitizer" not in r.stderr, r.stderr[-3000:]
        assert r.returncode == 0 and "accepted=" in r.stdout, (r.stdout + r.stderr)[-2000:]


def test_transport_survives_garbage_on_live_sockets_under_asan_ubsan(exes):
    """csrc/tests/fuzz_wire.cpp: raw clients throw random frames at a MasterNode, a fake master does the same to WorkerNodes."""
    exe = exes["fuzz_wire"]
    r = subprocess.run([str(exe), "1", "200"], capture_output=True, text=True, timeout=280)
    assert "runtime error" not in r.stderr and "AddressSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "0 threw" in r.stdout, (r.stdout + r.stderr)[-2000:]


def test_checkpoint_loaders_refuse_damaged_files_under_asan_ubsan(exes, tmp_path):
    """csrc/tests/fuzz_checkpoint.cpp: valid worker / master / ring files flipped, truncated and spliced at random, then loaded (and the
    restored worker is driven one step).  First run found a restored round state that sent the worker into a catch-up of billions
    of rounds; the loader now checks the state machine's invariants."""
    exe = exes["fuzz_checkpoint"]
    for seed in (1, 3):
        r = subprocess.run([str(exe), str(seed), "800", str(tmp_path)], capture_output=True, text=True, timeout=280)
        assert "runtime error" not in r.stderr and "AddressSanitizer" not in r.stderr, r.stderr[-3000:]
        assert r.returncode == 0 and "refused=" in r.stdout, (r.stdout + r.stderr)[-2000:]


def test_random_clusters_keep_the_invariants_under_asan_ubsan(exes):
    """csrc/tests/fuzz_cluster.cpp: whole in-process clusters with random geometry (uneven and empty blocks, dataSize 0), thresholds,
    ring and pipeline depth, message loss, parked workers and late joiners.  Every flushed element must be count x input, nothing
    is flushed twice, and a fault-free run at thresholds 1.0 flushes every round on every worker with count = P.  (First run found
    a pipeline depth above maxLag forcing partial sums in a fault-free run -- the master now caps it -- and dataSize = 0 never
    completing.)"""
    exe = exes["fuzz_cluster"]
    for seed in (1, 2):
        r = subprocess.run([str(exe), str(seed), "250"], capture_output=True, text=True, timeout=280)
        assert "runtime error" not in r.stderr and "AddressSanitizer" not in r.stderr, r.stderr[-3000:]
        assert r.returncode == 0 and "violations=0" in r.stdout, (r.stdout + r.stderr)[-2000:]


def test_tcp_cluster_is_clean_under_asan_ubsan(exes):
    exe = exes["asan_cluster"]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=280)
    assert "runtime error" not in r.stderr and "AddressSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "finished=1 flushed=180 bad=0" in r.stdout, (r.stdout + r.stderr)[-2000:]


def test_device_round_loop_is_clean_under_asan_ubsan(exes):
    """csrc/core/device_loop.cpp end to end in C++ (master + 3 workers over TCP, fake device plane with its own threads,
    out-of-order flushes, one rank that never hears StartAllreduce)."""
    exe = exes["asan_device_loop"]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=280)
    assert "runtime error" not in r.stderr and "AddressSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "finished=1 sunk=240 duplicates=0 bad=0" in r.stdout, (r.stdout + r.stderr)[-2000:]
