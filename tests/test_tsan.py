"""ThreadSanitizer build of the coordinator + TCP transport + CPU workers (race detection)."""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
SRCS = ["worker.cpp", "master.cpp", "config.cpp", "checkpoint.cpp", "local_cluster.cpp", "tcp_transport.cpp"]


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_host_runtime_is_race_free_under_tsan(tmp_path):
    exe = tmp_path / "tsan_cluster"
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-fno-omit-frame-pointer", f"-I{ROOT / 'include'}",
           str(ROOT / "csrc/tests/tsan_cluster.cpp"), *[str(ROOT / "csrc/core" / s) for s in SRCS], "-o", str(exe), "-lpthread"]
    b = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if b.returncode != 0 and "tsan" in (b.stderr or "").lower():
        pytest.skip("libtsan is not installed in this image")
    assert b.returncode == 0, b.stderr[-2000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300, env={"TSAN_OPTIONS": "halt_on_error=1 exitcode=66"})
    if "FATAL: ThreadSanitizer: unexpected memory mapping" in r.stderr:
        pytest.skip("ThreadSanitizer cannot run in this container (address-space layout)")
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "finished=1 flushed=180 bad=0" in r.stdout, (r.stdout + r.stderr)[-2000:]


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_device_round_loop_is_race_free_under_tsan(tmp_path):
    """The C++ round loop of the device plane (csrc/core/device_loop.cpp) next to threads that play the GPU: completion words and
    the peer board are written by other threads while the loop polls them."""
    exe = tmp_path / "tsan_device_loop"
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-fno-omit-frame-pointer", f"-I{ROOT / 'include'}",
           str(ROOT / "csrc/tests/tsan_device_loop.cpp"), *[str(ROOT / "csrc/core" / s) for s in SRCS + ["device_loop.cpp"]], "-o", str(exe), "-lpthread"]
    b = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if b.returncode != 0 and "tsan" in (b.stderr or "").lower():
        pytest.skip("libtsan is not installed in this image")
    assert b.returncode == 0, b.stderr[-2000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300, env={"TSAN_OPTIONS": "halt_on_error=1 exitcode=66"})
    if "FATAL: ThreadSanitizer: unexpected memory mapping" in r.stderr:
        pytest.skip("ThreadSanitizer cannot run in this container (address-space layout)")
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-3000:]
    assert r.returncode == 0 and "finished=1 sunk=240 duplicates=0 bad=0" in r.stdout, (r.stdout + r.stderr)[-2000:]
