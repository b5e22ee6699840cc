"""The parts of the bench.py / __graft_entry__ contract that can be checked without a GPU."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_reference_arm_says_unavailable_in_one_json_line_and_exits_zero():
    """The reference is a Scala/sbt project (no JVM on the box): `--impl reference` must say so and exit 0,
    under a launcher as well (RANK / WORLD_SIZE set: only rank 0 prints)."""
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "3", "--warmup", "3"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-1000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and "unavailable" in d and "\n" not in d["unavailable"]
    r1 = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2"], capture_output=True, text=True, timeout=120,
                        cwd=ROOT, env={**__import__("os").environ, "RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r1.returncode == 0 and not [ln for ln in r1.stdout.splitlines() if ln.startswith("{")]


def test_graft_entry_has_build_and_smoke():
    sys.path.insert(0, str(ROOT))
    import __graft_entry__ as g

    assert callable(g.build) and callable(g.smoke)


def test_clock_sampler_degrades_without_a_gpu():
    import importlib.util

    spec = importlib.util.spec_from_file_location("bench_mod_for_test", ROOT / "bench.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    s = m.ClockSampler(0)
    s.sample(2)
    out = s.stop()
    assert set(out) >= {"sm_mhz", "sm_max_mhz", "reasons"} and out["reasons"] == []
