"""Python CLI entry points (cli/master.py, cli/worker.py) on the CPU data plane, and transport edge cases over real sockets."""
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput, _core as core

ROOT = Path(__file__).resolve().parents[1]


def _free_port():
    import socket

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_python_cli_master_and_cpu_workers():
    port = _free_port()
    common = ["--set", f"allreduce.transport.port={port}", "--set", "allreduce.th-reduce=1.0", "--set", "allreduce.th-complete=1.0",
              "--set", "allreduce.max-round=9", "--timeout", "60"]
    m = subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.cli.master", str(port), "2", "10", "2", *common],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT)
    ws = [subprocess.Popen([sys.executable, "-m", "akka_allreduce_b200.cli.worker", "0", "10", "--backend", "cpu", "--quiet", *common],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT) for _ in range(2)]
    mo, _ = m.communicate(timeout=120)
    assert m.returncode == 0 and "finished=1" in mo, mo[-1500:]
    for w in ws:
        wo, _ = w.communicate(timeout=120)
        assert w.returncode == 0 and "rounds_flushed=10" in wo, wo[-1500:]


def test_late_joiner_over_tcp_restarts_everyone_at_round_zero():
    """thAllreduce = 0.5 of 2: the master starts with one worker; when the second joins everybody is re-initialised and
    the round clock restarts (AllreduceMaster.scala:40-47).  Stale traffic of the first incarnation is dropped by generation."""
    n = 8
    last = 20000
    master = AllreduceMaster(2, 0.5, 1.0, 1.0, max_lag=1, data_size=n, max_round=last, max_chunk_size=4)
    master.start(60)
    seen = [[], []]

    def make(i):
        return AllreduceWorker(lambda req: AllReduceInput(np.full(n, float(i + 1), dtype=np.float32)),
                               lambda out: seen[i].append((out.iteration, float(out.data[0]), int(out.count.min()))), backend="cpu")

    w0, w1 = make(0), make(1)
    t0 = threading.Thread(target=w0.run, args=("127.0.0.1", master.port, 60))
    t0.start()
    deadline = time.time() + 20
    while not seen[0] and time.time() < deadline:   # the first worker runs alone for a while: its own value, one contributor
        time.sleep(0.001)
    t1 = threading.Thread(target=w1.run, args=("127.0.0.1", master.port, 60))
    t1.start()
    t0.join(90)
    t1.join(90)
    assert master.join(30)
    assert master.metrics()["init_broadcasts"] == 2
    solo = [s for s in seen[0] if s[2] == 1]
    duo = [s for s in seen[0] if s[2] == 2]
    assert solo and all(v == 1.0 for _, v, _ in solo)
    assert duo and all(v == 3.0 for _, v, _ in duo) and duo[-1][0] == last
    assert seen[1] and all(v == 3.0 and c == 2 for _, v, c in seen[1])
    # the second incarnation started again at round 0
    assert any(it == 0 for it, v, c in duo)


def test_settings_carry_the_device_plane_knobs():
    s = core.AllreduceSettings.from_config(core.Config.parse_string(
        """allreduce {
          dtype = bf16
          algo = nvls
          normalize = on
          quorum-policy = exact
          spin-timeout = 250ms
          pipeline-depth = 3
        }"""))
    assert (s.dtype, s.algo, s.normalize, s.quorum_policy) == ("bf16", "nvls", True, "exact")
    assert s.spin_timeout_us == 250000 and s.master.pipeline_depth == 3
