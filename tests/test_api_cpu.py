"""Public Python API on the CPU data plane (AllreduceMaster + AllreduceWorker(backend="cpu")) and the fd rendezvous."""
import multiprocessing as mp
import os
import threading

import numpy as np
import pytest

from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput, AllReduceInputRequest, AllReduceOutput, load_settings


def test_master_and_cpu_workers_through_the_public_api():
    P, n, rounds = 3, 12, 6
    master = AllreduceMaster(P, 1.0, 1.0, 1.0, max_lag=1, data_size=n, max_round=rounds - 1, max_chunk_size=3)
    master.start(60)
    got = [[] for _ in range(P)]

    def make(i):
        def source(req: AllReduceInputRequest) -> AllReduceInput:
            return AllReduceInput(np.arange(n, dtype=np.float32) * (i + 1) + req.iteration)

        def sink(out: AllReduceOutput) -> None:
            got[i].append((out.iteration, np.array(out.data), np.array(out.count)))

        return AllreduceWorker(source, sink, backend="cpu")

    workers = [make(i) for i in range(P)]
    threads = [threading.Thread(target=w.run, args=("127.0.0.1", master.port, 60)) for w in workers]
    for t in threads:
        t.start()
    for t in threads:
        t.join(90)
    assert master.join(30)
    ids = sorted(w.id for w in workers)
    assert ids == [0, 1, 2]
    # sources differ per worker: sum over workers of arange * (k+1) + it, whatever id each thread was given
    for i in range(P):
        assert [it for it, _, _ in got[i]] == list(range(rounds))
        for it, data, count in got[i]:
            np.testing.assert_array_equal(data, np.arange(n, dtype=np.float32) * 6 + 3 * it)
            assert (count == P).all()
    assert master.metrics()["rounds_started"] == rounds


def test_load_settings_layers_overrides_over_the_shipped_file():
    s = load_settings(overrides={"allreduce.total-workers": 8, "allreduce.data-size": 16777216, "allreduce.max-chunk-size": 1048576,
                                 "allreduce.th-reduce": 1.0, "allreduce.th-complete": 1.0, "allreduce.dtype": "bf16"})
    assert (s.master.total_workers, s.master.data_size, s.master.max_chunk_size) == (8, 16777216, 1048576)
    assert s.master.max_lag == 1 and s.dtype == "bf16" and s.port == 2551


def _rdv_proc(rank, world, session, q):
    from akka_allreduce_b200.parallel.symmetric import UnixRendezvous

    rdv = UnixRendezvous(rank, world, session, timeout=30)
    r, w = os.pipe()
    os.write(w, f"hello from {rank}".encode())
    os.close(w)
    fds = rdv.all_gather_fds(r)
    texts = []
    for k, fd in enumerate(fds):
        if k == rank:
            texts.append(None)
            continue
        texts.append(None)  # a pipe can be read once; only rank (k+1) % world reads rank k's pipe
    reader_of = (rank - 1) % world
    msg = os.read(fds[reader_of], 100).decode()
    blobs = rdv.all_gather_bytes(bytes([rank]) * (rank + 1))
    b = rdv.broadcast_fd(r if rank == 0 else None)
    rdv.barrier()
    q.put((rank, msg, [len(x) for x in blobs], b >= 0))
    rdv.close()


def test_unix_rendezvous_passes_fds_between_processes():
    world = 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    session = f"test_{os.getpid()}"
    procs = [ctx.Process(target=_rdv_proc, args=(r, world, session, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=60) for _ in range(world))
    for p in procs:
        p.join(30)
    for rank, msg, lens, ok in res:
        assert msg == f"hello from {(rank - 1) % world}"
        assert lens == [1, 2, 3] and ok


def test_numa_affinity_helper_is_best_effort():
    from akka_allreduce_b200.utils import affinity

    assert affinity._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert affinity.bind_to_gpu_numa(0) in ({},) or isinstance(affinity.bind_to_gpu_numa(0), dict)   # no GPU here: must not raise


def test_metrics_endpoint_serves_master_and_worker_counters():
    import urllib.request

    from akka_allreduce_b200.utils.metrics import MetricsServer, render_prometheus

    master = AllreduceMaster(1, 1.0, 1.0, 1.0, max_lag=1, data_size=4, max_round=3, max_chunk_size=2)
    master.start(30)
    w = AllreduceWorker(lambda req: AllReduceInput(np.ones(4, dtype=np.float32)), lambda out: None, backend="cpu")
    w.run("127.0.0.1", master.port, 30)
    assert master.join(10)
    srv = MetricsServer({"master": master.metrics, "worker": w.metrics}).start()
    try:
        text = urllib.request.urlopen(f"http://127.0.0.1:{srv.port}/metrics", timeout=5).read().decode()
    finally:
        srv.stop()
    assert 'aar_rounds_started{role="master"} 4' in text
    assert 'aar_rounds_completed{role="worker"} 4' in text
    assert render_prometheus({"algo": "nvls", "push_dropped": False}, labels={"rank": "3"}) == \
        'aar_algo_info{rank="3",value="nvls"} 1\naar_push_dropped{rank="3"} 0\n'
    # histograms (the device-plane worker's lag_hist) become one series per bucket; label values are escaped
    assert render_prometheus({"lag_hist": {0: 5, 1: 2}, "note": 'a"b'}, labels={"role": "worker"}) == \
        'aar_lag_hist{role="worker",bucket="0"} 5\naar_lag_hist{role="worker",bucket="1"} 2\naar_note_info{role="worker",value="a\\"b"} 1\n'


def test_log_level_applies_to_both_runtimes():
    from akka_allreduce_b200.utils import log

    log.set_level("ERROR")
    assert log.get_logger().level == 40
    log.set_level("WARNING")


def test_rendezvous_turns_away_out_of_range_and_duplicate_ranks():
    """The rendezvous hands out the fds of the job's device memory: a connection that claims a rank outside 1..world-1, or one
    that is already taken, is closed and the rendezvous completes with the real peers."""
    import os
    import socket
    import struct
    import threading

    from akka_allreduce_b200.parallel.symmetric import UnixRendezvous

    session = f"test_rdv_{os.getpid()}"
    box = {}
    t = threading.Thread(target=lambda: box.setdefault("r0", UnixRendezvous(0, 2, session, timeout=30)))
    t.start()
    addr = "\0aar_rdv_" + session
    import time

    for bogus in (7, 0, -3):
        for _ in range(200):
            s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            try:
                s.connect(addr)
                break
            except (FileNotFoundError, ConnectionRefusedError):
                s.close()
                time.sleep(0.01)
        s.sendall(struct.pack("i", bogus))
        s.settimeout(5)
        assert s.recv(1) == b"", "the rendezvous must close a connection with an impossible rank"
        s.close()
    r1 = UnixRendezvous(1, 2, session, timeout=30)
    t.join(30)
    assert not t.is_alive()
    got = {}
    t2 = threading.Thread(target=lambda: got.setdefault("a", box["r0"].all_gather_bytes(b"zero")))
    t2.start()
    assert r1.all_gather_bytes(b"one!") == [b"zero", b"one!"]
    t2.join(10)
    assert got["a"] == [b"zero", b"one!"]
    r1.close()
    box["r0"].close()
