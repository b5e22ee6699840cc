"""AllreduceMaster: the coordinator process (admission, ids, round clock, completion quorum).

Thin Python face of the C++ coordinator (csrc/core/master.cpp + tcp_transport.cpp); behaviour and
constructor arguments follow AllreduceMaster.scala:15-24, with the knobs the reference hard-codes in
its main (:105-109) taken from the `allreduce { }` block of the config file.
"""
from __future__ import annotations

import threading
from pathlib import Path

from . import _core as core


def load_settings(conf: str | Path | None = None, overrides: dict | None = None, role: str = "master") -> core.AllreduceSettings:
    """Layered configuration like the reference mains: explicit overrides, then the role, then the file
    (AllreduceMaster.scala:116-118)."""
    text = "\n".join(f"{k} = {v}" for k, v in (overrides or {}).items())
    cfg = core.Config.parse_string(text).with_fallback(core.Config.parse_string(f"akka.cluster.roles = [{role}]"))
    path = Path(conf) if conf else Path(__file__).resolve().parents[1] / "conf" / "application.conf"
    if path.exists():
        cfg = cfg.with_fallback(core.Config.parse_file(str(path)))
    return core.AllreduceSettings.from_config(cfg)


class AllreduceMaster:
    def __init__(self, total_workers: int, th_allreduce: float = 1.0, th_reduce: float = 1.0, th_complete: float = 1.0, max_lag: int = 1,
                 data_size: int = 10, max_round: int = 100, max_chunk_size: int = 2, *, pipeline_depth: int = 0, host: str = "127.0.0.1",
                 port: int = 0, auto_down_after_ms: float = 10000.0, start_round: int = 0, checkpoint: str | Path | None = None,
                 checkpoint_every: int = 100, resume: bool = False):
        """checkpoint: file the coordinator state is written to every `checkpoint_every` quorum rounds and when run() ends
        (replaced atomically).  resume=True continues the round numbering of that file if it exists: the first admission starts
        at the first round that had not reached its quorum (the reference always restarts at 0).  start_round sets that round
        by hand."""
        self.config = core.MasterConfig(total_workers, th_allreduce, th_reduce, th_complete, max_lag, data_size, max_round, max_chunk_size,
                                        pipeline_depth, start_round)
        self.node = core.MasterNode(self.config, host, port, auto_down_after_ms)
        self.start_round = start_round
        if checkpoint is not None:
            if resume and Path(checkpoint).exists():
                self.start_round = core.resume_master_from_checkpoint(str(checkpoint), self.node.master())
            self.node.set_checkpoint(str(checkpoint), checkpoint_every)
        self.host = host
        self._thread: threading.Thread | None = None
        self.finished = False

    @classmethod
    def from_settings(cls, s: core.AllreduceSettings, **kw):
        m = s.master
        return cls(m.total_workers, m.th_allreduce, m.th_reduce, m.th_complete, m.max_lag, m.data_size, m.max_round, m.max_chunk_size,
                   pipeline_depth=m.pipeline_depth, host=s.hostname, port=s.port, auto_down_after_ms=s.auto_down_after_ms,
                   **{"start_round": m.start_round, **kw})

    @property
    def port(self) -> int:
        return self.node.port

    def run(self, timeout_sec: float = 3600.0) -> bool:
        """Serve until every round reached its quorum, then tell the workers to shut down."""
        self.finished = self.node.run(timeout_sec)
        return self.finished

    def start(self, timeout_sec: float = 3600.0) -> None:
        """Run in a background thread (the event loop releases the GIL)."""
        self._thread = threading.Thread(target=self.run, args=(timeout_sec,), daemon=True)
        self._thread.start()

    def join(self, timeout: float | None = None) -> bool:
        if self._thread:
            self._thread.join(timeout)
        return self.finished

    def metrics(self) -> dict:
        return self.node.master().metrics()
