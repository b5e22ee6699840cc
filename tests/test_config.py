"""application.conf-equivalent parser (csrc/core/config.cpp) -- reference file: src/main/resources/application.conf:1-34."""
import os
from pathlib import Path

import pytest

from akka_allreduce_b200 import _core as core

ROOT = Path(__file__).resolve().parents[1]

REFERENCE_STYLE = """
akka {
  actor {
    provider = cluster
  }
  remote {
    log-remote-lifecycle-events = off
    netty.tcp {
      hostname = "127.0.0.1"
      port = 0
    }
  }

  cluster {
    seed-nodes = [
      "akka.tcp://ClusterSystem@127.0.0.1:2551",
      "akka.tcp://ClusterSystem@127.0.0.1:2552"]

    # auto downing is NOT safe for production deployments.
    auto-down-unreachable-after = 10s
  }
  loglevel = "INFO"
  log-dead-letters = 5
}

# Disable legacy metrics in akka-cluster.
akka.cluster.metrics.enabled=off
akka.extensions=["akka.cluster.metrics.ClusterMetricsExtension"]
akka.cluster.metrics.native-library-extract-folder=${user.dir}/target/native
"""


def test_parses_reference_style_file():
    c = core.Config.parse_string(REFERENCE_STYLE)
    assert c.get_string("akka.actor.provider") == "cluster"
    assert c.get_string("akka.remote.netty.tcp.hostname") == "127.0.0.1"
    assert c.get_int("akka.remote.netty.tcp.port") == 0
    assert c.get_list("akka.cluster.seed-nodes") == [
        "akka.tcp://ClusterSystem@127.0.0.1:2551",
        "akka.tcp://ClusterSystem@127.0.0.1:2552",
    ]
    assert c.get_duration_ms("akka.cluster.auto-down-unreachable-after") == 10000
    assert c.get_string("akka.loglevel") == "INFO"
    assert c.get_int("akka.log-dead-letters") == 5
    assert c.get_bool("akka.cluster.metrics.enabled") is False
    assert c.get_bool("akka.remote.log-remote-lifecycle-events") is False
    assert c.get_list("akka.extensions") == ["akka.cluster.metrics.ClusterMetricsExtension"]
    assert c.get_string("akka.cluster.metrics.native-library-extract-folder") == os.getcwd() + "/target/native"


def test_with_fallback_layers_like_the_reference_mains():
    # AllreduceMaster.scala:116-118: port override, then role, then the file
    file_cfg = core.Config.parse_string(REFERENCE_STYLE)
    layered = (
        core.Config.parse_string("akka.remote.netty.tcp.port=2551")
        .with_fallback(core.Config.parse_string("akka.cluster.roles = [master]"))
        .with_fallback(file_cfg)
    )
    assert layered.get_int("akka.remote.netty.tcp.port") == 2551
    assert layered.get_list("akka.cluster.roles") == ["master"]
    assert layered.get_string("akka.loglevel") == "INFO"


def test_shipped_conf_and_settings():
    c = core.Config.parse_file(str(ROOT / "conf" / "application.conf"))
    s = core.AllreduceSettings.from_config(c)
    # the reference's hard-coded defaults (AllreduceMaster.scala:105-114)
    assert s.master.total_workers == 2
    assert s.master.th_allreduce == pytest.approx(1.0)
    assert s.master.th_reduce == pytest.approx(0.9)
    assert s.master.th_complete == pytest.approx(0.8)
    assert s.master.max_lag == 1
    assert s.master.max_round == 100
    assert s.master.data_size == 10
    assert s.master.max_chunk_size == 2
    assert s.hostname == "127.0.0.1" and s.port == 2551
    assert s.auto_down_after_ms == 10000
    assert s.dtype == "fp32" and s.algo == "auto" and s.normalize is False


def test_substitution_optional_env_and_errors(monkeypatch):
    monkeypatch.setenv("AAR_TEST_ENV", "hello")
    c = core.Config.parse_string('a = ${AAR_TEST_ENV}\nb = ${?AAR_NOT_SET_ANYWHERE}x\nc { d : 3, e : "q" }\nf = ${c.d}')
    assert c.get_string("a") == "hello"
    assert c.get_string("b") == "x"
    assert c.get_int("c.d") == 3 and c.get_string("c.e") == "q" and c.get_int("f") == 3
    with pytest.raises(Exception, match="could not resolve"):
        core.Config.parse_string("a = ${nope.nothing}")
    with pytest.raises(Exception, match="line 1"):
        core.Config.parse_string("a {")
    with pytest.raises(Exception, match="not an integer"):
        core.Config.parse_string("a = xyz").get_int("a")


def test_render_round_trip():
    c = core.Config.parse_string(REFERENCE_STYLE)
    again = core.Config.parse_string(c.render())
    assert again.keys() == c.keys()
    for k in c.keys():
        assert again.get_list(k) == c.get_list(k)


def test_settings_validation():
    with pytest.raises(Exception, match="thresholds"):
        core.AllreduceSettings.from_config(core.Config.parse_string("allreduce.th-reduce = 1.5"))
    with pytest.raises(Exception, match="dtype"):
        core.AllreduceSettings.from_config(core.Config.parse_string("allreduce.dtype = fp8"))


def test_settings_reject_out_of_range_and_unknown_values():
    """Found by fuzzing fromConfig: a negative data-size, an integer that does not fit in 32 bits, a port above 65535 or an unknown
    algorithm / quorum policy used to be accepted (or narrowed to garbage)."""
    def settings(body):
        return core.AllreduceSettings.from_config(core.Config.parse_string("allreduce {\n" + body + "\n}"))

    for body in ("data-size = -1", "data-size = 99999999999", "max-round = 99999999999", "transport.port = 70000", "transport.port = -1",
                 "pipeline-depth = -2", "algo = fastest", "quorum-policy = most", "total-workers = 0", "max-chunk-size = 0",
                 "th-reduce = nan", "th-complete = 1.5", "start-round = -3", "transport.heartbeat-interval = 0s"):
        with pytest.raises(Exception):
            settings(body)
    s = settings("start-round = 7\n algo = pipelined\n quorum-policy = exact\n data-size = 0")
    assert s.master.start_round == 7 and s.algo == "pipelined" and s.quorum_policy == "exact" and s.master.data_size == 0


def test_every_kernel_family_the_engine_knows_is_accepted_by_the_config():
    """allreduce.algo is handed to RoundEngine(algo=...): the config must accept exactly the names the engine does."""
    from akka_allreduce_b200.parallel.engine import _ALGO_CODE

    for name in ["auto", *sorted(_ALGO_CODE)]:
        s = core.AllreduceSettings.from_config(core.Config.parse_string("allreduce { algo = " + name + " }"))
        assert s.algo == name
    with pytest.raises(Exception, match="nvls_onehop"):
        core.AllreduceSettings.from_config(core.Config.parse_string("allreduce { algo = ring }"))
