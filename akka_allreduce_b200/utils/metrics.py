"""Metrics export: the counters of the master, the workers and the device engines as Prometheus text over HTTP.

The reference switches on akka-cluster-metrics + Sigar (application.conf:26-34) but never reads them; an operator of an
allreduce service wants rounds/s, lag, forced completions, contributor counts and timeouts.  `MetricsServer` serves whatever the
registered providers return (dicts of name -> number/bool/str) at /metrics; no third-party dependency."""
from __future__ import annotations

import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Mapping


def _esc(v: object) -> str:
    """Label values: backslash, double quote and newline are escaped (Prometheus text format)."""
    return str(v).replace("\\", "\\\\").replace('"', '\\"').replace("\n", "\\n")


def render_prometheus(metrics: Mapping[str, object], prefix: str = "aar", labels: Mapping[str, str] | None = None) -> str:
    lab = ""
    if labels:
        lab = "{" + ",".join(f'{k}="{v}"' for k, v in sorted(labels.items())) + "}"
    lines = []
    for name, value in sorted(metrics.items()):
        key = f"{prefix}_{name}".replace("-", "_").replace(".", "_")
        if isinstance(value, bool):
            value = int(value)
        if isinstance(value, (int, float)):
            lines.append(f"{key}{lab} {value}")
        elif isinstance(value, Mapping):  # histograms (e.g. lag_hist): one series per bucket
            for k, v in sorted(value.items(), key=lambda kv: str(kv[0])):
                if isinstance(v, (int, float)) and not isinstance(v, bool):
                    extra = (lab[:-1] + "," if lab else "{") + f'bucket="{_esc(k)}"' + "}"
                    lines.append(f"{key}{extra} {v}")
        else:  # strings become info-style series
            extra = (lab[:-1] + "," if lab else "{") + f'value="{_esc(value)}"' + "}"
            lines.append(f"{key}_info{extra} 1")
    return "\n".join(lines) + "\n"


class MetricsServer:
    """`MetricsServer({"master": master.metrics, "worker": worker.metrics}, port=0).start()`; GET /metrics."""

    def __init__(self, providers: Mapping[str, Callable[[], Mapping[str, object]]], host: str = "127.0.0.1", port: int = 0):
        self.providers = dict(providers)
        outer = self

        class Handler(BaseHTTPRequestHandler):
            def do_GET(self):  # noqa: N802
                if self.path.rstrip("/") not in ("/metrics", ""):
                    self.send_error(404)
                    return
                body = outer.render().encode()
                self.send_response(200)
                self.send_header("Content-Type", "text/plain; version=0.0.4")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def log_message(self, *a):  # quiet
                pass

        self.httpd = ThreadingHTTPServer((host, port), Handler)
        self.port = self.httpd.server_address[1]
        self._thread: threading.Thread | None = None

    def render(self) -> str:
        out = []
        for role, fn in self.providers.items():
            try:
                out.append(render_prometheus(fn(), labels={"role": role}))
            except Exception as e:  # a provider that is not ready yet must not take the endpoint down
                out.append(f'aar_provider_error{{role="{role}",error="{type(e).__name__}"}} 1\n')
        return "".join(out)

    def start(self) -> "MetricsServer":
        self._thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self.httpd.shutdown()
        self.httpd.server_close()
