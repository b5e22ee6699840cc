"""python -m akka_allreduce_b200.cli.worker [port] [dataSize] [--backend cuda|cpu] [--device N] [--conf file] [--set k=v]...

Same positional arguments as the reference main (AllreduceWorker.scala:272-302): synthetic source data[i] = i + iteration,
sink reports the iteration.  One process per GPU on the device data plane."""
import argparse
import sys

import numpy as np

from ..api import AllReduceInput
from ..master import load_settings
from ..worker import AllreduceWorker


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("positional", nargs="*")
    ap.add_argument("--backend", default="cuda", choices=["cuda", "cpu"])
    ap.add_argument("--device", type=int, default=None)
    ap.add_argument("--conf")
    ap.add_argument("--set", action="append", default=[])
    ap.add_argument("--timeout", type=float, default=3600.0)
    ap.add_argument("--quiet", action="store_true")
    a = ap.parse_args(argv)
    s = load_settings(a.conf, dict(kv.split("=", 1) for kv in a.set), role="worker")
    listen_port = int(a.positional[0]) if a.positional else 0
    data_size = int(a.positional[1]) if len(a.positional) > 1 else 10
    flushed = [0]

    if a.backend == "cpu":
        def source(req):
            return AllReduceInput(np.arange(data_size, dtype=np.float32) + np.float32(req.iteration))
    else:
        import torch

        base = {}

        def source(req):
            dev = torch.cuda.current_device()
            if dev not in base:
                base[dev] = torch.arange(data_size, dtype=torch.float32, device=f"cuda:{dev}")
            x = base[dev] + float(req.iteration)
            return AllReduceInput(x if s.dtype == "fp32" else x.to(torch.bfloat16))

    def sink(out):
        flushed[0] += 1
        if not a.quiet:
            print(f"----Data output at #{out.iteration}: {len(out.data)} elements, contributors min {int(out.count.min())}", flush=True)

    w = AllreduceWorker(source, sink, backend=a.backend, device=a.device, dtype=s.dtype, algo=s.algo, normalize=s.normalize,
                        exact_quorum=(s.quorum_policy == "exact"), spin_timeout_ms=s.spin_timeout_us / 1e3)
    w.connect(s.hostname, s.port, listen_port, s.heartbeat_ms)
    ok = w.run(timeout_sec=a.timeout)
    w.close()
    print(f"worker {w.id}: shutdown={int(ok)} rounds_flushed={flushed[0]}", flush=True)
    return 0 if ok else 2


if __name__ == "__main__":
    sys.exit(main())
