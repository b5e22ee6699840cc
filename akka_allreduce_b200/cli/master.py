"""python -m akka_allreduce_b200.cli.master [port] [totalWorkers] [dataSize] [maxChunkSize] [--conf file] [--set k=v]...
                                           [--checkpoint file [--checkpoint-every rounds] [--resume]]

Same positional arguments as the reference main (AllreduceMaster.scala:101-138)."""
import argparse
import sys

from ..master import AllreduceMaster, load_settings


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("positional", nargs="*")
    ap.add_argument("--conf")
    ap.add_argument("--set", action="append", default=[])
    ap.add_argument("--timeout", type=float, default=3600.0)
    ap.add_argument("--checkpoint", help="coordinator state file, replaced atomically every --checkpoint-every quorum rounds and at exit")
    ap.add_argument("--checkpoint-every", type=int, default=100)
    ap.add_argument("--resume", action="store_true", help="continue the round numbering of --checkpoint if the file exists")
    a = ap.parse_args(argv)
    over = dict(kv.split("=", 1) for kv in a.set)
    keys = ["allreduce.transport.port", "allreduce.total-workers", "allreduce.data-size", "allreduce.max-chunk-size"]
    for k, v in zip(keys, a.positional):
        over[k] = v
    s = load_settings(a.conf, over, role="master")
    m = AllreduceMaster.from_settings(s, checkpoint=a.checkpoint, checkpoint_every=a.checkpoint_every, resume=a.resume)
    if m.start_round:
        print(f"master: resuming at round {m.start_round} from {a.checkpoint}", flush=True)
    print(f"master: port={m.port} workers={s.master.total_workers} dataSize={s.master.data_size} maxChunkSize={s.master.max_chunk_size}", flush=True)
    ok = m.run(a.timeout)
    print(f"master: finished={int(ok)} metrics={m.metrics()}", flush=True)
    return 0 if ok else 2


if __name__ == "__main__":
    sys.exit(main())
