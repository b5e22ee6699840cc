"""Rounds per second through the public API (AllreduceMaster + AllreduceWorker(backend="cuda")) for small messages, where the
host loop is what limits the rate: the round loop in Python against the round loop in C++ (csrc/core/device_loop.cpp).
Run under torch.distributed.run; the source hands out a tensor that is already on the device, the sink does nothing."""
import argparse
import json
import os
import struct
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput  # noqa: E402
from akka_allreduce_b200.parallel.symmetric import UnixRendezvous  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", default="16384,262144,4194304", help="elements (fp32) per round")
ap.add_argument("--rounds", type=int, default=3000)
ap.add_argument("--max-lag", type=int, default=3)
a = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
base_port = int(os.environ.get("MASTER_PORT", "29500")) + 100
sess = f"rate_{base_port}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")
res = {"world": world, "rounds": a.rounds, "max_lag": a.max_lag, "rows": []}
k = 0
for n in [int(s) for s in a.sizes.split(",")]:
    row = {"elems": n, "bytes": 4 * n}
    x = torch.ones(n, device=f"cuda:{lr}")
    for native in (False, True):
        k += 1
        master = None
        if rank == 0:
            master = AllreduceMaster(world, 1.0, 1.0, 1.0, a.max_lag, n, a.rounds - 1, max(1, min(1 << 20, n // world)), pipeline_depth=a.max_lag,
                                     port=base_port + k)
            master.start(300)
        rdv.barrier()
        w = AllreduceWorker(lambda req: AllReduceInput(x), lambda out: None, backend="cuda", device=lr, session=f"{sess}_{k}", native_loop=native)
        ok = w.run("127.0.0.1", base_port + k, timeout_sec=300)
        if master:
            master.join(30)
        m = w.metrics()
        rps = min(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", m.get("rounds_per_sec", 0.0) if ok else 0.0)))
        row["cpp_loop" if native else "python_loop"] = {"rounds_per_sec": round(rps, 1), "us_per_round": round(1e6 / rps, 2) if rps else None,
                                                        "algo": m.get("algo"), "delivered": m["rounds_delivered"]}
        w.close()
        rdv.barrier()
    row["cpp_vs_python"] = round(row["cpp_loop"]["rounds_per_sec"] / max(1e-9, row["python_loop"]["rounds_per_sec"]), 3)
    res["rows"].append(row)
if rank == 0:
    print(json.dumps(res))
