"""GPT-2 small data-parallel training step (BASELINE.json config 4): bf16 parameters and gradients, gradient buckets
averaged through the worker source/sink contract by this framework's kernels, overlapped with backward -- against the
same step with NCCL all_reduce on the same buckets.  Run under torch.distributed.run; prints tokens/s (whole job).
"""
import argparse
import json
import os
import struct
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from akka_allreduce_b200.models.gpt2 import GPT2, GPT2Config  # noqa: E402
from akka_allreduce_b200.parallel.ddp import ThresholdDDP  # noqa: E402
from akka_allreduce_b200.parallel.symmetric import UnixRendezvous  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--impl", default="both", choices=["aar", "nccl", "none", "both"])
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--seq", type=int, default=1024)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--bucket-mb", type=float, default=32.0)
ap.add_argument("--layers", type=int, default=12)
ap.add_argument("--comm-ctas", type=int, default=0, help="SM budget of the gradient allreduce kernels (0 = all SMs)")
ap.add_argument("--algo", default="auto")
ap.add_argument("--configs", default="", help="several framework arms in one process: algo:ctas[:priority[:pushCTAs:reduceCTAs]],... e.g. bulk:0:high,bulk:16:high:4:7")
a = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
lr = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
sess = f"g2_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
rdv = UnixRendezvous(rank, world, sess + "_c")


def gmax(x):
    return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))


def build():
    torch.manual_seed(0)
    m = GPT2(GPT2Config(n_layer=a.layers)).to(dev).to(torch.bfloat16)
    return m


def data(step):
    g = torch.Generator(device="cpu").manual_seed(1000 * rank + step)
    x = torch.randint(0, 50257, (a.batch, a.seq + 1), generator=g)
    return x[:, :-1].to(dev), x[:, 1:].to(dev)


res = {"world": world, "batch_per_gpu": a.batch, "seq": a.seq, "layers": a.layers, "bucket_mb": a.bucket_mb}


def run(impl, key=None, algo=None, ctas=None, prio=None, split=None):
    key = key or impl
    model = build()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, fused=True)
    ddp = None
    if impl == "aar":
        if prio:
            os.environ["AAR_STREAM_PRIORITY"] = prio
        ddp = ThresholdDDP(model, rank, world, lr, bucket_mb=a.bucket_mb, session=sess + "_" + key.replace(":", "_"),
                           comm_ctas=a.comm_ctas if ctas is None else ctas, algo=algo or a.algo, comm_split=split)
        buckets = ddp.buckets
    else:
        import torch.distributed as dist

        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=dev)
        # same bucket layout, NCCL on a side stream launched from the same hooks
        ddp = ThresholdDDP(model, rank, 1, lr, bucket_mb=a.bucket_mb, session=sess + "_" + impl)  # world=1 engines are unused below
        comm = torch.cuda.Stream()
        works = []

        def launch(b):
            if impl == "none":        # no gradient exchange at all: the compute-only floor of the step (not a training run)
                return
            comm.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(comm):
                dist.all_reduce(b.buffer)
                b.buffer.div_(world)

        ddp._launch = launch

        def finish():
            torch.cuda.current_stream().wait_stream(comm)
            ddp.step += 1
            ddp._arm()

        ddp.finish = finish
    losses = []
    times = []
    for step in range(a.warmup + a.steps):
        x, y = data(step)
        torch.cuda.synchronize(); rdv.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loss = model(x, y)
        loss.backward()
        ddp.finish()
        opt.step()
        ddp.zero_grad()
        e1.record()
        torch.cuda.synchronize()
        if step >= a.warmup:
            times.append(e0.elapsed_time(e1))
        losses.append(float(loss))
    ms = gmax(sum(times) / len(times))
    toks = world * a.batch * a.seq / (ms * 1e-3)
    res[key] = {"ms_per_step": round(ms, 3), "tokens_per_sec": round(toks, 1), "loss_first": round(losses[0], 4), "loss_last": round(losses[-1], 4),
                 "buckets": len(ddp.buckets), "bucket_elems": ddp.buckets[0].buffer.numel(),
                 "algo": ddp.buckets[0].engine.algo if impl == "aar" else impl, "params": model.num_params()}
    ddp.close()
    del model, opt
    torch.cuda.empty_cache()


if a.configs:
    for cfg in a.configs.split(","):
        f = cfg.split(":")
        algo, ctas, prio = f[0], (f[1] if len(f) > 1 else "0"), (f[2] if len(f) > 2 else "high")
        split = (int(f[3]), int(f[4])) if len(f) > 4 else None     # algo:ctas:priority:pushCTAs:reduceCTAs
        run("aar", key=cfg, algo=algo, ctas=int(ctas), prio=prio or "high", split=split)
    run("nccl")
    run("none")
    for cfg in a.configs.split(","):
        res[cfg]["vs_nccl"] = round(res["nccl"]["ms_per_step"] / res[cfg]["ms_per_step"], 3)
    floor = res["none"]["ms_per_step"]
    for k in list(a.configs.split(",")) + ["nccl"]:
        res[k]["exposed_comm_ms"] = round(res[k]["ms_per_step"] - floor, 3)
else:
    for impl in (["aar", "nccl"] if a.impl == "both" else [a.impl]):
        run(impl)
if "aar" in res and "nccl" in res:
    res["aar_vs_nccl"] = round(res["nccl"]["ms_per_step"] / res["aar"]["ms_per_step"], 3)
if rank == 0:
    print(json.dumps(res))
rdv.barrier()
