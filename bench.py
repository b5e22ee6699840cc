#!/usr/bin/env python
"""Headline benchmark (BASELINE.json config 2): N workers on N B200s, dataSize = 16 Mi fp32 (64 MiB per worker per
round), maxChunkSize = 1 Mi, thresholds = 1.0, maxLag = 1, synthetic data, compared with NCCL allreduce of the same
buffers on the same box.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

Prints ONE JSON line (rank 0).  `value` is the whole-job aggregate: N x payload bytes x rounds/s, in GB/s (weak
scaling: every GPU contributes a 64 MiB buffer every round); `busbw_GBps` (NCCL convention, algbw x 2(N-1)/N) and
`rounds_per_sec` ride along, as does the NCCL measurement taken in the same process.  Device-timed with CUDA events
on the launching stream, max over ranks.  `e2e` is the same workload through the public API (AllreduceMaster +
AllreduceWorker with a data source that copies the step's input from pinned host memory and a sink that reads the
step's contributor counts and a checksum slice back to the host).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="aar", choices=["aar", "reference", "nccl"])
    ap.add_argument("--data-size", type=int, default=16 * 1024 * 1024)
    ap.add_argument("--chunk", type=int, default=1024 * 1024)
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--max-lag", type=int, default=1)
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--grid", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--view", action="store_true", help="zero-copy output (sink reads the ring slot)")
    ap.add_argument("--stage-input", action="store_true", help="zero-copy input on the in-switch path: the inputs live in the engine's "
                    "multicast-bound staging slots (RoundEngine.stage_input); not the default, reported in config")
    ap.add_argument("--nvls-ctas", default="", help="nStage:nReduce for the NVLS kernel")
    ap.add_argument("--split", default="", help="nPush:nReduce for the pipelined P2P kernel")
    ap.add_argument("--no-overlap", action="store_true", help="one round in flight at a time (no use of the maxLag window)")
    return ap.parse_args()


def reference_unavailable():
    # The reference is a Scala/sbt/Akka project; it is not a Python distribution, so the offline pip install has
    # nothing to build, and the box has no JVM, sbt or ivy cache (DESIGN.md "Reference arm").
    print(json.dumps({"impl": "reference", "unavailable": "reference is a Scala/sbt Akka project: no setup.py/pyproject for pip, "
                      "and no java/sbt/ivy cache on the box to run it (see DESIGN.md)"}))


class ClockSampler:
    """SM clocks and throttle reasons DURING the timed region (B200_PROFILING.md clocks line).

    Sampled through NVML inside this process at a moment the caller chooses: right after the K timed rounds have been enqueued,
    while the GPU is still executing them.  A free-running `nvidia-smi -lms` beside the benchmark is what this replaces: its
    queries take driver locks at random moments, and one landing inside a 3 ms timed region stalls the launches of that rank
    (measured: 189 us instead of 167 us per round in one run out of five at 4 GPUs).  Without NVML the old sampler is the
    fallback."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, gpu_index: int):
        self.sm, self.mx, self.reasons = [], [], set()
        self.nvml = self.handle = self.proc = None
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = [v for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip()]
            phys = int(visible[gpu_index]) if gpu_index < len(visible) and visible[gpu_index].strip().isdigit() else gpu_index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
            self.sample()          # resolves the entry points before the timed region; dropped below
            self.sm.clear(), self.mx.clear(), self.reasons.clear()
        except Exception:
            self.nvml = None
            self._start_nvidia_smi(gpu_index)

    def sample(self, n: int = 1) -> None:
        """Take n samples now (the caller has just enqueued the timed work and has not synchronised yet)."""
        if self.nvml is None:
            return
        nv = self.nvml
        for _ in range(n):
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)))
                self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(self.handle, nv.NVML_CLOCK_SM)))
                get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                mask = int(get(self.handle))
                self.reasons.update(name for bit, name in self.REASONS if mask & bit)
            except Exception:
                return

    def _start_nvidia_smi(self, gpu_index: int) -> None:
        self.path = f"/tmp/aar_clocks_{os.getpid()}.csv"
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(gpu_index)],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.nvml is not None:
            if self.sm:
                out.update({"sm_mhz": statistics.median(self.sm), "sm_max_mhz": max(self.mx), "samples": len(self.sm),
                            "how": "NVML, sampled while the timed rounds were executing"})
            out["reasons"] = sorted(self.reasons)
            try:
                self.nvml.nvmlShutdown()
            except Exception:
                pass
            return out
        if not self.proc:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                p = [x.strip() for x in line.split(",")]
                if len(p) < 9:
                    continue
                try:
                    sm.append(float(p[1]))
                    mx.append(float(p[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out["sm_mhz"] = statistics.median(sm)
            out["sm_max_mhz"] = max(mx)
            out["samples"] = len(sm)
            out["how"] = "nvidia-smi -lms 100 beside the benchmark"
        out["reasons"] = sorted(reasons)
        return out


def main():
    a = parse_args()
    if a.impl == "reference":
        if int(os.environ.get("RANK", "0")) == 0:   # one line per job, also when launched with one process per GPU
            reference_unavailable()
        return 0

    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != a.gpus:
        if a.gpus > 1 and world == 1:
            print(f"bench.py --gpus {a.gpus} must be launched with torch.distributed.run (one process per GPU)", file=sys.stderr)
            return 2
        a.gpus = world
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from akka_allreduce_b200.utils.affinity import bind_to_gpu_numa

    numa = bind_to_gpu_numa(local_rank)   # pinned source buffers of the end-to-end run land next to this GPU's PCIe root
    dtype = torch.float32 if a.dtype == "fp32" else torch.bfloat16
    es = 4 if a.dtype == "fp32" else 2
    nbytes = a.data_size * es
    N, K, W = a.gpus, a.steps, a.warmup

    from akka_allreduce_b200.ops import lib as L
    from akka_allreduce_b200.parallel.engine import ONESHOT_MAX_BYTES, RoundEngine
    from akka_allreduce_b200.parallel.symmetric import SymmetricMemory, UnixRendezvous

    session = f"bench_{os.environ.get('MASTER_PORT', '0')}_{os.getppid() if world > 1 else os.getpid()}"
    rdv = UnixRendezvous(rank, world, session + "_ctl")

    def max_over_ranks(x: float) -> float:
        import struct

        return max(struct.unpack("d", b)[0] for b in rdv.all_gather_bytes(struct.pack("d", x)))

    def global_sync():
        torch.cuda.synchronize(dev)
        rdv.barrier()

    result = {}
    sampler = ClockSampler(local_rank) if rank == 0 else None

    # synthetic inputs: rank r contributes (r + 1) * base + round-dependent offset; two buffers alternate so the
    # working set (inputs + rings, > 300 MB per rank) stays far above the 126 MB L2
    g = torch.Generator(device="cpu").manual_seed(1234)
    base = torch.randn(a.data_size, generator=g).to(dtype)
    xs = [(base * (rank + 1) + k).to(dev) for k in range(2)]

    # ---------------------------------------------------------------- device-timed rounds
    if a.impl == "aar":
        layout = L.compute_layout(N, a.data_size, a.chunk, a.max_lag, L.AAR_F32 if a.dtype == "fp32" else L.AAR_BF16, ONESHOT_MAX_BYTES // es, True)
        if N == 1:
            eng = RoundEngine.single_process(a.data_size, a.chunk, a.max_lag, dtype=dtype, device=local_rank, grid=a.grid, algo=a.algo)
            symm = None
        else:
            symm = SymmetricMemory(UnixRendezvous(rank, world, session + "_mem"), layout.regionBytes, local_rank, want_multicast=True)
            eng = RoundEngine.from_symmetric(symm, layout, grid=a.grid, algo=a.algo, overlap=False if a.no_overlap else None)
        if a.view:
            eng.base_flags |= L.FLAG_VIEW_OUTPUT
        if a.nvls_ctas:
            eng.nvls_stage_ctas, eng.nvls_reduce_ctas = (int(v) for v in a.nvls_ctas.split(":"))
        if a.split:
            ps, rs = (int(v) for v in a.split.split(":"))
            eng.split = eng.bulk_split = eng.nvls_bulk_split = ps | (rs << 16)
        staged = False
        if a.stage_input and a.max_lag == 1 and eng.stage_input(0) is not None:   # slot of round r is r % 2 = index of its input
            xs = [eng.stage_input(k).copy_(xs[k]) for k in range(2)]
            staged = True
        global_sync()
        rnd = 0
        for _ in range(W):
            eng.enqueue_round(rnd, xs[rnd % 2])
            rnd += 1
        eng.wait_all()
        global_sync()
        launches0 = eng.launches
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(K):
            out, counts = eng.enqueue_round(rnd, xs[rnd % 2])
            rnd += 1
        eng.wait_all()   # the launching stream joins the round streams before the end event
        t1.record()
        if sampler:
            sampler.sample(3)   # the host is ahead of the device: the timed rounds are still executing
        global_sync()
        ms = max_over_ranks(t0.elapsed_time(t1) / K)
        launches = eng.launches - launches0
        # correctness of the last timed round, outside the timed region: the FULL vector against the ordered fp32 sum over the
        # ranks (peer-index order from 0.0f, AllreduceWorker.scala:245-249).  Bit-exact on the P2P paths; the in-switch sum does not
        # promise that order, so there: equal up to fp32 reassociation AND the same bits on every rank.
        import struct

        last = (rnd - 1) % 2
        acc = torch.zeros(a.data_size, dtype=torch.float32, device=dev)
        for r in range(N):
            acc += (base * (r + 1) + last).to(dev).to(torch.float32)   # base is already in `dtype`: the same rounding as the inputs
        want = acc.to(dtype)
        del acc
        if a.view:
            lay = eng.layout
            got = eng._region_view(lay.reduceOff + ((rnd - 1) % lay.L) * lay.P * lay.blkStride * lay.elemSize, a.data_size * lay.elemSize, dtype)
        else:
            got = out
        in_switch = eng.algo in ("nvls", "nvls_bulk", "nvls_onehop")
        if in_switch:
            ok = bool(torch.allclose(got.float(), want.float(), rtol=2e-5 if dtype == torch.float32 else 2e-2, atol=1e-4))
        else:
            ok = bool(torch.equal(got, want))
        digest = struct.pack("q", int(got.view(torch.int16 if dtype == torch.bfloat16 else torch.int32).to(torch.int64).sum().item()))
        ok = ok and len(set(rdv.all_gather_bytes(digest))) == 1          # every rank holds the same bits
        cmin, cmax = int(counts.min().item()), int(counts.max().item())
        ok = ok and cmin == N and cmax == N
        del want
        result.update({"algo": eng.algo, "status_bits": eng.status(), "verified": ok, "verified_how": "full vector vs ordered fp32 sum: " + ("allclose + identical bits on every rank (in-switch sum)" if in_switch else "bit-exact") , "multicast": bool(symm and symm.has_multicast),
                       "multicast_error": (symm.mc_error if symm else "")})
        gpu_launches = launches * N
    else:  # plain NCCL allreduce of the same buffers: the baseline named by BASELINE.json
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
        bufs = [x.clone() for x in xs]
        for i in range(W):
            dist.all_reduce(bufs[i % 2])
        global_sync()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(K):
            dist.all_reduce(bufs[i % 2])
        t1.record()
        if sampler:
            sampler.sample(3)
        global_sync()
        ms = max_over_ranks(t0.elapsed_time(t1) / K)
        gpu_launches = 0
        result.update({"algo": "nccl"})

    rounds_per_sec = 1e3 / ms
    algbw = nbytes / (ms * 1e-3) / 1e9
    value = N * algbw
    busbw = algbw * 2 * (N - 1) / N

    # ---------------------------------------------------------------- NCCL on the same box, same buffers
    nccl = None
    if a.impl == "aar" and N > 1 and not a.no_nccl:
        try:
            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=dev)
            bufs = [x.clone() for x in xs]
            for i in range(W):
                dist.all_reduce(bufs[i % 2])
            global_sync()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(K):
                dist.all_reduce(bufs[i % 2])
            e1.record()
            global_sync()
            nms = max_over_ranks(e0.elapsed_time(e1) / K)
            nalg = nbytes / (nms * 1e-3) / 1e9
            nccl = {"ms_per_step": nms, "algbw_GBps": nalg, "busbw_GBps": nalg * 2 * (N - 1) / N, "version": ".".join(map(str, torch.cuda.nccl.version()))}
        except Exception as e:  # the comparison must never break the headline number
            nccl = {"error": str(e)[:200]}

    # ---------------------------------------------------------------- end to end through the public API
    e2e = None
    if a.impl == "aar" and not a.no_e2e:
        try:
            e2e = run_e2e(a, rank, world, local_rank, dtype, nbytes, rdv, max_over_ranks, global_sync)
        except Exception as e:
            e2e = {"error": f"{type(e).__name__}: {e}"[:300]}

    clocks = sampler.stop() if sampler else None
    if rank == 0:
        line = {
            "metric": "allreduce aggregate payload GB/s at dataSize=16Mi fp32 per worker (N x bytes x rounds/s); bus GB/s and rounds/sec alongside, vs NCCL",
            "value": value, "unit": "GB/s", "n_gpus": N, "steps": K, "warmup": W, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
            "impl": a.impl,
            "config": {"model": "threshold allreduce, BASELINE.json config 2 (N workers, dataSize=16Mi, maxChunkSize=1Mi, thresholds=1.0, maxLag=1)",
                       "global_batch": N * a.data_size, "seq_len": a.data_size, "parallelism": f"dp{N}", "data_size": a.data_size,
                       "max_chunk_size": a.chunk, "max_lag": a.max_lag, "th_reduce": 1.0, "th_complete": 1.0,
                       "l2_policy": "inputs+rings larger than L2 (2 x 64 MiB inputs alternate, 256 MiB rings per rank)", "view_output": a.view, "stage_input": bool(a.impl == "aar" and a.stage_input), "rounds_in_flight": 2 if (a.impl == "aar" and N > 1 and a.max_lag >= 1 and not a.no_overlap) else 1},
            "rounds_per_sec": rounds_per_sec, "algbw_GBps": algbw, "busbw_GBps": busbw,
            "gpu_launches": gpu_launches, "clocks": clocks, "numa_binding": numa, "nccl": nccl,
            "vs_nccl": (nccl["ms_per_step"] / ms) if nccl and "ms_per_step" in nccl else None,
            "e2e": e2e,
        }
        line.update(result)
        print(json.dumps(line), flush=True)
    try:
        rdv.barrier()   # leave together; a rank that lost its peers (e.g. after an end-to-end failure elsewhere) still exits 0
    except Exception:
        pass
    return 0


def run_e2e(a, rank, world, local_rank, dtype, nbytes, rdv, max_over_ranks, global_sync):
    """The same rounds through AllreduceMaster + AllreduceWorker: per step the source copies the step's input from pinned
    host memory to the device, the sink reads the contributor counts and a 4 KiB slice of the result back to the host."""
    import torch

    from akka_allreduce_b200 import AllreduceMaster, AllreduceWorker, AllReduceInput

    N, K, W = a.gpus, a.steps, a.warmup
    dev = torch.device("cuda", local_rank)
    total = W + K
    port = int(os.environ.get("MASTER_PORT", "29500")) + 101
    master = None
    if rank == 0:
        master = AllreduceMaster(N, 1.0, 1.0, 1.0, a.max_lag, a.data_size, total - 1, a.chunk, pipeline_depth=a.max_lag, port=port)
        master.start(120.0)
    rdv.barrier()
    pinned = [torch.randn(a.data_size).to(dtype).pin_memory() for _ in range(2)]
    staging = [torch.empty(a.data_size, dtype=dtype, device=dev) for _ in range(a.max_lag + 2)]
    host_counts = None
    host_head = torch.empty(1024, dtype=dtype).pin_memory()
    copy_stream = torch.cuda.Stream(device=dev)
    read_stream = torch.cuda.Stream(device=dev)
    marks = {}
    bytes_in = nbytes
    state = {"d2h": 0}

    def source(req):
        x = staging[req.iteration % len(staging)]
        with torch.cuda.stream(copy_stream):
            x.copy_(pinned[req.iteration % 2], non_blocking=True)
        torch.cuda.current_stream(dev).wait_stream(copy_stream)
        return AllReduceInput(x)

    def sink(out):
        nonlocal host_counts
        if host_counts is None:
            host_counts = torch.empty(out.count.numel(), dtype=torch.int32).pin_memory()
        # The worker calls the sink after the host has SEEN the round flushed (completion word written by the round's last CTA
        # behind a system fence), so the result may be read on any stream.  A stream of its own: on the current stream the
        # read-back would queue behind the NEXT round's host -> device copy, which the source has already enqueued there.
        with torch.cuda.stream(read_stream):
            host_counts.copy_(out.count, non_blocking=True)
            host_head.copy_(out.data[:1024], non_blocking=True)
        read_stream.synchronize()
        state["d2h"] = host_counts.numel() * 4 + host_head.numel() * host_head.element_size()
        if int(host_counts.min()) != N:
            raise RuntimeError(f"round {out.iteration}: contributor counts {host_counts.tolist()} != {N}")
        if out.iteration == W - 1 or (W == 0 and out.iteration == 0):
            marks["t0"] = time.perf_counter()
        if out.iteration == total - 1:
            marks["t1"] = time.perf_counter()

    info = {}
    worker = AllreduceWorker(source, sink, backend="cuda", device=local_rank, dtype=a.dtype, algo=a.algo, grid=a.grid,
                             session=f"e2e_{port}_{os.getppid() if world > 1 else os.getpid()}",
                             on_init=lambda w: info.update(algo=w.engine.algo, rounds_in_flight=2 if w.engine.overlap else 1))
    ok = worker.run("127.0.0.1", port, timeout_sec=120.0)
    if master:
        master.join(30)
    worker.close()
    global_sync()
    if not ok or "t1" not in marks or "t0" not in marks:
        return {"error": f"e2e did not finish (shutdown={ok}, delivered={worker.delivered})"}
    steps = K if W > 0 else K - 1
    sec = max_over_ranks(marks["t1"] - marks["t0"])
    ms = sec * 1e3 / steps

    # ---- anchors for the end-to-end number, same buffers, same copy stream, host-timed the same way
    # (1) floor: the host -> device copy of the step's input alone (what no allreduce implementation can go below)
    def timed_loop(step):
        for i in range(W):
            step(i)
        torch.cuda.synchronize(dev)
        rdv.barrier()
        t0 = time.perf_counter()
        for i in range(W, W + K):
            step(i)
        torch.cuda.synchronize(dev)
        return max_over_ranks(time.perf_counter() - t0) * 1e3 / K

    def h2d_only(i):
        x = staging[i % len(staging)]
        with torch.cuda.stream(copy_stream):
            x.copy_(pinned[i % 2], non_blocking=True)
        torch.cuda.current_stream(dev).wait_stream(copy_stream)
        host_head.copy_(x[:1024], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()

    floor_ms = timed_loop(h2d_only)
    # (2) NCCL with the identical source and sink: copy in, all_reduce, read the same slice back, once per step
    nccl_ms = None
    if N > 1 and not a.no_nccl:
        try:
            import torch.distributed as dist

            if dist.is_initialized():
                def nccl_step(i):
                    x = staging[i % len(staging)]
                    with torch.cuda.stream(copy_stream):
                        x.copy_(pinned[i % 2], non_blocking=True)
                    torch.cuda.current_stream(dev).wait_stream(copy_stream)
                    dist.all_reduce(x)
                    host_head.copy_(x[:1024], non_blocking=True)
                    torch.cuda.current_stream(dev).synchronize()

                nccl_ms = timed_loop(nccl_step)
        except Exception:
            nccl_ms = None
    return {"value": N * nbytes / (ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms, "rounds_per_sec": 1e3 / ms,
            "h2d_bytes_per_step": bytes_in, "d2h_bytes_per_step": state["d2h"],
            "path": "AllreduceMaster(pipeline_depth=maxLag) + AllreduceWorker(backend=cuda); source: pinned host -> device copy of the whole "
                    "input per round (copy stream); sink: contributor counts + a 1024-element slice read back per round on a stream of its own, "
                    "synchronised every round (the reduced vector itself stays on the device, where the next consumer is)", "algo": info.get("algo", a.algo),
            "rounds_in_flight": info.get("rounds_in_flight", 1),
            "h2d_floor_ms_per_step": floor_ms, "vs_h2d_floor": ms / floor_ms if floor_ms else None,
            "nccl_same_source_sink_ms_per_step": nccl_ms, "vs_nccl": (nccl_ms / ms) if nccl_ms else None}


if __name__ == "__main__":
    sys.exit(main())
