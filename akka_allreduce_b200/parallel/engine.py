"""One rank's handle on the device data plane: launches the round kernels on a CUDA stream, owns the L-deep
ring of output / count buffers, the host doorbell (newest round announced by the master) and the host-visible
completion word.

This is the part of AllreduceWorker.scala:84-104 (StartAllreduce -> fetch -> scatter) and :253-268 (complete) that
touches data, with the actor mailbox replaced by stream order: rounds of one rank run in order on its stream, ranks
are decoupled from each other by the thresholds, the device-resident ring and the flow-control words.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Sequence

import torch

from .. import _core as core
from ..ops import lib as L
from .local_gpu import _TORCH2AAR

# Measured crossover (profiles/): below this many bytes the one-hop one-shot kernel wins over two-shot.
# AAR_ONESHOT_MAX_BYTES overrides it (crossover sweeps; every rank must use the same value, it sizes the layout).
ONESHOT_MAX_BYTES = int(os.environ.get("AAR_ONESHOT_MAX_BYTES", 256 * 1024))


_DONE_WORD0 = 16   # host-mapped words: [0] doorbell, [_DONE_WORD0 + slot] completion per ring slot
_ALGO_CODE = {"twoshot": 0, "pipelined": 1, "oneshot": 2, "nvls": 3, "bulk": 4, "nvls_bulk": 5, "nvls_onehop": 6}
_IN_SWITCH = ("nvls", "nvls_bulk", "nvls_onehop")
_TABLE = None


def selector_table() -> dict:
    """The measured selection table (written by bench/sweep.py --write-table on the 8-GPU box, committed next to this file):
    {"<P>": {"fp32" | "bf16": [[max_bytes, algo, split], ...]}} -- first row whose max_bytes covers the message wins."""
    global _TABLE
    if _TABLE is None:
        import json
        from pathlib import Path

        path = Path(os.environ.get("AAR_SELECTOR_TABLE", Path(__file__).with_name("selector_table.json")))
        try:
            _TABLE = json.loads(path.read_text())
        except (OSError, ValueError):
            _TABLE = {}
    return _TABLE


def pick_algo(P: int, elem_size: int, nbytes: int, *, bulk_ok: bool, nvls_ok: bool, oneshot_ok: bool, normalize: bool = False,
              table: dict | None = None) -> tuple[str, int]:
    """`algo="auto"` as a pure function: (kernel family, role split) for a message of `nbytes` at world size P.  First row of the
    measured table that covers the size AND whose kernel this layout / these thresholds can run; without a usable row, small
    messages take the one-hop kernel and the rest the bulk engine kernels.  bulk_ok: 16-byte aligned geometry; nvls_ok: multicast
    mapping + staging + full thresholds + no exact-quorum policy (the switch sums every bound GPU); oneshot_ok: the layout has
    one-shot buffers."""
    if P == 1:
        return ("bulk" if bulk_ok else "twoshot"), 0
    rows = (selector_table() if table is None else table).get(str(P), {}).get("fp32" if elem_size == 4 else "bf16", [])
    for max_bytes, name, split in rows:
        if nbytes > max_bytes:
            continue
        if name == "oneshot" and not oneshot_ok:
            continue
        if (name in ("nvls_bulk", "nvls_onehop") and not nvls_ok) or (name == "nvls" and not (nvls_ok and not normalize)):
            continue
        if name in ("bulk", "nvls_bulk", "nvls_onehop") and not bulk_ok:
            continue
        return name, int(split)
    if oneshot_ok and nbytes <= ONESHOT_MAX_BYTES:
        return "oneshot", 0
    if nvls_ok and P >= 8:
        return "nvls_bulk", 0
    return ("bulk" if bulk_ok else "pipelined"), 0


class RoundEngine:
    def __init__(self, rank: int, region_ptrs: Sequence[int], layout: L.AarLayout, *, th_reduce: float = 1.0, th_complete: float = 1.0,
                 normalize: bool = False, exact_quorum: bool = False, timeout_ms: float = 0.0, device: int = 0, grid: int = 0,
                 algo: str = "auto", mc_ptr: int = 0, keepalive=None, overlap: bool | None = None, max_inflight: int | None = None):
        self.lib = L.load()
        self.rank, self.P = rank, layout.P
        self.layout = layout
        self.device = torch.device("cuda", device)
        self.dtype = torch.float32 if layout.dtype == L.AAR_F32 else torch.bfloat16
        self.grid = grid
        self.keepalive = keepalive
        self.geo = core.Geometry(layout.P, layout.dataSize, layout.maxChunk)
        self.data_size = layout.dataSize
        self.max_lag = layout.L - 1
        self.mc_ptr = mc_ptr
        p = L.AarRankParams()
        p.rank = rank
        p.minRequired = core.Geometry.min_required(th_reduce, layout.P)
        p.minChunksRequired = self.geo.min_chunks_required(th_complete)
        p.maxLag = self.max_lag
        self.full_thresholds = p.minRequired >= layout.P and p.minChunksRequired >= layout.totalChunks
        self.bulk_split = 0          # role split of the bulk kernels (nA | nB << 16, 0 = the library's heuristic)
        self.nvls_bulk_split = 0
        self.algo = self._select(algo, normalize, exact_quorum)
        with torch.cuda.device(self.device):
            self.outputs = [torch.empty(self.data_size, dtype=self.dtype, device=self.device) for _ in range(layout.L)]
            self.counts = [torch.zeros(layout.totalChunks, dtype=torch.int32, device=self.device) for _ in range(layout.L)]
            host = C.POINTER(C.c_uint)()
            dev = C.POINTER(C.c_uint)()
            L.check(self.lib.aar_host_words_alloc(_DONE_WORD0 + layout.L, C.byref(host), C.byref(dev)), "aar_host_words_alloc")
        self._host_words, self._dev_words = host, dev
        p.flags = (L.FLAG_NORMALIZE if normalize else 0) | (L.FLAG_EXACT_QUORUM if exact_quorum else 0)
        # Strict protocol by default (docs/PROTOCOL.md; all variants verified on B200 at the start of round 2).  The bulk kernels
        # are always strict; on the register kernels:
        #   * partial quorum: the reduce-broadcast obeys the scatter's per-destination slot flow control (AAR_STRICT=0 opts out);
        #   * every chunk required: the eager gather takes ONE decision per chunk, so a wait that gives up (timeout, doorbell)
        #     can never leave a chunk half copied.
        strict = os.environ.get("AAR_STRICT", "1") not in ("", "0")
        if strict and not self.full_thresholds:
            p.flags |= L.FLAG_BCAST_FC
        if strict and p.minChunksRequired >= layout.totalChunks:
            p.flags |= L.FLAG_EAGER_DEC
        if os.environ.get("AAR_ELIDE_LATE", "0") not in ("", "0") and not self.full_thresholds:
            p.flags |= L.FLAG_ELIDE_LATE | L.FLAG_BCAST_FC   # late-work elision lives in the strict register kernels
        # Bulk kernels: own contribution read in place, own reduced block written straight to the output -- only when there
        # are peers (a world of one keeps the full three-pass loop-back, which is what one rank of a larger world does per round)
        if self.algo == "bulk" and layout.P > 1 and os.environ.get("AAR_SELF_DIRECT", "1") not in ("", "0"):
            p.flags |= L.FLAG_SELF_DIRECT
        p.timeoutNs = int(timeout_ms * 1e6)
        if self.algo == "nvls_onehop" and not self.full_thresholds:
            # in-switch (every rank is summed), and its give-up is decided per CTA, not once per round (docs/PROTOCOL.md): only
            # where no round is ever forced in normal operation.  A spin timeout stays allowed as a liveness guard -- a round that
            # runs into it sets the status bits, so its output is flagged whatever it contains.
            raise ValueError("algo=nvls_onehop is offered for thresholds 1.0 only")
        dev_addr = C.cast(dev, C.c_void_p).value
        p.doorbell = dev_addr            # word 0: newest round announced by the master
        p.hostDone = dev_addr + 4 * _DONE_WORD0   # one word per ring slot: round + 1 of the last round flushed in that slot
        self._in_order = 0               # rounds known flushed in order (host-side scan of the slot words)
        for j in range(layout.P):
            p.region[j] = region_ptrs[j]
        p.mcRegion = mc_ptr
        self.params = p
        self.base_flags = p.flags
        self._lay = C.byref(self.layout)
        self._pref = C.byref(self.params)
        self.rounds_enqueued = 0
        self.launches = 0
        self.fallback_launches = 0
        self._nvtx = os.environ.get("AAR_NVTX", "0") not in ("", "0")   # one NVTX range per round launch (off by default)
        self._slot_round = [-1] * layout.L
        self._bounce: dict[int, torch.Tensor] = {}
        # CTA role split of the register pipelined kernels (measured on 8xB200 in round 1)
        self.split = 32 | (48 << 16)      # nPush | nReduce << 16; the rest gather
        self.nvls_reduce_ctas = 8
        self.nvls_stage_ctas = 48
        # The maxLag window lets two consecutive rounds of one worker be in flight: they alternate between two internal
        # streams, so the push of round r+1 starts on the SMs that round r's push CTAs have already left while round r is
        # still reducing and gathering.  Needs two ring slots (maxLag >= 1) and peers to overlap with.
        self.overlap = (layout.L >= 2 and layout.P > 1) if overlap is None else (overlap and layout.L >= 2)
        # rounds of this rank in flight at once: 2 by default; a deeper ring (maxLag > 1) may use more, which is what lets a rank
        # whose round r starts late keep up with rounds r+1 .. r+maxLag meanwhile (straggler mode)
        if max_inflight is None:
            # measured, 8 GPUs, thresholds 0.75, maxLag 4, one rank 500 us late per round: 93 % of the straggler-free rate with
            # four rounds in flight against 84 % with two (profiles/r2_straggler_8gpu_*.json).  Only with a partial quorum: at
            # thresholds 1.0 every round waits for the late rank and deeper overlap buys nothing.
            max_inflight = 4 if not self.full_thresholds else 2
        self.inflight = max(2, min(int(max_inflight), layout.L, 8)) if self.overlap else 1
        if (self.inflight > 2 or (self.algo == "nvls_onehop" and self.overlap)) and grid == 0:
            # every round in flight must be able to hold ALL its CTAs at once: a kernel that waits for peers while it occupies the
            # SMs an earlier round of this rank still needs is a resource deadlock (seen at 4 in flight with full-size grids)
            sms = C.c_int(0)
            L.check(self.lib.aar_device_sm_count(device, C.byref(sms)), "aar_device_sm_count")
            self.grid = grid = max(3, sms.value // self.inflight)
        # The per-round host work -- order after the caller's stream, launch, record the slot's event -- is ONE C call
        # (csrc/cuda/aar_engine.cu); AAR_NATIVE_ENGINE=0 keeps the Python choreography (NVTX ranges, explicit streams).
        self._native = None
        with torch.cuda.device(self.device):
            prio = 0 if os.environ.get("AAR_STREAM_PRIORITY", "high") in ("default", "0") else -1   # as csrc/cuda/aar_engine.cu
            self._streams = [torch.cuda.Stream(device=self.device, priority=prio) for _ in range(self.inflight)] if self.overlap else []
            # one event per round in a ring of L + streams entries (see aar_engine.cu: what a later round or a consumer waits for)
            self._ring = layout.L + max(1, self.inflight if self.overlap else 1)
            self._events = [torch.cuda.Event() for _ in range(self._ring)]
            self._event_round = [-1] * self._ring
            if os.environ.get("AAR_NATIVE_ENGINE", "1") not in ("", "0") and not self._nvtx:
                regs = (C.c_void_p * layout.P)(*[region_ptrs[j] for j in range(layout.P)])
                h = C.c_void_p(0)
                L.check(self.lib.aar_engine_create(self._lay, rank, regs, C.c_void_p(mc_ptr or None), p.minRequired, p.minChunksRequired,
                                                   self.base_flags, p.timeoutNs, _ALGO_CODE[self.algo], grid, self.inflight if self.overlap else 0, C.byref(h)), "aar_engine_create")
                self._native = h
        self._native_flags = self.base_flags   # what the native engine was created with; later additions travel per round
        self._native_tuning = None

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def single_process(cls, data_size: int, max_chunk: int, max_lag: int = 1, dtype: torch.dtype = torch.float32, device: int = 0, **kw):
        """A world of one rank (N=1 runs and the smoke test): the region is a plain device allocation."""
        lay = L.compute_layout(1, data_size, max_chunk, max_lag, _TORCH2AAR[dtype], ONESHOT_MAX_BYTES // (4 if dtype == torch.float32 else 2), False)
        with torch.cuda.device(device):
            region = torch.zeros(lay.regionBytes, dtype=torch.uint8, device=f"cuda:{device}")
        return cls(0, [region.data_ptr()], lay, device=device, keepalive=region, **kw)

    @classmethod
    def from_symmetric(cls, symm, layout: L.AarLayout, **kw):
        return cls(symm.rank, symm.ptrs, layout, device=symm.device, mc_ptr=symm.mc_ptr, keepalive=symm, **kw)

    def _bulk_geometry(self) -> bool:
        """The bulk kernels move 16-byte aligned pieces: blocks, chunks and the vector length must be multiples of 16 bytes."""
        lay = self.layout
        V = 16 // lay.elemSize
        return lay.stepSize % V == 0 and lay.maxChunk % V == 0 and lay.dataSize % V == 0 and lay.blkStride % V == 0

    def _select(self, algo: str, normalize: bool, exact_quorum: bool) -> str:
        """auto: from the measured table (world size, dtype, bytes) -> kernel family + role split; thresholds < 1, normalisation
        and the exact-quorum policy rule the in-switch path out (the switch sums every bound GPU)."""
        lay = self.layout
        nbytes = lay.dataSize * lay.elemSize
        bulk_ok = self._bulk_geometry()
        nvls_ok = bool(self.mc_ptr and lay.stageOff and bulk_ok and self.full_thresholds and not exact_quorum)   # normalisation: sum / P in registers
        if algo == "auto":
            name, split = pick_algo(lay.P, lay.elemSize, nbytes, bulk_ok=bulk_ok, nvls_ok=nvls_ok, oneshot_ok=bool(lay.oneshotOff),
                                    normalize=normalize)
            if name == "bulk":
                self.bulk_split = split
            elif name == "nvls_bulk":
                self.nvls_bulk_split = split
            return name
        if algo == "oneshot" and not lay.oneshotOff:
            raise ValueError("algo=oneshot needs a layout with one-shot buffers")
        if algo in _IN_SWITCH and not (self.mc_ptr and lay.stageOff):
            raise ValueError(f"algo={algo} needs a multicast mapping and the staging area")
        if algo in ("bulk", "nvls_bulk", "nvls_onehop") and not bulk_ok:
            raise ValueError(f"algo={algo} needs blocks and chunks that are multiples of 16 bytes")
        if algo not in _ALGO_CODE:
            raise ValueError(f"unknown algo {algo}")
        return algo

    # ------------------------------------------------------------------ rounds
    def set_doorbell(self, newest_round: int) -> None:
        """Host side of StartAllreduce: kernels that wait compare this with their round (catch-up when > maxLag behind)."""
        if self._native is not None:
            self.lib.aar_engine_set_doorbell(self._native, int(newest_round))
            return
        self._host_words[0] = max(int(self._host_words[0]), newest_round) if newest_round >= 0 else 0

    def round_flushed(self, rnd: int) -> bool:
        """Has round `rnd` been flushed (its last CTA wrote the slot's completion word)?  Rounds may complete out of order: the
        device retires per ring slot, so round r+1 can be here before round r (AllreduceWorker.scala:253-268)."""
        if self._native is not None:
            return bool(self.lib.aar_engine_round_flushed(self._native, int(rnd)))
        return int(self._host_words[_DONE_WORD0 + rnd % self.layout.L]) >= rnd + 1

    def rounds_flushed(self) -> int:
        """Rounds flushed IN ORDER from the start: every round below the returned number is done."""
        if self._native is not None:
            return int(self.lib.aar_engine_rounds_flushed(self._native))
        while self.round_flushed(self._in_order):
            self._in_order += 1
        return self._in_order

    def prepare_round(self, rnd: int, x: torch.Tensor, out: torch.Tensor | None = None):
        """Everything enqueue_round does before the launch: validate the input, pick the output, bounce an unaligned view,
        forward changed tuning knobs.  Returns (x, output) -- the tensors the launch will read and write."""
        if x.dtype != self.dtype or x.numel() != self.data_size or not x.is_contiguous() or x.device != self.device:
            raise ValueError(f"input must be a contiguous {self.dtype} tensor of {self.data_size} elements on {self.device} "
                             f"(AllreduceWorker.scala:174-176 rejects a size that differs from initialisation)")
        slot = rnd % self.layout.L
        output = self.outputs[slot] if out is None else out
        if self.algo == "oneshot" and output.data_ptr() == x.data_ptr():
            # the one-shot kernel has no barrier between its push and its reduce: a CTA may store output tile t while another
            # CTA of this rank has not pushed that part of the input to a slower peer yet
            raise ValueError("in-place rounds (out aliasing the input) are not supported by the one-shot kernel")
        if self.algo in ("bulk", "nvls_bulk", "nvls_onehop") and x.data_ptr() % 16:
            # the bulk copy engine needs 16-byte aligned addresses: bounce an oddly offset view through an aligned buffer
            b = self._bounce.get(slot)
            if b is None:
                b = self._bounce[slot] = torch.empty_like(self.outputs[slot])
            b.copy_(x)
            x = b
        if self._native is not None:
            tuning = (self.split, self.nvls_reduce_ctas | (self.nvls_stage_ctas << 8), self.nvls_bulk_split if self.algo in ("nvls_bulk", "nvls_onehop") else self.bulk_split)
            if tuning != self._native_tuning:   # knobs changed since the last round (benchmarks sweep them)
                L.check(self.lib.aar_engine_tune(self._native, tuning[0], tuning[1]), "aar_engine_tune")
                L.check(self.lib.aar_engine_tune_bulk(self._native, tuning[2]), "aar_engine_tune_bulk")
                self._native_tuning = tuning
        return x, output

    def note_launched(self, rnd: int) -> None:
        """Bookkeeping after a launch (also called by the C++ round loop's source adapter, which launches through the C entry
        point itself)."""
        self._slot_round[rnd % self.layout.L] = rnd
        self.launches += 1
        self.rounds_enqueued = max(self.rounds_enqueued, rnd + 1)

    def native_plane(self):
        """The data plane as the C++ round loop wants it (csrc/core/device_loop.h): engine handle + the addresses of the four C
        entry points.  None on the Python launch path."""
        if self._native is None:
            return None
        from .. import _core as core

        def addr(fn):
            return C.cast(fn, C.c_void_p).value

        handle = self._native.value if hasattr(self._native, "value") else int(self._native)
        return core.DevicePlane(engine=handle, enqueue=addr(self.lib.aar_engine_enqueue_ex), doorbell=addr(self.lib.aar_engine_set_doorbell),
                                flushed=addr(self.lib.aar_engine_round_flushed), synchronize=addr(self.lib.aar_engine_synchronize),
                                slots=self.layout.L, max_lag=self.max_lag, data_size=self.data_size,
                                extra_flags=self.base_flags & ~self._native_flags)

    def enqueue_round(self, rnd: int, x: torch.Tensor, *, stream: torch.cuda.Stream | None = None, force: bool = False,
                      skip_scatter: bool = False, out: torch.Tensor | None = None, delay_ns: int = 0):
        """Launch round `rnd` with input `x`; returns (output, counts) tensors that are valid after `wait_round(rnd)` (or once
        the explicit `stream` reaches the end of the round) and are reused maxLag+1 rounds later.  `x` must stay alive
        and unmodified until then."""
        x, output = self.prepare_round(rnd, x, out)
        slot = rnd % self.layout.L
        if self._native is not None:
            if stream is not None:
                raise NotImplementedError("the native engine picks its own stream (AAR_NATIVE_ENGINE=0 for explicit streams)")
            extra = (L.FLAG_FORCE if force else 0) | (L.FLAG_SKIP_SCATTER if skip_scatter else 0) | (self.base_flags & ~self._native_flags)
            L.check(self.lib.aar_engine_enqueue_ex(self._native, rnd, C.c_void_p(x.data_ptr()), C.c_void_p(output.data_ptr()),
                                                   C.c_void_p(self.counts[slot].data_ptr()), extra,
                                                   C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream), int(delay_ns)), "aar_engine_enqueue")
            self.note_launched(rnd)
            return output, self.counts[slot]
        p = self.params
        p.round = rnd
        p.input = x.data_ptr()
        p.output = output.data_ptr()
        p.outCounts = self.counts[slot].data_ptr()
        p.flags = self.base_flags | (L.FLAG_FORCE if force else 0) | (L.FLAG_SKIP_SCATTER if skip_scatter else 0)
        cur = torch.cuda.current_stream(self.device)
        if stream is None and self.overlap:
            stream = self._streams[rnd % self.inflight]
            stream.wait_stream(cur)          # x was produced on the caller's stream
            p.flags |= L.FLAG_OVERLAPPED
            # every round <= rnd - L of this rank must be done before this one starts (retirement is per ring slot: nothing else
            # keeps one stream from running ahead of the other, and the catch-up predicate relies on the window)
            for k in range(1, self.inflight):
                q = rnd - self.layout.L
                while q >= 0 and q % self.inflight != (rnd + k) % self.inflight:
                    q -= 1
                if q >= 0 and self._event_round[q % self._ring] == q:
                    stream.wait_event(self._events[q % self._ring])
        run_stream = stream or cur
        q = rnd - self.layout.L           # the output / count buffers of this slot belong to round rnd - L until it is done
        if q >= 0 and self._event_round[q % self._ring] == q:
            run_stream.wait_event(self._events[q % self._ring])
        s = C.c_void_p(run_stream.cuda_stream)
        if self._nvtx:
            torch.cuda.nvtx.range_push(f"aar.round{rnd}.{self.algo}.slot{slot}")
        if delay_ns > 0:  # straggler injection: this rank's round starts late
            L.check(self.lib.aar_launch_delay(int(delay_ns), s), "aar_launch_delay")
        if self.algo == "twoshot":
            L.check(self.lib.aar_launch_round_fused(self._pref, self._lay, self.grid, s), "aar_launch_round_fused")
        elif self.algo == "pipelined":
            L.check(self.lib.aar_launch_round_pipelined(self._pref, self._lay, self.grid, self.split, s), "aar_launch_round_pipelined")
        elif self.algo == "oneshot":
            L.check(self.lib.aar_launch_oneshot(self._pref, self._lay, 2, self.grid, s), "aar_launch_oneshot")
        elif self.algo in ("bulk", "nvls_bulk", "nvls_onehop"):
            which = {"bulk": 0, "nvls_bulk": 1, "nvls_onehop": 2}[self.algo]
            if self.lib.aar_bulk_supported(self._pref, self._lay, which):
                L.check(self.lib.aar_launch_round_bulk(self._pref, self._lay, which, 7, self.grid, self.nvls_bulk_split if which else self.bulk_split, s),
                        "aar_launch_round_bulk")
            elif which == 0:   # unaligned input / output pointer or geometry: the register kernels take anything
                p.flags &= ~L.FLAG_SELF_DIRECT
                if self.layout.P > 1:
                    L.check(self.lib.aar_launch_round_pipelined(self._pref, self._lay, self.grid, self.split, s), "aar_launch_round_pipelined")
                else:
                    L.check(self.lib.aar_launch_round_fused(self._pref, self._lay, self.grid, s), "aar_launch_round_fused")
                self.fallback_launches = getattr(self, "fallback_launches", 0) + 1
            else:
                raise L.AarError("algo=nvls_bulk: " + (self.lib.aar_last_error() or b"").decode())
        elif self.algo == "nvls":
            L.check(self.lib.aar_launch_nvls(self._pref, self._lay, 7 | (self.nvls_reduce_ctas << 8) | (self.nvls_stage_ctas << 16), self.grid, s), "aar_launch_nvls")
        else:
            raise ValueError(f"unknown algo {self.algo}")
        self._events[rnd % self._ring].record(run_stream)
        self._event_round[rnd % self._ring] = rnd
        if self._nvtx:
            torch.cuda.nvtx.range_pop()
        self.launches += 1
        self.rounds_enqueued = max(self.rounds_enqueued, rnd + 1)
        return output, self.counts[slot]

    def stage_input(self, rnd: int, stream: torch.cuda.Stream | None = None) -> torch.Tensor | None:
        """Zero-copy input for the in-switch path: the slot of this rank's multicast-bound staging area that round `rnd` will
        reduce from.  A source that writes its data here (H2D copy, the producing kernel's output, ...) and then passes the
        SAME tensor to `enqueue_round(rnd, x)` saves the kernel's 1x read + 1x write staging pass (the stage role then only
        publishes the arrival flags).  Returns None when the engine has no staging area (P2P paths copy from anywhere).
        The slot was last read by the peers' reduce of round rnd - L; every owner had finished that before this rank's round
        rnd - L could retire, so ordering after that round's event is enough: `stream` (default: current) is made to wait."""
        lay = self.layout
        if self.algo not in _IN_SWITCH or not lay.stageOff:
            return None
        if rnd >= lay.L:
            self.wait_round(rnd - lay.L, stream)
        V = 16 // lay.elemSize
        padded = (lay.dataSize + V - 1) // V * V
        off = lay.stageOff + (rnd % lay.L) * padded * lay.elemSize
        return self._region_view(off, lay.dataSize * lay.elemSize, self.dtype)

    def wait_round(self, rnd: int, stream: torch.cuda.Stream | None = None) -> None:
        """Make `stream` (default: the current stream) wait for round `rnd`; call before consuming its output on a stream."""
        if self._native is not None:
            st = stream or torch.cuda.current_stream(self.device)
            L.check(self.lib.aar_engine_wait(self._native, rnd, C.c_void_p(st.cuda_stream)), "aar_engine_wait")
            return
        if self._event_round[rnd % self._ring] >= rnd:   # == rnd: its event; > rnd: a later round launched behind its completion
            (stream or torch.cuda.current_stream(self.device)).wait_event(self._events[rnd % self._ring])

    def synchronize(self) -> None:
        """Host waits until every launched round of this rank has finished."""
        if self._native is not None:
            L.check(self.lib.aar_engine_synchronize(self._native), "aar_engine_synchronize")
            return
        for ev, r in zip(self._events, self._event_round):
            if r >= 0:
                ev.synchronize()

    def wait_all(self, stream: torch.cuda.Stream | None = None) -> None:
        st = stream or torch.cuda.current_stream(self.device)
        if self._native is not None:
            for r in self._slot_round:
                if r >= 0:
                    L.check(self.lib.aar_engine_wait(self._native, r, C.c_void_p(st.cuda_stream)), "aar_engine_wait")
            return
        for ev, r in zip(self._events, self._event_round):
            if r >= 0:
                st.wait_event(ev)

    def allreduce(self, x: torch.Tensor, rnd: int | None = None) -> torch.Tensor:
        """Convenience: one round, result tensor (stream-ordered)."""
        rnd = self.rounds_enqueued if rnd is None else rnd
        out, _ = self.enqueue_round(rnd, x)
        self.wait_round(rnd)
        return out

    def clear_status(self) -> None:
        """Zero the sticky status bits of this rank's region (benchmarks that reuse one region for several engines)."""
        torch.cuda.synchronize(self.device)
        self._region_view(self.layout.statusOff, 4, torch.int32).zero_()
        torch.cuda.synchronize(self.device)

    def status(self) -> int:
        """Device status bits (timeouts, dropped pushes) of this rank; synchronises."""
        torch.cuda.synchronize(self.device)
        t = torch.empty(2, dtype=torch.int32, device=self.device)
        holder = type("H", (), {})()
        holder.__cuda_array_interface__ = {"shape": (2,), "typestr": "<i4", "data": (self.params.region[self.rank] + self.layout.statusOff, False), "version": 3}
        t.copy_(torch.as_tensor(holder, device=self.device))
        return int(t[0].item())

    # ------------------------------------------------------------------ checkpoint of the device-resident ring
    def _region_view(self, off: int, nbytes: int, dtype: torch.dtype) -> torch.Tensor:
        holder = type("H", (), {})()
        holder.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (self.params.region[self.rank] + off, False), "version": 3}
        return torch.as_tensor(holder, device=self.device).view(dtype)

    def snapshot(self) -> "core.RingSnapshot":
        """Host copy of this rank's ring slots and flag words (layout of SURVEY.md section 5: temporalBuffer[L][P][block] +
        arrival words per slot), to be written with `_core.save_ring_snapshot`.  Call with no round in flight."""
        torch.cuda.synchronize(self.device)
        lay = self.layout
        ring_bytes = lay.L * lay.P * lay.blkStride * lay.elemSize
        s = core.RingSnapshot()
        s.rank, s.P, s.L, s.data_size, s.max_chunk_size = self.rank, lay.P, lay.L, lay.dataSize, lay.maxChunk
        s.round = self.rounds_enqueued
        to_f32 = (lambda t: t.float()) if self.dtype == torch.bfloat16 else (lambda t: t)
        s.scatter_ring = to_f32(self._region_view(lay.scatterOff, ring_bytes, self.dtype)).cpu().tolist() if ring_bytes <= (64 << 20) else []
        s.reduce_ring = to_f32(self._region_view(lay.reduceOff, ring_bytes, self.dtype)).cpu().tolist() if ring_bytes <= (64 << 20) else []
        (s_off, s_bytes), (r_off, r_bytes) = self._flag_sections()
        s.scatter_flags = [v & 0xFFFFFFFF for v in self._region_view(s_off, s_bytes, torch.int32).cpu().tolist()]
        s.reduce_flags = [v & 0xFFFFFFFFFFFFFFFF for v in self._region_view(r_off, r_bytes, torch.int64).cpu().tolist()]
        return s

    def _flag_sections(self):
        """Where this engine's kernels keep their arrival words: per chunk (register kernels) or per piece (bulk kernels) -- the
        "countFilled" of the reference's DataBuffer (DataBuffer.scala:27), saved and restored with the ring."""
        lay = self.layout
        if self.algo in ("bulk", "nvls_bulk"):
            return ((lay.pflagSOff, lay.L * lay.C * lay.piecesPerChunk * L.AAR_MAXP * 4), (lay.pflagROff, lay.L * lay.P * lay.C * lay.piecesPerChunk * 8))
        return ((lay.sflagOff, lay.L * lay.C * L.AAR_MAXP * 4), (lay.rflagOff, lay.L * lay.P * lay.C * 8))

    def start_at(self, rnd: int) -> None:
        """Begin the round numbering at `rnd` on a fresh engine (a master that resumed from a checkpoint): every rank counts as
        having retired all earlier rounds, so the slot flow control of round `rnd` has nothing to wait for.  Every rank calls
        this on its own region before the first launch (the worker does, between mapping the regions and its first round)."""
        self._mark_retired_below(rnd)
        self.rounds_enqueued = rnd
        torch.cuda.synchronize(self.device)

    def _mark_retired_below(self, rnd: int) -> None:
        """Every round below `rnd` counts as retired by every rank: the newest-round words, the per-slot retirement words (what
        the slot flow control reads) and the host's completion words."""
        lay = self.layout
        self._region_view(lay.peerBaseOff, L.AAR_MAXP * 4, torch.int32)[: lay.P] = rnd
        slot_base = self._region_view(lay.slotBaseOff, lay.L * L.AAR_MAXP * 4, torch.int32).view(lay.L, L.AAR_MAXP)
        for r in range(max(0, rnd - lay.L), rnd):
            slot_base[r % lay.L, : lay.P] = r + 1
            self._host_words[_DONE_WORD0 + r % lay.L] = r + 1
        self._region_view(lay.statusOff, 8, torch.int32)[1] = rnd
        self._in_order = rnd
        if self._native is not None:
            L.check(self.lib.aar_engine_start_at(self._native, rnd), "aar_engine_start_at")

    def restore(self, s: "core.RingSnapshot") -> None:
        """Inverse of snapshot(): the ring payload and the arrival words go back into this rank's region; round numbering
        continues at s.round.  The geometry must match."""
        lay = self.layout
        if (s.P, s.L, s.data_size, s.max_chunk_size) != (lay.P, lay.L, lay.dataSize, lay.maxChunk):
            raise ValueError("ring snapshot was taken with a different geometry")
        ring_bytes = lay.L * lay.P * lay.blkStride * lay.elemSize
        if s.scatter_ring:
            self._region_view(lay.scatterOff, ring_bytes, self.dtype).copy_(torch.tensor(s.scatter_ring, dtype=torch.float32).to(self.dtype))
            self._region_view(lay.reduceOff, ring_bytes, self.dtype).copy_(torch.tensor(s.reduce_ring, dtype=torch.float32).to(self.dtype))
        (s_off, s_bytes), (r_off, r_bytes) = self._flag_sections()
        if len(s.scatter_flags) * 4 != s_bytes or len(s.reduce_flags) * 8 != r_bytes:
            raise ValueError("ring snapshot was taken by a different kernel family (per-chunk vs per-piece arrival words)")
        self._region_view(s_off, s_bytes, torch.int32).copy_(
            torch.tensor([v - (1 << 32) if v >= (1 << 31) else v for v in s.scatter_flags], dtype=torch.int32))
        self._region_view(r_off, r_bytes, torch.int64).copy_(
            torch.tensor([v - (1 << 64) if v >= (1 << 63) else v for v in s.reduce_flags], dtype=torch.int64))
        # A snapshot is taken with no round in flight, so every peer had retired every round below s.round: restore the
        # flow-control words accordingly (round s.round would otherwise wait for retirements that happened before the save).
        self._mark_retired_below(s.round)
        self.rounds_enqueued = s.round
        torch.cuda.synchronize(self.device)

    def metrics(self) -> dict:
        """Structured counters of this rank's data plane (the reference only has log lines)."""
        st = self.status()
        return {"algo": self.algo, "rounds_enqueued": self.rounds_enqueued, "rounds_flushed": self.rounds_flushed(), "launches": self.launches,
                "timeout_scatter": bool(st & L.ST_TIMEOUT_SCATTER), "timeout_reduce": bool(st & L.ST_TIMEOUT_REDUCE),
                "timeout_gather": bool(st & L.ST_TIMEOUT_GATHER), "push_dropped": bool(st & L.ST_PUSH_DROPPED),
                "rounds_in_flight": self.inflight, "multicast": bool(self.mc_ptr)}

    def close(self) -> None:
        if self._native is not None:
            self.lib.aar_engine_destroy(self._native)
            self._native = None
        if self._host_words:
            self.lib.aar_host_words_free(self._host_words)
            self._host_words = None
