"""Single-device cluster emulation: P ranks whose symmetric regions all live on one GPU.

Purpose: exercise every kernel and the whole threshold / ring / catch-up logic on ONE device, in one
process, deterministically.  Kernels of different ranks wait on each other through flags, so on a
single GPU they are launched in phase order (all K1, then all K2, then all K3) -- no launch ever waits
for a later one (B200_PROFILING.md: never run mutually-waiting kernels as separate launches on one GPU).
This mirrors how the reference tests one worker against a fake cluster (AllreduceSpec.scala:757-763),
except that all P ranks are real here.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Sequence

import torch

from .. import _core as core
from ..ops import lib as L

_TORCH2AAR = {torch.float32: L.AAR_F32, torch.bfloat16: L.AAR_BF16}


class LocalGpuCluster:
    def __init__(self, P: int, data_size: int, max_chunk: int, max_lag: int = 1, th_reduce: float = 1.0, th_complete: float = 1.0,
                 dtype: torch.dtype = torch.float32, normalize: bool = False, exact_quorum: bool = False, timeout_ms: float = 2000.0,
                 device: str | torch.device = "cuda:0", grid: int = 0, oneshot: bool = False, bcast_fc: bool = False, eager_dec: bool = False,
                 elide_late: bool = False, impl: str = "reg", self_direct: bool = False):
        self.lib = L.load()
        self.P, self.data_size, self.max_chunk, self.max_lag = P, data_size, max_chunk, max_lag
        self.dtype = dtype
        self.device = torch.device(device)
        self.grid = grid
        self.geo = core.Geometry(P, data_size, max_chunk)
        self.min_required = core.Geometry.min_required(th_reduce, P)
        self.min_chunks_required = self.geo.min_chunks_required(th_complete)
        self.base_flags = (L.FLAG_NORMALIZE if normalize else 0) | (L.FLAG_EXACT_QUORUM if exact_quorum else 0) | (L.FLAG_BCAST_FC if bcast_fc else 0) | (L.FLAG_EAGER_DEC if eager_dec else 0) | ((L.FLAG_ELIDE_LATE | L.FLAG_BCAST_FC) if elide_late else 0)
        self.timeout_ns = int(timeout_ms * 1e6)
        # impl = "bulk": the phases run on the bulk async-copy engine kernels (csrc/cuda/aar_bulk.cu) wherever the geometry and the
        # pointers allow it, rank by rank; anything else falls back to the register kernels, exactly as RoundEngine does
        self.impl = impl
        if self_direct:
            self.base_flags |= L.FLAG_SELF_DIRECT
        import os
        self.base_flags |= int(os.environ.get("AAR_DEBUG_FLAGS", "0"), 0)
        self.layout = L.compute_layout(P, data_size, max_chunk, max_lag, _TORCH2AAR[dtype], data_size if oneshot else 0, False)
        with torch.cuda.device(self.device):
            self.regions = [torch.zeros(self.layout.regionBytes, dtype=torch.uint8, device=self.device) for _ in range(P)]
            self.outputs = [torch.empty(data_size, dtype=dtype, device=self.device) for _ in range(P)]
            self.counts = [torch.zeros(self.layout.totalChunks, dtype=torch.int32, device=self.device) for _ in range(P)]
        self.params = (L.AarRankParams * P)()
        for r in range(P):
            p = self.params[r]
            p.rank = r
            p.minRequired = self.min_required
            p.minChunksRequired = self.min_chunks_required
            p.maxLag = max_lag
            p.timeoutNs = self.timeout_ns
            p.output = self.outputs[r].data_ptr()
            p.outCounts = self.counts[r].data_ptr()
            for j in range(P):
                p.region[j] = self.regions[j].data_ptr()

    # ------------------------------------------------------------------ rounds
    def _prep(self, rnd: int, inputs: Sequence[torch.Tensor], force: bool, skip: Iterable[int]):
        skip = set(skip)
        for r in range(self.P):
            x = inputs[r]
            assert x.dtype == self.dtype and x.numel() == self.data_size and x.is_cuda and x.is_contiguous()
            p = self.params[r]
            p.round = rnd
            p.input = x.data_ptr()
            p.flags = self.base_flags | (L.FLAG_FORCE if force else 0) | (L.FLAG_SKIP_SCATTER if r in skip else 0)
        mask = 0
        for r in skip:
            mask |= 1 << r
        return mask

    def _phase(self, r: int, phase: int, stream) -> None:
        """One phase (1 scatter, 2 reduce + broadcast, 4 gather) of rank r on the kernels `impl` selects."""
        lay = C.byref(self.layout)
        pr = C.byref(self.params[r])
        if self.impl == "bulk" and self.lib.aar_bulk_supported(pr, lay, 0):
            L.check(self.lib.aar_launch_round_bulk(pr, lay, 0, phase, self.grid, 0, stream), "aar_launch_round_bulk")
            self.bulk_launches = getattr(self, "bulk_launches", 0) + 1
            return
        fn = {1: self.lib.aar_launch_scatter, 2: self.lib.aar_launch_reduce_bcast, 4: self.lib.aar_launch_gather}[phase]
        L.check(fn(pr, lay, self.grid, stream), "phase launch")

    def run_round(self, rnd: int, inputs: Sequence[torch.Tensor], skip: Iterable[int] = (), force: bool = False):
        """One allreduce round for all ranks: returns (outputs, counts) -- per-rank tensors owned by the cluster."""
        mask = self._prep(rnd, inputs, force, skip)
        if self.impl == "bulk":
            st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            with torch.cuda.device(self.device):
                for phase in (1, 2, 4):  # phase order matters on a single device: no kernel may wait for a later launch
                    for r in range(self.P):
                        if phase == 1 and (mask >> r) & 1:
                            continue
                        self._phase(r, phase, st)
            return self.outputs, self.counts
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self.lib.aar_local_round(self.params, self.P, C.byref(self.layout), mask, self.grid, C.c_void_p(stream)), "aar_local_round")
        return self.outputs, self.counts

    def run_phases(self, rnd: int, inputs: Sequence[torch.Tensor], scatter: Iterable[int] = (), reduce: Iterable[int] = (),
                   gather: Iterable[int] = (), force: bool = False):
        """Launch selected phases for selected ranks (fault-injection tests script arrival orders with this)."""
        self._prep(rnd, inputs, force, ())
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        lay = C.byref(self.layout)
        with torch.cuda.device(self.device):
            for r in scatter:
                self._phase(r, 1, stream)
            for r in reduce:
                self._phase(r, 2, stream)
            for r in gather:
                self._phase(r, 4, stream)

    def run_oneshot(self, rnd: int, inputs: Sequence[torch.Tensor], skip: Iterable[int] = (), force: bool = False):
        skip = set(skip)
        self._prep(rnd, inputs, force, skip)
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        lay = C.byref(self.layout)
        with torch.cuda.device(self.device):
            for r in range(self.P):
                if r not in skip:
                    L.check(self.lib.aar_launch_oneshot(C.byref(self.params[r]), lay, 0, self.grid, stream), "oneshot push")
            for r in range(self.P):
                L.check(self.lib.aar_launch_oneshot(C.byref(self.params[r]), lay, 1, self.grid, stream), "oneshot reduce")
        return self.outputs, self.counts

    # ------------------------------------------------------------------ introspection
    def _section(self, rank: int, off: int, nbytes: int, dtype: torch.dtype) -> torch.Tensor:
        return self.regions[rank][off:off + nbytes].view(dtype)

    def status(self, rank: int) -> int:
        return int(self._section(rank, self.layout.statusOff, 64, torch.int32)[0].item())

    def rounds_flushed(self, rank: int) -> int:
        return int(self._section(rank, self.layout.statusOff, 64, torch.int32)[1].item())

    def contributor_mask(self, rank: int, rnd: int, chunk: int) -> int:
        """Decision word of K2 for one of `rank`'s chunks: which sources were summed."""
        lay = self.layout
        words = self._section(rank, lay.k2DecOff, lay.L * lay.C * 8, torch.int64)
        w = int(words[(rnd % lay.L) * lay.C + chunk].item())
        return (w & 0xFFFFFFFF) if (w >> 32) == rnd + 1 else -1

    def slot_base(self, rank: int) -> list[list[int]]:
        """Per ring slot: round + 1 of the last round each peer retired in that slot, as seen by `rank` (the slot flow control)."""
        lay = self.layout
        return self._section(rank, lay.slotBaseOff, lay.L * L.AAR_MAXP * 4, torch.int32).view(lay.L, L.AAR_MAXP)[:, : self.P].tolist()

    def peer_base(self, rank: int) -> list[int]:
        return self._section(rank, self.layout.peerBaseOff, L.AAR_MAXP * 4, torch.int32)[: self.P].tolist()
