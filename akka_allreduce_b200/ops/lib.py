"""ctypes binding of the sm_100a data plane (libaar_cuda.so, C ABI in include/akka_allreduce_b200.h).

The library is built in-tree by `akka_allreduce_b200.utils.build` and loaded from the package directory.
On a machine with a CUDA device a missing library is an error, not a fallback: nothing in this package
re-implements the kernels in PyTorch.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

AAR_MAXP = 16
AAR_F32, AAR_BF16 = 0, 1
FLAG_NORMALIZE, FLAG_EXACT_QUORUM, FLAG_FORCE, FLAG_SKIP_SCATTER, FLAG_VIEW_OUTPUT, FLAG_TRACE, FLAG_OVERLAPPED = 1, 2, 4, 8, 16, 32, 64
FLAG_SELF_DIRECT = 1024  # bulk kernels: own contribution read in place, own reduced block straight to the output
FLAG_ELIDE_LATE = 512  # opt-in (AAR_ELIDE_LATE=1): a rank a whole round late skips work nobody will read; implies FLAG_BCAST_FC
FLAG_EAGER_DEC = 256  # opt-in (AAR_STRICT=1): eager gather decides once per chunk (no half-flushed chunk when a wait gives up)
FLAG_BCAST_FC = 128  # opt-in (AAR_BCAST_FC=1): reduce-broadcast obeys the slot flow control at thresholds < 1
TRACE_CTAS, TRACE_MARKS = 512, 8
ST_TIMEOUT_SCATTER, ST_TIMEOUT_REDUCE, ST_TIMEOUT_GATHER, ST_PUSH_DROPPED, ST_BULK_FAULT = 1, 2, 4, 8, 16


class AarLayout(C.Structure):
    _fields_ = [
        ("P", C.c_int), ("L", C.c_int), ("C", C.c_int),
        ("dataSize", C.c_int), ("maxChunk", C.c_int), ("stepSize", C.c_int), ("blkStride", C.c_int), ("totalChunks", C.c_int),
        ("elemSize", C.c_int), ("dtype", C.c_int),
        ("tileElems", C.c_int), ("tilesPerChunk", C.c_int),
        ("scatterOff", C.c_uint64), ("reduceOff", C.c_uint64), ("sflagOff", C.c_uint64), ("rflagOff", C.c_uint64),
        ("peerBaseOff", C.c_uint64), ("k1DecOff", C.c_uint64), ("k1CtrOff", C.c_uint64), ("k2DecOff", C.c_uint64),
        ("k2CtrOff", C.c_uint64), ("k3DecOff", C.c_uint64), ("k3CtrOff", C.c_uint64), ("countsOff", C.c_uint64),
        ("statusOff", C.c_uint64), ("oneshotOff", C.c_uint64), ("oneshotFlagOff", C.c_uint64), ("stageOff", C.c_uint64),
        ("barrierOff", C.c_uint64), ("traceOff", C.c_uint64), ("ctrlBytes", C.c_uint64), ("regionBytes", C.c_uint64),
        ("egDecOff", C.c_uint64),
        ("tilesPerPiece", C.c_int), ("piecesPerChunk", C.c_int), ("pflagSOff", C.c_uint64), ("pflagROff", C.c_uint64),
        ("slotBaseOff", C.c_uint64),
    ]


class AarRankParams(C.Structure):
    _fields_ = [
        ("rank", C.c_int), ("round", C.c_int), ("minRequired", C.c_int), ("minChunksRequired", C.c_int), ("maxLag", C.c_int),
        ("flags", C.c_uint), ("timeoutNs", C.c_ulonglong),
        ("input", C.c_void_p), ("output", C.c_void_p), ("outCounts", C.c_void_p), ("doorbell", C.c_void_p), ("hostDone", C.c_void_p),
        ("region", C.c_void_p * AAR_MAXP), ("mcRegion", C.c_void_p),
    ]


class AarVmm(C.Structure):
    _fields_ = [("handle", C.c_uint64), ("ptr", C.c_void_p), ("bytes", C.c_size_t), ("fd", C.c_int), ("device", C.c_int)]


class AarError(RuntimeError):
    pass


_lib = None


def lib_path() -> Path:
    import os

    override = os.environ.get("AAR_CUDA_LIB")   # e.g. an instrumented build (bench/prof_reduce.py)
    return Path(override) if override else Path(__file__).resolve().parents[1] / "libaar_cuda.so"


def load():
    """Load libaar_cuda.so (building it first if the sources are newer and nvcc is around)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        from ..utils import build

        build.build_cuda()
    if not path.exists():
        raise AarError(f"{path} is missing: run `python -m akka_allreduce_b200.utils.build`")
    lib = C.CDLL(str(path))
    lib.aar_last_error.restype = C.c_char_p
    a, b = C.c_int(0), C.c_int(0)
    lib.aar_abi_sizes(C.byref(a), C.byref(b))
    if (a.value, b.value) != (C.sizeof(AarLayout), C.sizeof(AarRankParams)):
        raise AarError(f"{path} was built from a different header: AarLayout {a.value} vs {C.sizeof(AarLayout)} bytes, "
                       f"AarRankParams {b.value} vs {C.sizeof(AarRankParams)} bytes -- rebuild (python -m akka_allreduce_b200.utils.build)")
    P = C.POINTER
    sigs = {
        "aar_layout_compute": [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, P(AarLayout)],
        "aar_launch_scatter": [P(AarRankParams), P(AarLayout), C.c_int, C.c_void_p],
        "aar_launch_reduce_bcast": [P(AarRankParams), P(AarLayout), C.c_int, C.c_void_p],
        "aar_launch_gather": [P(AarRankParams), P(AarLayout), C.c_int, C.c_void_p],
        "aar_launch_round_fused": [P(AarRankParams), P(AarLayout), C.c_int, C.c_void_p],
        "aar_launch_round_pipelined": [P(AarRankParams), P(AarLayout), C.c_int, C.c_int, C.c_void_p],
        "aar_launch_round_bulk": [P(AarRankParams), P(AarLayout), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p],
        "aar_bulk_supported": [P(AarRankParams), P(AarLayout), C.c_int],
        "aar_launch_oneshot": [P(AarRankParams), P(AarLayout), C.c_int, C.c_int, C.c_void_p],
        "aar_launch_nvls": [P(AarRankParams), P(AarLayout), C.c_int, C.c_int, C.c_void_p],
        "aar_local_round": [P(AarRankParams), C.c_int, P(AarLayout), C.c_uint, C.c_int, C.c_void_p],
        "aar_local_allreduce_static": [P(C.c_void_p), P(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p],
        "aar_region_alloc_local": [C.c_size_t, P(C.c_void_p)],
        "aar_region_free_local": [C.c_void_p],
        "aar_vmm_granularity": [C.c_int, C.c_int, C.c_int, P(C.c_size_t)],
        "aar_vmm_create": [C.c_int, C.c_size_t, P(AarVmm)],
        "aar_vmm_import": [C.c_int, C.c_int, C.c_size_t, P(AarVmm)],
        "aar_vmm_release": [P(AarVmm)],
        "aar_mc_supported": [C.c_int],
        "aar_mc_create": [C.c_int, C.c_size_t, P(C.c_uint64), P(C.c_int)],
        "aar_mc_import": [C.c_int, P(C.c_uint64)],
        "aar_mc_add_device": [C.c_uint64, C.c_int],
        "aar_mc_bind_and_map": [C.c_uint64, C.c_int, P(AarVmm), C.c_size_t, P(C.c_void_p)],
        "aar_mc_release": [C.c_uint64, C.c_void_p, C.c_size_t],
        "aar_host_words_alloc": [C.c_size_t, P(P(C.c_uint)), P(P(C.c_uint))],
        "aar_host_words_free": [P(C.c_uint)],
        "aar_device_sm_count": [C.c_int, P(C.c_int)],
        "aar_enable_peer_access": [C.c_int, C.c_int],
        "aar_launch_delay": [C.c_ulonglong, C.c_void_p],
        "aar_bench_all2all": [C.c_int, C.c_int, C.c_int, P(C.c_void_p), C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p],
        "aar_bench_link": [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p],
        "aar_engine_create": [P(AarLayout), C.c_int, P(C.c_void_p), C.c_void_p, C.c_int, C.c_int, C.c_uint, C.c_ulonglong, C.c_int, C.c_int, C.c_int,
                              P(C.c_void_p)],
        "aar_engine_tune": [C.c_void_p, C.c_int, C.c_int],
        "aar_engine_enqueue": [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint, C.c_void_p],
        "aar_engine_enqueue_ex": [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint, C.c_void_p, C.c_ulonglong],
        "aar_engine_tune_bulk": [C.c_void_p, C.c_int],
        "aar_engine_start_at": [C.c_void_p, C.c_int],
        "aar_engine_wait": [C.c_void_p, C.c_int, C.c_void_p],
        "aar_engine_synchronize": [C.c_void_p],
        "aar_engine_set_doorbell": [C.c_void_p, C.c_int],
        "aar_engine_rounds_flushed": [C.c_void_p],
        "aar_engine_round_flushed": [C.c_void_p, C.c_int],
        "aar_engine_destroy": [C.c_void_p],
        "aar_bench_litmus_bulk_flag": [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p],
        "aar_bench_bulk_copy_multi": [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p],
        "aar_bench_bulk_copy": [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue  # optional entry points (later kernels) are checked at call time
        fn.argtypes = argtypes
        fn.restype = C.c_int
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().aar_last_error()
        raise AarError(f"{what or 'libaar_cuda'} failed ({rc}): {msg.decode() if msg else 'no detail'}")


def compute_layout(P: int, data_size: int, max_chunk: int, max_lag: int, dtype: int = AAR_F32, oneshot_max_elems: int = 0,
                   with_stage: bool = False) -> AarLayout:
    lay = AarLayout()
    check(load().aar_layout_compute(P, data_size, max_chunk, max_lag, dtype, oneshot_max_elems, int(with_stage), C.byref(lay)), "aar_layout_compute")
    return lay
