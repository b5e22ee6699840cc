"""Partitioning / threshold arithmetic (csrc/core/geometry.h) and the host ring (csrc/core/data_buffer.h).

The reference has no direct DataBuffer tests; thresholds are the float32 values
listed in SURVEY.md section 2.4.
"""
import numpy as np
import pytest

from akka_allreduce_b200 import _core as core


@pytest.mark.parametrize(
    "th,P,want",
    [(0.9, 2, 1), (0.75, 4, 3), (0.7, 3, 2), (0.6, 2, 1), (0.75, 8, 6), (0.9, 8, 7), (1.0, 8, 8), (0.1, 1, 1), (0.8, 4, 3)],
)
def test_min_required_float32(th, P, want):
    assert core.Geometry.min_required(th, P) == want


@pytest.mark.parametrize(
    "th,P,data,chunk,want",
    [
        (0.8, 2, 6, 2, 3),      # 0.8f * 2 * 2
        (0.7, 3, 9, 1, 6),      # 0.7f * 3 * 3
        (0.75, 3, 9, 2, 4),     # 0.75f * 3 * 2
        (0.8, 4, 8, 2, 3),      # 0.8f * 4 * 1
        (0.75, 8, 8 * 16, 1, 96),
        (0.9, 8, 8 * 16, 1, 115),
        (1.0, 2, 3, 2, 2),
    ],
)
def test_min_chunks_required_float32(th, P, data, chunk, want):
    assert core.Geometry(P, data, chunk).min_chunks_required(th) == want


def test_partition_matches_reference_formulas():
    g = core.Geometry(2, 10, 2)
    assert g.step_size == 5
    assert [(g.block_start(i), g.block_end(i)) for i in range(2)] == [(0, 5), (5, 10)]
    assert [g.chunk_len(0, c) for c in range(g.num_chunks(0))] == [2, 2, 1]
    assert g.total_chunks() == 6


def test_uneven_and_empty_blocks_are_legal():
    g = core.Geometry(4, 9, 2)  # reference: ArrayIndexOutOfBounds (B2)
    assert [g.block_len(i) for i in range(4)] == [3, 3, 3, 0]
    assert [g.num_chunks(i) for i in range(4)] == [2, 2, 2, 0]
    assert g.total_chunks() == 6 and g.min_chunks_required(1.0) == 6  # B3: reachable at thComplete = 1
    g = core.Geometry(8, 7, 4)
    assert [g.block_len(i) for i in range(8)] == [1] * 7 + [0]
    assert g.chunk_index(3, 0) == 3


def test_appendix_b_config2_sizes():
    g = core.Geometry(8, 16 * 1024 * 1024, 1024 * 1024)
    assert g.max_block_len() == 2097152 and g.max_num_chunks() == 2 and g.total_chunks() == 16


def test_ring_store_count_retire():
    b = core.DataBuffer(4, 3, 2, 2, [2, 2, 2])  # rowLen 4, 3 peers, depth 2, chunk 2
    assert b.store(0, 1, 0, [1.0, 2.0])
    assert not b.store(0, 1, 0, [5.0, 6.0])  # duplicate (src, chunk): count unchanged, data refreshed
    assert b.count_chunk(0, 0) == 1 and b.mask_chunk(0, 0) == 0b010
    assert list(b.row(0, 1)[:2]) == [5.0, 6.0]
    assert b.store(1, 0, 1, [9.0, 9.0], aux=3)
    assert b.slot_of(1) == 1 and b.aux(1, 0, 1) == 3
    assert b.count_round(0) == 1 and b.count_round(1) == 1
    b.retire(0)
    assert b.count_chunk(0, 0) == 0
    # round 2 reuses slot 0; nothing of round 0 leaks into it
    assert b.store(2, 2, 1, [7.0, 8.0])
    assert b.mask_chunk(2, 0) == 0 and b.mask_chunk(2, 1) == 0b100
    with pytest.raises(Exception):
        b.store(2, 0, 1, [1.0, 2.0, 3.0])  # overflows the row


def test_c_abi_structs_match_the_ctypes_declarations():
    """The ctypes mirror of include/akka_allreduce_b200.h is checked against sizeof() as compiled (lib.load() refuses a mismatch);
    new layout sections are appended so that every older offset stays put."""
    import ctypes as C

    from akka_allreduce_b200.ops import lib as L

    lib = L.load()
    a, b = C.c_int(0), C.c_int(0)
    assert lib.aar_abi_sizes(C.byref(a), C.byref(b)) == 0
    assert (a.value, b.value) == (C.sizeof(L.AarLayout), C.sizeof(L.AarRankParams))
    lay = L.compute_layout(8, 16 << 20, 1 << 20, 1, L.AAR_F32, 65536, True)
    # BASELINE.json config 2: the offsets the round-1 measurements ran with
    assert (lay.reduceOff, lay.stageOff, lay.sflagOff, lay.traceOff, lay.regionBytes) == (134217728, 268435456, 402653184, 402656768, 404750336)
    offs = [getattr(lay, f) for f, _ in L.AarLayout._fields_ if f.endswith("Off") and getattr(lay, f)]
    # round 2 appended the piece flags of the bulk kernels behind the eager-gather decision words
    # ... and the per-slot retirement words (rounds may retire out of order) behind those
    assert len(set(offs)) == len(offs) and max(offs) == lay.slotBaseOff and lay.egDecOff < lay.pflagSOff < lay.pflagROff < lay.slotBaseOff
    assert lay.pflagSOff >= lay.egDecOff + lay.L * lay.P * lay.C * 8
    assert lay.slotBaseOff >= lay.pflagROff + lay.L * lay.P * lay.C * lay.piecesPerChunk * 8
    assert lay.slotBaseOff + lay.L * 16 * 4 <= lay.regionBytes
    assert (lay.tilesPerPiece, lay.piecesPerChunk) == (8, 16)


def test_c_header_is_plain_c_and_the_native_engine_example_links(tmp_path):
    """examples/native_round.c: a C (not C++) caller of include/akka_allreduce_b200.h -- layout, local region, native round engine.
    On a box without a GPU it must stop at the first CUDA call with a clean error (exit 3), with a GPU it runs four rounds."""
    import shutil
    import subprocess
    from pathlib import Path

    import pytest

    from akka_allreduce_b200.ops import lib as L

    root = Path(__file__).resolve().parents[1]
    cuda = Path("/usr/local/cuda")
    if shutil.which("gcc") is None or not (cuda / "include/cuda_runtime.h").exists():
        pytest.skip("needs gcc and the CUDA headers")
    L.load()
    so = L.lib_path()
    exe = tmp_path / "native_round"
    b = subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", f"-I{root / 'include'}", f"-I{cuda / 'include'}", str(root / "examples/native_round.c"), "-o", str(exe),
                        str(so), f"-L{cuda / 'lib64'}", "-lcudart", f"-Wl,-rpath,{so.parent}"], capture_output=True, text=True, timeout=120)
    assert b.returncode == 0, b.stderr[-2000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert r.returncode in (0, 3), (r.stdout + r.stderr)[-1000:]
    if r.returncode == 0:
        assert "4 rounds flushed, 0 mismatches" in r.stdout
    else:
        assert "failed" in r.stderr
