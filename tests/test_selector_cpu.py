"""`algo="auto"` without a GPU: the shipped selector table (measured on the 8-GPU box, bench/sweep.py --write-table) is well formed
and pick_algo honours what a layout / a threshold setting can run."""
import json
from pathlib import Path

from akka_allreduce_b200.parallel.engine import _ALGO_CODE, ONESHOT_MAX_BYTES, pick_algo, selector_table

MiB = 1 << 20
ALL = dict(bulk_ok=True, nvls_ok=True, oneshot_ok=True)


def test_shipped_table_is_well_formed():
    t = selector_table()
    assert {"2", "4", "8"} <= set(t)
    for P in ("2", "4", "8"):
        for dt in ("fp32", "bf16"):
            rows = t[P][dt]
            assert rows and rows[-1][0] >= 1 << 40, "the last row covers every size"
            sizes = [r[0] for r in rows]
            assert sizes == sorted(sizes)
            assert all(r[1] in _ALGO_CODE and isinstance(r[2], int) for r in rows)
            assert any(r[1] in ("bulk", "pipelined", "twoshot") and r[0] >= 1 << 40 for r in rows), "a P2P row as the last resort (thresholds < 1)"
    raw = json.loads(Path(selector_table.__code__.co_filename).with_name("selector_table.json").read_text())
    assert raw == t


def test_headline_shapes_pick_what_was_measured():
    assert pick_algo(8, 4, 64 * MiB, **ALL)[0] == "nvls_bulk"
    assert pick_algo(2, 4, 64 * MiB, **ALL)[0] == "bulk"
    assert pick_algo(4, 4, 64 * MiB, **ALL)[0] == "bulk"
    assert pick_algo(2, 4, 256 * 1024, **ALL)[0] == "oneshot"
    assert pick_algo(1, 4, 64 * MiB, **ALL) == ("bulk", 0)
    assert pick_algo(1, 4, 4004, bulk_ok=False, nvls_ok=False, oneshot_ok=False) == ("twoshot", 0)


def test_rows_a_layout_cannot_run_are_skipped():
    # thresholds < 1 / exact quorum / no multicast: nothing in-switch, at any size
    for P in (2, 4, 8):
        for nbytes in (64 * 1024, MiB, 64 * MiB, 1 << 30):
            name, _ = pick_algo(P, 4, nbytes, bulk_ok=True, nvls_ok=False, oneshot_ok=True)
            assert name in ("bulk", "oneshot"), (P, nbytes, name)
    # no one-shot buffers in the layout (gradient buckets): the next usable row
    assert pick_algo(2, 4, 256 * 1024, bulk_ok=True, nvls_ok=True, oneshot_ok=False)[0] == "nvls_bulk"
    assert pick_algo(2, 4, 256 * 1024, bulk_ok=True, nvls_ok=False, oneshot_ok=False)[0] == "bulk"
    # geometry that is not 16-byte aligned: register kernels only
    assert pick_algo(8, 4, 64 * MiB, bulk_ok=False, nvls_ok=False, oneshot_ok=False)[0] == "pipelined"
    assert pick_algo(4, 4, 4096, bulk_ok=False, nvls_ok=False, oneshot_ok=True)[0] == "oneshot"


def test_without_a_table_the_fallback_rules_apply():
    assert pick_algo(8, 4, 64 * MiB, table={}, **ALL)[0] == "nvls_bulk"
    assert pick_algo(4, 4, 64 * MiB, table={}, **ALL)[0] == "bulk"
    assert pick_algo(4, 4, ONESHOT_MAX_BYTES, table={}, **ALL)[0] == "oneshot"
    assert pick_algo(4, 4, ONESHOT_MAX_BYTES + 4, table={}, **ALL)[0] == "bulk"
    # a table row with a split: handed through
    assert pick_algo(2, 2, MiB, table={"2": {"bf16": [[1 << 40, "bulk", 16 | (60 << 16)]]}}, **ALL) == ("bulk", 16 | (60 << 16))
    # the register in-switch kernel cannot normalise
    t = {"8": {"fp32": [[1 << 40, "nvls", 0], [1 << 40, "bulk", 0]]}}
    assert pick_algo(8, 4, MiB, table=t, normalize=True, **ALL)[0] == "bulk" and pick_algo(8, 4, MiB, table=t, **ALL)[0] == "nvls"
