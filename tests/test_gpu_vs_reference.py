"""GPU (-m gpu): our CUDA path vs the UNMODIFIED reference GPU pipeline on the same inputs.

oracle/_ref/libcudasp_gecc_ref.so is the reference's own src/cudasp_gpu.cu compiled for sm_100 (recipe:
oracle/Makefile, target `ref`); oracle/gecc_ref.py drives it the way the reference's ProcessBatch does.
This pins parity to the reference itself, not only to the CPU restatement: the match flags the reference
produces must equal our match ids, bit for bit, on the SQL test rows and on synthetic tables with planted
base and label matches.
"""
import numpy as np
import pytest

import cudasp_b200_loader
from oracle import gecc_ref as GR
from oracle import ref_scan as R
from tests import synth

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not GR.available(), reason="oracle/_ref not built (needs /root/reference at build time)")]


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1
    return m


def _csr(rows):
    tw = np.frombuffer(b"".join(r[2] for r in rows), np.uint8).reshape(len(rows), 64).copy()
    outs, off = [], [0]
    for r in rows:
        outs.extend(int(v) for v in (r[3] or []) if v is not None)
        off.append(len(outs))
    return tw, np.array(outs, dtype=np.int64), np.array(off, dtype=np.uint32)


def test_reference_sql_rows(M, golden):
    """The hand-made rows of test/sql/cudasp.test through the reference binary and through ours."""
    rows = [(bytes.fromhex(r["txid"]), r["height"], bytes.fromhex(r["tweak"]), r["outputs"]) for r in golden["sql"]["test_data"]]
    tw, outs, off = _csr(rows)
    seen = set()
    for q in golden["sql"]["queries"]:
        if q["table"] != "test_data":
            continue
        key = (q["scan_sk"], q["spend"], tuple(q["labels"]))
        if key in seen:
            continue
        seen.add(key)
        sk, spend, labels = bytes.fromhex(q["scan_sk"]), bytes.fromhex(q["spend"]), [bytes.fromhex(l) for l in q["labels"]]
        n = q["rows_visible"]
        flags, _ = GR.scan_batch(sk, spend, labels, tw[:n], outs, off[: n + 1])
        with M.Query(sk, spend, labels, batch_rows=64) as qq:
            ids = qq.scan_batch(tw[:n], outs, off[: n + 1])
        assert np.array_equal(np.nonzero(flags)[0].astype(np.uint32), ids), q


@pytest.mark.parametrize("rows,L", [(5000, 0), (20000, 1), (12345, 3)])
def test_reference_synthetic_tables(M, rows, L):
    T = synth.make_table(rows, 9100 + L, n_labels=L, n_base=25, n_per_label=7 if L else 0, batch_rows=rows, specials=True)
    flags, _ = GR.scan(T["sk"], T["spend"], T["labels"], T["tweaks"], T["outputs"], T["offsets"], batch_rows=rows)
    ref_ids = np.nonzero(flags)[0].astype(np.uint32)
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=8192) as q:
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
    assert np.array_equal(ref_ids, T["expected"]), "the reference itself disagrees with the planted matches"
    assert np.array_equal(ids, ref_ids)


def test_reference_replicated_rows(M, golden):
    """test/sql/cudasp.test:120-154 -- 20 000 identical rows, every one matches."""
    rep = golden["sql"]["replicated"]
    q = [x for x in golden["sql"]["queries"] if x["table"] != "test_data"][0]
    n = 20000
    rows = [(b"", rep["height"], bytes.fromhex(rep["tweak"]), rep["outputs"])] * n
    tw, outs, off = _csr(rows)
    sk, spend, labels = bytes.fromhex(q["scan_sk"]), bytes.fromhex(q["spend"]), [bytes.fromhex(l) for l in q["labels"]]
    flags, _ = GR.scan_batch(sk, spend, labels, tw, outs, off)
    with M.Query(sk, spend, labels, batch_rows=50000) as qq:
        ids = qq.scan_batch(tw, outs, off)
    assert int(flags.sum()) == n and len(ids) == n
