"""GPU (-m gpu): the columnar (Arrow C data interface) entry of the hot path, include/cudasp_arrow.h, against the planted
match sets, the row-wise C-ABI entry and the oracle's row semantics."""
import numpy as np
import pyarrow as pa
import pytest

import cudasp_b200_loader
from oracle import ref_scan as R
from tests import synth
from tests.test_gpu_compressed import compress

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def arrow_table(T, keys=None, width=64, list_type=pa.list_(pa.int64()), fixed=True):
    rows = len(T["offsets"]) - 1
    keys = T["tweaks"] if keys is None else keys
    k = pa.FixedSizeBinaryArray.from_buffers(pa.binary(width), rows, [None, pa.py_buffer(np.ascontiguousarray(keys).tobytes())])
    if not fixed:
        k = k.cast(pa.binary())
    off_t = pa.int64() if pa.types.is_large_list(list_type) else pa.int32()
    cls = pa.LargeListArray if pa.types.is_large_list(list_type) else pa.ListArray
    outs = cls.from_arrays(pa.array(T["offsets"].astype(np.int64), off_t), pa.array(T["outputs"], pa.int64()))
    txid = pa.array([i.to_bytes(4, "little") * 8 for i in range(rows)], pa.binary())
    return pa.table({"txid": txid, "height": pa.array((800000 + np.arange(rows) // 3000).astype(np.int32)), "tweak_key": k, "outputs": outs})


@pytest.mark.parametrize("rows,L,batch,chunk", [(1, 0, 300000, None), (5000, 2, 1024, 777), (200000, 0, 300000, 65536)])
@pytest.mark.parametrize("fmt", ["xy64", "sec1", "binary"])
def test_arrow_table_match_sets(M, rows, L, batch, chunk, fmt):
    T = synth.make_table(rows, 777 + rows, n_labels=L, n_base=min(40, rows), n_per_label=5 if L else 0, batch_rows=batch, specials=False)
    tbl = arrow_table(T, compress(T["tweaks"]), 33) if fmt == "sec1" else arrow_table(T, fixed=fmt != "binary")
    if chunk:
        tbl = pa.Table.from_batches(tbl.to_batches(max_chunksize=chunk))
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=batch) as q:
        ids = M.scan_table(q, tbl)
        assert q.launch_count() > 0
        one = M.scan_record_batch(q, tbl.combine_chunks().to_batches()[0])
        cols = M.scan_columns(q, tbl.column("tweak_key").combine_chunks(), tbl.column("outputs").combine_chunks())
    assert np.array_equal(ids, T["expected"].astype(np.uint64))
    assert np.array_equal(one, ids) and np.array_equal(cols, ids)


def test_arrow_null_rules_equal_row_function(M):
    """Same rows through the Arrow entry and through the reference-shaped row interface (and the oracle's scan_rows)."""
    rows = 3000
    T = synth.make_table(rows, 2024, n_labels=1, n_base=300, n_per_label=100, batch_rows=512, specials=True)
    off = T["offsets"].astype(np.int64)
    rng = np.random.default_rng(5)
    ck = compress(T["tweaks"])
    py_rows = []
    for i in range(rows):
        outs = [int(v) for v in T["outputs"][off[i]: off[i + 1]]]
        key = T["tweaks"][i].tobytes() if (i // 400) % 2 == 0 else ck[i].tobytes()
        r = rng.random()
        if r < 0.05:
            key = None
        elif r < 0.10:
            outs = None
        elif r < 0.20 and outs:
            outs[int(rng.integers(0, len(outs)))] = None
        elif r < 0.22:
            key = key[:31]
        elif r < 0.24 and len(key) == 33:
            key = b"\x07" + key[1:]
        py_rows.append((i.to_bytes(4, "little"), 800000 + i, key, outs))
    tbl = pa.table({"txid": pa.array([r[0] for r in py_rows], pa.binary()), "height": pa.array([r[1] for r in py_rows], pa.int32()),
                    "tweak_key": pa.array([r[2] for r in py_rows], pa.binary()),
                    "outputs": pa.array([r[3] for r in py_rows], pa.list_(pa.int64()))})
    got = M.cudasp_scan_arrow(tbl, T["sk"], T["spend"], T["labels"], batch_size=512)
    got = sorted(zip(got["txid"].to_pylist(), got["height"].to_pylist(), got["tweak_key"].to_pylist()))
    want_rows = sorted(M.cudasp_scan(py_rows, T["sk"], T["spend"], T["labels"], batch_size=512))
    assert got == want_rows and len(got) > 100
    assert got == sorted(R.scan_rows(T["sk"], T["spend"], T["labels"], py_rows))


def test_arrow_parquet_reader_stream(M, tmp_path):
    import pyarrow.parquet as pq
    rows = 150000
    T = synth.make_table(rows, 99, n_labels=0, n_base=60, batch_rows=300000, specials=False)
    path = str(tmp_path / "t.parquet")
    pq.write_table(arrow_table(T), path, row_group_size=20000, compression="NONE")
    pf = pq.ParquetFile(path)
    reader = pa.RecordBatchReader.from_batches(pf.schema_arrow, pf.iter_batches(batch_size=20000))
    with M.Query(T["sk"], T["spend"], [], batch_rows=300000) as q:
        ids, seen = M.scan_stream(q, reader)
        ids2, seen2 = M.scan_parquet(q, path, threads=3)
    assert seen == rows and np.array_equal(ids, T["expected"].astype(np.uint64))
    assert seen2 == rows and np.array_equal(ids2, ids)


def test_arrow_argument_errors(M):
    T = synth.make_table(10, 1, n_base=2, specials=False)
    with M.Query(T["sk"], T["spend"], []) as q:
        bad = pa.record_batch({"tweak_key": pa.array([b"x" * 64], pa.binary()), "outputs": pa.array([[1.0]], pa.list_(pa.float64()))})
        with pytest.raises(M.CudaspError, match="int64"):
            M.scan_record_batch(q, bad)
        with pytest.raises(KeyError):
            M.scan_record_batch(q, bad, tweak_col="nope")
