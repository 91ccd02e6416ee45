"""GPU (-m gpu): cudasp_scan_parquet (include/cudasp_parquet.h) -- Parquet pages decoded by the library itself -- returns exactly the
planted rows of a synthetic table (oracle-made candidates), with the txid / height / tweak_key bytes of the file, for several writer
settings, both key formats and NULLs; and the DuckDB table function of the same name returns what cudasp_scan over read_parquet does."""
import os
import subprocess

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

import cudasp_b200_loader
from tests import synth
from tests.test_gpu_compressed import compress

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DUCKDB = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin", "duckdb")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def table_of(T, rows, sec1=False, null_every=0):
    off = T["offsets"].astype(np.int64)
    keys = compress(T["tweaks"]) if sec1 else T["tweaks"]
    kb = 33 if sec1 else 64
    key_list = [keys[i].tobytes() for i in range(rows)]
    txids = [b"%032d" % i for i in range(rows)]
    heights = [800000 + i // 3000 for i in range(rows)]
    lists = [T["outputs"][off[i]: off[i + 1]].tolist() for i in range(rows)]
    expected = set(int(x) for x in T["expected"])
    if null_every:
        for i in range(0, rows, null_every):
            which = (i // null_every) % 4
            if which == 0:
                key_list[i] = None
            elif which == 1:
                txids[i] = None
            elif which == 2:
                heights[i] = None
            else:
                lists[i] = None
            expected.discard(i)
    assert all(len(k) == kb for k in key_list if k is not None)
    t = pa.table({"txid": pa.array(txids, pa.binary()), "height": pa.array(heights, pa.int32()), "tweak_key": pa.array(key_list, pa.binary()),
                  "outputs": pa.array(lists, pa.list_(pa.int64()))})
    want = [(0, i, txids[i], heights[i], key_list[i]) for i in sorted(expected)]
    return t, want


@pytest.mark.parametrize("settings", [dict(compression="NONE", use_dictionary=False), dict(compression="SNAPPY", use_dictionary=True),
                                      dict(compression="NONE", use_dictionary=True, data_page_version="2.0", data_page_size=20000)])
@pytest.mark.parametrize("sec1,null_every", [(False, 0), (True, 0), (False, 37)])
def test_native_parquet_scan_returns_the_planted_rows(M, tmp_path, settings, sec1, null_every):
    rows = 60000
    T = synth.make_table(rows, 7700, n_labels=2, n_base=40, n_per_label=10, batch_rows=20000)
    t, want = table_of(T, rows, sec1, null_every)
    path = str(tmp_path / "t.parquet")
    pq.write_table(t, path, row_group_size=7000, **settings)
    with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=20000) as q:
        got, n = M.scan_parquet_files(q, [path], threads=4, with_rows=True)
        assert n == rows and got == want
        got2 = M.scan_parquet_files(q, [path, path], threads=0)
        assert got2 == want + [(1,) + w[1:] for w in want]


def test_unsupported_codec_falls_back_when_asked(M, tmp_path):
    rows = 5000
    T = synth.make_table(rows, 7701, n_labels=0, n_base=12, batch_rows=5000)
    t, want = table_of(T, rows)
    path = str(tmp_path / "z.parquet")
    pq.write_table(t, path, compression="ZSTD")
    with M.Query(T["sk"], T["spend"], [], devices=[0], batch_rows=5000) as q:
        with pytest.raises(M.CudaspError, match="codec"):
            M.scan_parquet_files(q, [path])
        assert M.scan_parquet_files(q, [path], fallback=True) == want


@pytest.mark.skipif(not os.path.exists(DUCKDB), reason="duckdb_ext/bin/duckdb not built")
def test_duckdb_table_function_equals_cudasp_scan_over_read_parquet(M, tmp_path):
    rows = 50000
    T = synth.make_table(rows, 7702, n_labels=1, n_base=30, n_per_label=15, batch_rows=300000)
    t, want = table_of(T, rows, null_every=41)
    path = str(tmp_path / "t.parquet")
    pq.write_table(t, path, row_group_size=9000, compression="SNAPPY")

    def blob(b):
        return "'" + "".join("\\x%02x" % c for c in b) + "'::BLOB"
    keys = "%s, %s, [%s]" % (blob(T["sk"]), blob(T["spend"]), ", ".join(blob(l) for l in T["labels"]))
    sql = ("SELECT count(*), sum(height), min(hex(txid)), max(hex(txid)), sum(octet_length(tweak_key)) FROM cudasp_scan_parquet('%s', %s);\n"
           "SELECT count(*), sum(height), min(hex(txid)), max(hex(txid)), sum(octet_length(tweak_key)) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM read_parquet('%s')), %s);\n"
           "SELECT count(*) FROM cudasp_scan_parquet(['%s', '%s'], %s, threads := 3, batch_size := 20000);\n" % (path, keys, path, keys, path, path, keys))
    p = subprocess.run([DUCKDB, "-csv", "-noheader"], input=sql, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr
    lines = p.stdout.strip().splitlines()
    assert lines[0] == lines[1]
    assert int(lines[0].split(",")[0]) == len(want) and int(lines[2]) == 2 * len(want)
