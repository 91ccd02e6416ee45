"""CPU: the DuckDB extension's host logic (duckdb_ext/cudasp_extension.cpp) through the REAL DuckDB engine, with the GPU library
replaced by a test double of the C ABI (tests/hostcheck/mock_cabi.cpp, selected with LD_LIBRARY_PATH).  The double's stand-in
rule -- a row matches iff one of its outputs equals the int64 at key bytes [1, 9) -- makes the expected result a closed form, so
what is checked here is everything AROUND the scan: bind errors, NULL rules (reference src/cudasp_extension.cu:195-221), the
column-wise fast path against the row-wise path (dictionary vectors, odd txid sizes, mixed key lengths, list gaps), batch cutting,
slot recycling under many worker threads, emit_early.  The EC path itself is covered by the -m gpu tests."""
import os
import struct
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin")
DUCKDB = os.path.join(BIN, "duckdb")
MOCK_SRC = os.path.join(ROOT, "tests", "hostcheck", "mock_cabi.cpp")
MOCK_DIR = os.path.join(ROOT, "tests", "hostcheck", "mock_cabi")
pytestmark = pytest.mark.skipif(not os.path.exists(DUCKDB), reason="duckdb_ext/bin/duckdb not built (tools/build_duckdb.sh)")

KEYS = "repeat('\\x01',32)::BLOB, repeat('\\x02',64)::BLOB, []::BLOB[]"


@pytest.fixture(scope="module")
def mock():
    so = os.path.join(MOCK_DIR, "libcudasp_b200.so")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(MOCK_SRC):
        os.makedirs(MOCK_DIR, exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, MOCK_SRC, "-lpthread"])
    return MOCK_DIR


def run(mock, sql, threads=4, devices=2, stats=None):
    env = dict(os.environ, LD_LIBRARY_PATH=mock, MOCK_CUDASP_DEVICES=str(devices))
    if stats:
        env["MOCK_CUDASP_STATS"] = stats
    p = subprocess.run([DUCKDB, "-csv", "-noheader"], input="PRAGMA threads=%d;\n%s" % (threads, sql), capture_output=True, text=True,
                       env=env, timeout=600)
    return p.returncode, [l for l in p.stdout.strip().splitlines() if l], p.stderr


# a table whose matching rows are known in closed form: key byte 0 is the format-neutral 0x04 (64 B) / 0x02 (33 B) prefix, bytes 1..8
# hold the little-endian int64 `i`, and outputs contains i exactly when i % 7 == 3
TABLE = """
CREATE TABLE t AS SELECT
  ('\\x' || lpad(to_hex(i), 8, '0') || repeat('ab', 28))::BLOB AS txid_hex, i AS i FROM range({n}) r(i);
CREATE TABLE src AS SELECT
  from_hex(lpad(to_hex(i), 8, '0') || repeat('ab', 28)) AS txid,
  (800000 + i // 3000)::INTEGER AS height,
  from_hex('04' || {le} || repeat('cd', 55)) AS tweak_key,
  CASE WHEN i % 7 = 3 THEN [i + 100000, i, -i] ELSE [i + 100000, i + 1]::BIGINT[] END AS outputs,
  i
FROM range({n}) r(i);
"""
# little-endian hex of the BIGINT i (8 bytes): byte k = (i >> 8k) & 255
LE = " || ".join("lpad(to_hex((i >> %d) & 255), 2, '0')" % (8 * k) for k in range(8))


def make(n):
    return TABLE.format(n=n, le=LE)


def expected(n):
    ids = [i for i in range(n) if i % 7 == 3]
    return len(ids), sum(800000 + i // 3000 for i in ids)


def test_plain_table_many_threads_and_batch_sizes(mock, tmp_path):
    n = 200000
    want = expected(n)
    for threads, batch in [(1, 300000), (4, 300000), (8, 7000), (16, 100)]:
        stats = str(tmp_path / ("stats_%d_%d" % (threads, batch)))
        rc, out, err = run(mock, make(n) + "SELECT count(*), sum(height) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM src), "
                           "%s, batch_size := %d);" % (KEYS, batch), threads=threads, stats=stats)
        assert rc == 0, err
        assert out[-1] == "%d,%d" % want
        st = eval(open(stats).read().strip().splitlines()[-1])
        assert st["rows"] == n and st["max_rows"] <= batch


def test_returned_columns_are_the_input_bytes(mock):
    n = 5000
    rc, out, err = run(mock, make(n) + "SELECT count(*) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM src), %s) s "
                       "JOIN src ON (s.txid = src.txid AND s.height = src.height AND s.tweak_key = src.tweak_key) WHERE src.i %% 7 = 3;" % KEYS)
    assert rc == 0, err
    assert int(out[-1]) == expected(n)[0]


def test_null_rules_and_odd_rows_equal_the_closed_form(mock):
    """NULL txid / height / tweak_key drop the row; a NULL list is empty; NULL children are skipped; keys that are neither 64 nor
    33 bytes never match; txids of other lengths are returned as they are; 33-byte keys cut the batch (one format per batch)."""
    n = 30000
    sql = make(n) + """
CREATE TABLE odd AS SELECT
  CASE WHEN i % 11 = 0 THEN NULL WHEN i % 13 = 0 THEN from_hex('00' || lpad(to_hex(i), 8, '0')) ELSE txid END AS txid,
  CASE WHEN i % 17 = 0 THEN NULL ELSE height END AS height,
  CASE WHEN i % 19 = 0 THEN NULL WHEN i % 23 = 0 THEN from_hex('02' || """ + LE + """ || repeat('ee', 24)) WHEN i % 29 = 0 THEN from_hex('0405') ELSE tweak_key END AS tweak_key,
  CASE WHEN i % 31 = 0 THEN NULL WHEN i % 37 = 0 THEN [NULL, i, NULL]::BIGINT[] WHEN i % 41 = 0 THEN []::BIGINT[] ELSE outputs END AS outputs,
  i
FROM src;
SELECT count(*), sum(height), sum(octet_length(txid)), sum(octet_length(tweak_key)) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM odd), """ + KEYS + """, batch_size := 4096);
"""
    rc, out, err = run(mock, sql, threads=4)
    assert rc == 0, err
    cnt = hs = tl = kl = 0
    for i in range(n):
        if i % 11 == 0 or i % 17 == 0 or i % 19 == 0:
            continue                                        # NULL txid / height / key: the row is dropped
        if i % 23 != 0 and i % 29 == 0:
            continue                                        # 2-byte key: can never match
        if i % 31 == 0 or (i % 37 != 0 and i % 41 == 0):
            continue                                        # NULL or empty list
        if not (i % 37 == 0 or i % 7 == 3):
            continue                                        # outputs do not hold i
        cnt += 1
        hs += 800000 + i // 3000
        tl += 5 if i % 13 == 0 else 32
        kl += 33 if i % 23 == 0 else 64
    assert out[-1] == "%d,%d,%d,%d" % (cnt, hs, tl, kl)


def test_parquet_dictionary_pages_take_the_column_wise_path(mock, tmp_path):
    """pyarrow writes the first pages of every column dictionary-encoded: DuckDB hands those over as dictionary vectors."""
    pa = pytest.importorskip("pyarrow")
    import numpy as np
    import pyarrow.parquet as pq
    n = 60000
    keys = bytearray(64 * n)
    for i in range(n):
        keys[64 * i] = 4
        keys[64 * i + 1: 64 * i + 9] = struct.pack("<q", i)
    tw = pa.FixedSizeBinaryArray.from_buffers(pa.binary(64), n, [None, pa.py_buffer(bytes(keys))]).cast(pa.binary())
    tx = pa.array([b"%032d" % i for i in range(n)], pa.binary())
    outs = pa.array([[i + 5, i] if i % 7 == 3 else [i + 5] for i in range(n)], pa.list_(pa.int64()))
    tbl = pa.table({"txid": tx, "height": pa.array((800000 + np.arange(n) // 3000).astype(np.int32)), "tweak_key": tw, "outputs": outs})
    path = str(tmp_path / "t.parquet")
    pq.write_table(tbl, path, row_group_size=20000)
    env = dict(os.environ, LD_LIBRARY_PATH=mock, CUDASP_TRACE="1", MOCK_CUDASP_DEVICES="1")
    sql = "PRAGMA threads=1;\nSELECT count(*), sum(height) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM read_parquet('%s')), %s);" % (path, KEYS)
    p = subprocess.run([DUCKDB, "-csv", "-noheader"], input=sql, capture_output=True, text=True, env=env, timeout=600)
    assert p.returncode == 0, p.stderr
    assert p.stdout.strip().splitlines()[-1] == "%d,%d" % expected(n)
    assert "%d rows (%d column-wise)" % (n, n) in p.stderr, p.stderr


def test_emit_early_and_bind_errors(mock):
    n = 50000
    rc, out, err = run(mock, make(n) + "SELECT count(*), sum(height) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM src), %s, "
                       "batch_size := 2048, emit_early := true);" % KEYS, threads=8)
    assert rc == 0, err
    assert out[-1] == "%d,%d" % expected(n)
    rc, out, err = run(mock, "SELECT * FROM cudasp_scan((SELECT 1), repeat('\\x01',31)::BLOB, repeat('\\x02',64)::BLOB, []::BLOB[]);")
    assert rc != 0 and "scan_private_key must be exactly 32 bytes, got 31 bytes" in err
    rc, out, err = run(mock, make(10) + "SELECT * FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM src), repeat('\\x00',32)::BLOB, "
                       "repeat('\\x02',64)::BLOB, []::BLOB[]);")
    assert rc != 0 and "Invalid Input Error" in err and "all-zero scan key" in err       # CUDASP_ERR_KEY -> InvalidInputException


def test_cudasp_scan_parquet_bind_and_refusal(mock, tmp_path):
    """The Parquet sibling: same key checks and texts as cudasp_scan; the C ABI's CUDASP_ERR_UNSUPPORTED becomes an error that points
    at cudasp_scan over read_parquet (the double refuses every file; the decoders themselves are tested in test_parquet_ingest.py)."""
    rc, out, err = run(mock, "SELECT * FROM cudasp_scan_parquet('x.parquet', repeat('\\x01',31)::BLOB, repeat('\\x02',64)::BLOB, []::BLOB[]);")
    assert rc != 0 and "scan_private_key must be exactly 32 bytes, got 31 bytes" in err
    rc, out, err = run(mock, "SELECT * FROM cudasp_scan_parquet(42, repeat('\\x01',32)::BLOB, repeat('\\x02',64)::BLOB, []::BLOB[]);")
    assert rc != 0
    rc, out, err = run(mock, "SELECT * FROM cudasp_scan_parquet(['a.parquet', 'b.parquet'], repeat('\\x01',32)::BLOB, repeat('\\x02',64)::BLOB, "
                       "[repeat('\\x03',63)::BLOB]);")
    assert rc != 0 and "Each label key must be exactly 64 bytes, got 63 bytes" in err
    rc, out, err = run(mock, "SELECT * FROM cudasp_scan_parquet('x.parquet', repeat('\\x01',32)::BLOB, repeat('\\x02',64)::BLOB, []::BLOB[], threads := 2);")
    assert rc != 0 and "use cudasp_scan((SELECT txid, height, tweak_key, outputs FROM read_parquet(...))" in err
    # glob patterns are expanded at bind time, in sorted order, next to plain paths (the double reports the paths it was handed)
    for name in ("part_2.parquet", "part_0.parquet", "part_1.parquet", "other.txt"):
        (tmp_path / name).write_bytes(b"")
    rc, out, err = run(mock, "SELECT * FROM cudasp_scan_parquet(['%s/part_*.parquet', 'plain.parquet'], %s);" % (tmp_path, KEYS))
    assert rc != 0 and "paths: part_0.parquet part_1.parquet part_2.parquet plain.parquet" in err
    rc, out, err = run(mock, "SELECT * FROM cudasp_scan_parquet('%s/none_*.parquet', %s);" % (tmp_path, KEYS))
    assert rc != 0 and "No files found" in err
