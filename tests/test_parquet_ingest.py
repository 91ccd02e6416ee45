"""CPU: the Parquet front-end (csrc/parquet_ingest.cpp, include/cudasp_parquet.h) against the test double of the staging core
(tests/hostcheck/mock_staging.cpp; its match rule: a row matches iff one output equals int64_le(key[1:9])).  Under test: the page
decoders -- footer, data pages V1 / V2, PLAIN and dictionary encodings, RLE / bit-packed levels, Snappy, rows spanning pages -- and
the reference's NULL rules (src/cudasp_extension.cu:195-221) on files written by pyarrow with many writer settings."""
import ctypes as C
import os
import subprocess
import time

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "csrc")
SO = os.path.join(ROOT, "tests", "hostcheck", "libparquetcheck.so")
SRCS = [os.path.join(CSRC, "parquet_ingest.cpp"), os.path.join(CSRC, "arrow_ingest.cpp"), os.path.join(ROOT, "tests", "hostcheck", "mock_staging.cpp")]
DEPS = SRCS + [os.path.join(CSRC, "staging_internal.hpp"), os.path.join(ROOT, "include", "cudasp_arrow.h"), os.path.join(ROOT, "include", "cudasp_parquet.h")]
ERR_UNSUPPORTED, ERR_IO = -5, -6


class Result(C.Structure):
    _fields_ = [("n_matches", C.c_uint64), ("rows_scanned", C.c_uint64), ("file_index", C.POINTER(C.c_uint32)),
                ("row_in_file", C.POINTER(C.c_uint64)), ("heights", C.POINTER(C.c_int32)), ("txids", C.POINTER(C.c_uint8)),
                ("txid_offsets", C.POINTER(C.c_uint64)), ("tweak_keys", C.POINTER(C.c_uint8)), ("key_offsets", C.POINTER(C.c_uint64))]


@pytest.fixture(scope="module")
def P():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(d) for d in DEPS):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-o", SO] + SRCS + ["-lpthread"])
    L = C.CDLL(SO)
    L.mock_query_create.restype = C.c_void_p
    L.mock_query_create.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32]
    L.mock_query_destroy.argtypes = [C.c_void_p]
    L.mock_busy_slots.argtypes = [C.c_void_p]
    L.mock_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    L.cudasp_last_error.restype = C.c_char_p
    L.cudasp_last_error.argtypes = [C.c_void_p]
    L.cudasp_scan_parquet.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.c_int, C.POINTER(C.POINTER(Result))]
    L.cudasp_parquet_result_free.argtypes = [C.POINTER(Result)]

    def scan(paths, piece=97, cap_rows=128, cap_outputs=1024, stats=None, threads=1, slots=3):
        q = L.mock_query_create(piece, cap_rows, cap_outputs, slots)
        try:
            arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
            out = C.POINTER(Result)()
            rc = L.cudasp_scan_parquet(q, arr, len(paths), threads, C.byref(out))
            if rc:
                return rc, (L.cudasp_last_error(q) or b"").decode()
            r = out.contents
            got = []
            for i in range(r.n_matches):
                t0, t1, k0, k1 = r.txid_offsets[i], r.txid_offsets[i + 1], r.key_offsets[i], r.key_offsets[i + 1]
                got.append((r.file_index[i], r.row_in_file[i], bytes(r.txids[t0:t1]), r.heights[i], bytes(r.tweak_keys[k0:k1])))
            rows = r.rows_scanned
            L.cudasp_parquet_result_free(out)
            assert L.mock_busy_slots(q) == 0, "a staging slot leaked"
            if stats is not None:
                st = (C.c_uint64 * 6)()
                L.mock_stats(q, st)
                stats.update(submits=st[0], rows=st[1])
            return 0, (rows, got)
        finally:
            L.mock_query_destroy(q)
    return scan


def make_rows(n, seed, p_null=0.0, key_sizes=(64,), txid_sizes=(32,)):
    """-> python rows (txid, height, key, outputs) and the expected matches under the reference's NULL rules + the double's rule."""
    rng = np.random.default_rng(seed)
    rows, expect = [], []
    for i in range(n):
        ks = int(rng.choice(key_sizes))
        key = rng.integers(0, 256, ks, dtype=np.uint8).tobytes()
        want = int.from_bytes(key[1:9], "little", signed=True) if ks >= 9 else 0
        outs = [int(x) for x in rng.integers(-2**62, 2**62, int(rng.integers(0, 7)))]
        hit = rng.random() < 0.25
        if hit:
            outs.insert(int(rng.integers(0, len(outs) + 1)), want)
        if rng.random() < p_null and outs:                       # a NULL child (skipped)
            j = int(rng.integers(0, len(outs)))
            outs[j] = None
        lst = outs
        if rng.random() < p_null:
            lst = None                                           # NULL list = empty
        txid = rng.integers(0, 256, int(rng.choice(txid_sizes)), dtype=np.uint8).tobytes() if rng.random() >= p_null else None
        height = int(800000 + i) if rng.random() >= p_null else None
        k = key if rng.random() >= p_null else None
        rows.append((txid, height, k, lst))
        ok = txid is not None and height is not None and k is not None and len(k) in (64, 33) and lst is not None and want in [o for o in lst if o is not None]
        if ok:
            expect.append((i, txid, height, k))
    return rows, expect


def table_of(rows, key_type=pa.binary(), txid_type=pa.binary()):
    return pa.table({"txid": pa.array([r[0] for r in rows], txid_type), "height": pa.array([r[1] for r in rows], pa.int32()),
                     "tweak_key": pa.array([r[2] for r in rows], key_type), "outputs": pa.array([r[3] for r in rows], pa.list_(pa.int64()))})


WRITER_SETTINGS = [
    dict(compression="NONE", use_dictionary=False),
    dict(compression="NONE", use_dictionary=True),
    dict(compression="NONE", use_dictionary=True, data_page_version="2.0"),
    dict(compression="SNAPPY", use_dictionary=False, data_page_version="2.0"),
    dict(compression="SNAPPY", use_dictionary=True),
    dict(compression="NONE", use_dictionary=False, data_page_size=700),           # many pages per chunk; V1 rows may span pages
    dict(compression="SNAPPY", use_dictionary=True, data_page_size=900, dictionary_pagesize_limit=2000),   # dictionary overflow -> PLAIN
]


@pytest.mark.parametrize("settings", range(len(WRITER_SETTINGS)))
@pytest.mark.parametrize("nulls", [0.0, 0.07])
def test_parquet_decoders_equal_the_closed_form(P, tmp_path, settings, nulls):
    rows, expect = make_rows(3000, 100 + settings, p_null=nulls, key_sizes=(64, 64, 64, 33, 12), txid_sizes=(32, 32, 5))
    path = str(tmp_path / "t.parquet")
    pq.write_table(table_of(rows), path, row_group_size=1100, **WRITER_SETTINGS[settings])
    rc, res = P([path])
    assert rc == 0, res
    n, got = res
    assert n == 3000
    assert [(g[1], g[2], g[3], g[4]) for g in got] == expect and all(g[0] == 0 for g in got)
    # the same file twice: file order then row order
    rc, res = P([path, path])
    assert rc == 0 and res[0] == 6000
    assert [(g[0], g[1]) for g in res[1]] == [(0, e[0]) for e in expect] + [(1, e[0]) for e in expect]


def test_fixed_size_columns_required_fields_and_positions(P, tmp_path):
    rows, expect = make_rows(2500, 7)
    t = table_of(rows, key_type=pa.binary(64), txid_type=pa.binary(32))
    # required (non-nullable) fields: no definition levels at all; other names: the first four columns by position
    schema = pa.schema([pa.field("a", pa.binary(32), nullable=False), pa.field("b", pa.int32(), nullable=False),
                        pa.field("c", pa.binary(64), nullable=False), pa.field("d", pa.list_(pa.field("item", pa.int64(), nullable=False)), nullable=False),
                        pa.field("extra", pa.int64())])
    t2 = pa.table([t["txid"], t["height"], t["tweak_key"], t["outputs"].cast(schema.field("d").type), pa.array(range(2500), pa.int64())], schema=schema)
    for i, kw in enumerate([dict(use_dictionary=False), dict(use_dictionary=True, data_page_version="2.0", compression="SNAPPY")]):
        path = str(tmp_path / ("f%d.parquet" % i))
        pq.write_table(t2, path, row_group_size=999, **kw)
        rc, res = P([path])
        assert rc == 0, res
        assert res[0] == 2500 and [(g[1], g[2], g[3], g[4]) for g in res[1]] == expect


def test_unsupported_and_malformed_files_are_refused(P, tmp_path):
    rows, _ = make_rows(200, 3)
    path = str(tmp_path / "z.parquet")
    pq.write_table(table_of(rows), path, compression="ZSTD")
    rc, msg = P([path])
    assert rc == ERR_UNSUPPORTED and "codec" in msg
    path2 = str(tmp_path / "few.parquet")
    pq.write_table(pa.table({"txid": pa.array([b"x"]), "height": pa.array([1], pa.int32())}), path2)
    rc, msg = P([path2])
    assert rc == ERR_UNSUPPORTED and "four columns" in msg
    path3 = str(tmp_path / "wrong.parquet")
    pq.write_table(pa.table({"txid": pa.array([b"x"]), "height": pa.array([1.5]), "tweak_key": pa.array([b"k"]), "outputs": pa.array([[1]], pa.list_(pa.int64()))}), path3)
    rc, msg = P([path3])
    assert rc == ERR_UNSUPPORTED and "INT32" in msg
    junk = str(tmp_path / "junk.parquet")
    open(junk, "wb").write(b"not a parquet file at all")
    assert P([junk])[0] == ERR_IO
    assert P([str(tmp_path / "missing.parquet")])[0] == ERR_IO
    # a truncated copy of a good file: refused, never read out of bounds
    good = str(tmp_path / "good.parquet")
    pq.write_table(table_of(rows), good, use_dictionary=False, compression="NONE")
    data = open(good, "rb").read()
    cut = str(tmp_path / "cut.parquet")
    open(cut, "wb").write(data[: len(data) // 2] + data[-200:])
    assert P([cut])[0] in (ERR_IO, ERR_UNSUPPORTED)
    assert P([])[0] == 0
    # hand-made footers (found by tests/hostcheck/parquet_fuzz.py): a map of 2^40 one-bit values, which the thrift skipper once
    # walked element by element, and structs nested 100 000 deep, which it once followed recursively
    def with_footer(name, footer):
        f = str(tmp_path / name)
        open(f, "wb").write(b"PAR1" + footer + len(footer).to_bytes(4, "little") + b"PAR1")
        return f
    t0 = time.time()
    rc, msg = P([with_footer("map.parquet", bytes([0xFB, 0x80, 0x80, 0x80, 0x80, 0x80, 0x20, 0x11, 0x00]))])
    assert rc == ERR_IO and "map longer" in msg
    rc, msg = P([with_footer("deep.parquet", bytes([0x1C]) * 100000)])
    assert rc == ERR_IO and "nesting" in msg
    assert time.time() - t0 < 5


DUCKDB = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin", "duckdb")
MOCK_DIR = os.path.join(ROOT, "tests", "hostcheck", "mock_cabi")


@pytest.mark.skipif(not os.path.exists(DUCKDB), reason="duckdb_ext/bin/duckdb not built (tools/build_duckdb.sh)")
@pytest.mark.parametrize("codec", ["uncompressed", "snappy"])
def test_files_written_by_duckdb(P, tmp_path, codec):
    """DuckDB's own Parquet writer (other page sizes, dictionary policy and level layout than pyarrow's)."""
    rows, expect = make_rows(40000, 11, p_null=0.03, key_sizes=(64, 64, 33), txid_sizes=(32,))
    src = str(tmp_path / "src.parquet")
    pq.write_table(table_of(rows), src)
    dst = str(tmp_path / "duck.parquet")
    env = dict(os.environ, LD_LIBRARY_PATH=MOCK_DIR)
    if not os.path.exists(os.path.join(MOCK_DIR, "libcudasp_b200.so")):
        os.makedirs(MOCK_DIR, exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", os.path.join(MOCK_DIR, "libcudasp_b200.so"),
                               os.path.join(ROOT, "tests", "hostcheck", "mock_cabi.cpp"), "-lpthread"])
    sql = "COPY (SELECT txid, height, tweak_key, outputs FROM read_parquet('%s')) TO '%s' (FORMAT parquet, COMPRESSION %s, ROW_GROUP_SIZE 16384);" % (src, dst, codec)
    p = subprocess.run([DUCKDB], input=sql, capture_output=True, text=True, env=env, timeout=300)
    assert p.returncode == 0, p.stderr
    rc, res = P([dst], piece=4000, cap_rows=4096, cap_outputs=40000)
    assert rc == 0, res
    assert res[0] == 40000 and [(g[1], g[2], g[3], g[4]) for g in res[1]] == expect


def test_pieces_span_row_groups(P, tmp_path):
    """Row groups smaller than a piece: the rows of consecutive row groups (and files) fill the same pieces, and a match is
    still reported with its own file, row, txid, height and key."""
    rows, expect = make_rows(2000, 77)
    path = str(tmp_path / "small_groups.parquet")
    pq.write_table(table_of(rows), path, row_group_size=50, compression="NONE")
    assert pq.ParquetFile(path).metadata.num_row_groups == 40
    st = {}
    rc, res = P([path, path], piece=128, cap_rows=128, cap_outputs=4096, stats=st)
    assert rc == 0, res
    assert res[0] == 4000
    assert [(g[0],) + g[1:] for g in res[1]] == [(0,) + e for e in expect] + [(1,) + e for e in expect]
    assert st["rows"] == 4000 and st["submits"] == -(-4000 // 128)         # 32 pieces, not 80


@pytest.mark.parametrize("key_type", [pa.binary(), pa.binary(64), pa.binary(33)])
@pytest.mark.parametrize("settings", [0, 2, 3, 5])
def test_dense_pages_staged_where_they_lie(P, tmp_path, key_type, settings):
    """The fast paths: pages of plain 64- / 33-byte keys without NULLs are staged straight from the file mapping (fixed-length
    values packed, BYTE_ARRAY values behind their 4-byte lengths), and pages of outputs in which every level entry is a value
    find their row starts by scanning the one-bit repetition levels -- RLE runs of 0 (one-element lists), RLE runs of 1 (one
    long list) and bit-packed stretches."""
    rng = np.random.default_rng(500 + settings)
    w = 33 if key_type == pa.binary(33) else 64
    lens = np.concatenate([np.ones(3000, np.int64), [700], rng.integers(1, 10, 4000), np.ones(77, np.int64), [2, 1, 300], rng.integers(1, 4, 2500)])
    n = len(lens)
    keys = rng.integers(0, 256, (n, w), dtype=np.uint8)
    want = keys[:, 1:9].copy().view("<i8").reshape(n)
    off = np.zeros(n + 1, np.int64)
    off[1:] = np.cumsum(lens)
    vals = rng.integers(-2**62, 2**62, int(off[-1]), dtype=np.int64)
    hit = np.flatnonzero(rng.random(n) < 0.1)
    vals[off[hit] + rng.integers(0, 1 << 30, len(hit)) % lens[hit]] = want[hit]
    txid = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    height = (700000 + np.arange(n)).astype(np.int32)
    tbl = pa.table({"txid": pa.FixedSizeBinaryArray.from_buffers(pa.binary(32), n, [None, pa.py_buffer(txid.tobytes())]).cast(pa.binary()),
                    "height": pa.array(height),
                    "tweak_key": pa.FixedSizeBinaryArray.from_buffers(pa.binary(w), n, [None, pa.py_buffer(keys.tobytes())]).cast(key_type),
                    "outputs": pa.ListArray.from_arrays(pa.array(off.astype(np.int32)), pa.array(vals))})
    path = str(tmp_path / "dense.parquet")
    pq.write_table(tbl, path, row_group_size=4100, **WRITER_SETTINGS[settings])
    st = {}
    rc, res = P([path], piece=1000, cap_rows=1000, cap_outputs=2000, stats=st)
    assert rc == 0, res
    assert res[0] == n and st["rows"] == n
    assert [(g[1], g[2], g[3], g[4]) for g in res[1]] == [(int(i), txid[i].tobytes(), int(height[i]), keys[i].tobytes()) for i in hit]


@pytest.mark.parametrize("threads,slots", [(4, 6), (8, 3), (3, 1)])
def test_decoder_threads_share_one_query(P, tmp_path, threads, slots):
    """Several decoder threads, each with its own scan over the row groups it takes, on one query: the result is the
    single-threaded one (file order, then row order) whatever the interleaving -- also with fewer slots than threads, where a
    thread holding an open piece makes the others wait."""
    rows, expect = make_rows(6000, 321, p_null=0.03, key_sizes=(64, 64, 33))
    paths = []
    for i, st in enumerate((0, 4)):
        path = str(tmp_path / ("f%d.parquet" % i))
        pq.write_table(table_of(rows), path, row_group_size=250, **WRITER_SETTINGS[st])
        paths.append(path)
    st = {}
    rc, res = P(paths, piece=128, cap_rows=128, cap_outputs=4096, stats=st, threads=threads, slots=slots)
    assert rc == 0, res
    assert res[0] == 12000
    assert [(g[0],) + g[1:] for g in res[1]] == [(0,) + e for e in expect] + [(1,) + e for e in expect]
