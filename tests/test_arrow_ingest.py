"""CPU: the Arrow ingest front-end (csrc/arrow_ingest.cpp, include/cudasp_arrow.h) against a test double of the
staging core (tests/hostcheck/mock_staging.cpp).  What is under test is the host logic -- the reference's NULL rules
(src/cudasp_extension.cu:195-221), row maps, piece cutting, key-format switches, slicing/offsets, slot recycling --
not the EC arithmetic (the double's match rule: a row matches iff one output equals int64_le(key[1:9]))."""
import ctypes as C
import os
import subprocess

import numpy as np
import pyarrow as pa
import pytest

import cudasp_b200_loader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "csrc")
SO = os.path.join(ROOT, "tests", "hostcheck", "libarrowcheck.so")
SRCS = [os.path.join(CSRC, "arrow_ingest.cpp"), os.path.join(ROOT, "tests", "hostcheck", "mock_staging.cpp")]
DEPS = SRCS + [os.path.join(CSRC, "staging_internal.hpp"), os.path.join(ROOT, "include", "cudasp_arrow.h")]


@pytest.fixture(scope="module")
def A():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(d) for d in DEPS):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-o", SO] + SRCS)
    cudasp_b200_loader.load()
    from importlib import import_module
    mod = import_module("cudasp_b200.host.arrow")
    L = C.CDLL(SO)
    L.mock_query_create.restype = C.c_void_p
    L.mock_query_create.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32]
    L.mock_query_destroy.argtypes = [C.c_void_p]
    L.mock_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    L.mock_busy_slots.argtypes = [C.c_void_p]

    class Ctx:
        lib, arrow = L, mod

        class query:
            def __init__(self, piece=64, cap_rows=64, cap_outputs=256, slots=3):
                self.h = L.mock_query_create(piece, cap_rows, cap_outputs, slots)

            def __enter__(self):
                return self

            def __exit__(self, *a):
                assert L.mock_busy_slots(self.h) == 0, "a staging slot leaked"
                L.mock_query_destroy(self.h)

            def stats(self):
                s = (C.c_uint64 * 6)()
                L.mock_stats(self.h, s)
                return dict(zip(["submits", "rows", "submits_sec1", "outputs", "max_in_flight", "reserves"], list(s)))
    return Ctx


def make_rows(n, seed, p_match=0.2, key_sizes=(64,), p_null_key=0.0, p_null_list=0.0, p_null_child=0.0, p_bad_len=0.0):
    """Python-level rows + the expected matching row indices under the reference's NULL rules and the double's rule."""
    rng = np.random.default_rng(seed)
    keys, lists, expect = [], [], []
    for i in range(n):
        size = int(rng.choice(key_sizes))
        if rng.random() < p_bad_len:
            size = int(rng.choice([0, 12, 32, 65]))
        key = rng.integers(0, 256, size, dtype=np.uint8).tobytes()
        want = int.from_bytes(key[1:9], "little", signed=True) if size >= 9 else 0
        outs = [int(x) for x in rng.integers(-2**62, 2**62, int(rng.integers(0, 6)))]
        hit = rng.random() < p_match
        if hit:
            outs.insert(int(rng.integers(0, len(outs) + 1)), want)
        if rng.random() < p_null_child and outs:
            j = int(rng.integers(0, len(outs)))
            if outs[j] == want:
                hit = want in outs[:j] + outs[j + 1:]
            outs[j] = None
        lst = outs
        if rng.random() < p_null_list:
            lst, hit = None, False
        if rng.random() < p_null_key:
            key, hit = None, False
        if key is not None and len(key) not in (33, 64):
            hit = False
        if hit and (lst is None or want not in [o for o in lst if o is not None]):
            hit = False
        keys.append(key)
        lists.append(lst)
        if hit:
            expect.append(i)
    return keys, lists, np.array(expect, np.uint64)


def batch_of(keys, lists, key_type=pa.binary(), list_type=pa.list_(pa.int64())):
    return pa.record_batch({"txid": pa.array([b"t%d" % i for i in range(len(keys))], pa.binary()),
                            "tweak_key": pa.array(keys, key_type), "outputs": pa.array(lists, list_type)})


@pytest.mark.parametrize("key_type,width", [(pa.binary(64), 64), (pa.binary(33), 33), (pa.binary(), 64), (pa.large_binary(), 33)])
@pytest.mark.parametrize("list_type", [pa.list_(pa.int64()), pa.large_list(pa.int64())])
def test_dense_batches_all_column_types(A, key_type, width, list_type):
    keys, lists, expect = make_rows(1000, 1, key_sizes=(width,))
    b = batch_of(keys, lists, key_type, list_type)
    with A.query(piece=128, cap_rows=128, cap_outputs=64, slots=3) as q:
        ids = A.arrow.scan_record_batch(q.h, b, _lib=A.lib)
        st = q.stats()
    assert np.array_equal(ids, expect)
    assert st["rows"] == 1000 and st["submits"] == 8            # ceil(1000/128) dense pieces, nothing dropped
    assert st["submits_sec1"] == (8 if width == 33 else 0)
    assert st["max_in_flight"] == 3 and st["reserves"] > 0       # slots recycled; small output staging grown on demand


def test_null_rules_and_row_map(A):
    keys, lists, expect = make_rows(3000, 2, p_null_key=0.1, p_null_list=0.1, p_null_child=0.3, p_bad_len=0.05)
    assert len(expect) > 100
    b = batch_of(keys, lists)
    with A.query(piece=100, cap_rows=100, cap_outputs=40, slots=2) as q:
        ids = A.arrow.scan_record_batch(q.h, b, _lib=A.lib)
        st = q.stats()
    assert np.array_equal(ids, expect)
    n_live = sum(1 for k in keys if k is not None and len(k) in (33, 64))
    assert st["rows"] == n_live                                  # dropped rows never reach staging
    n_children = sum(len([o for o in l if o is not None]) for k, l in zip(keys, lists)
                     if k is not None and len(k) in (33, 64) and l is not None)
    assert st["outputs"] == n_children                           # NULL children / NULL lists are not staged


def test_mixed_key_formats_cut_pieces(A):
    keys, lists, expect = make_rows(2000, 3, key_sizes=(33, 64))
    with A.query(piece=256, cap_rows=256, cap_outputs=4096, slots=4) as q:
        ids = A.arrow.scan_record_batch(q.h, batch_of(keys, lists), _lib=A.lib)
        st = q.stats()
    assert np.array_equal(ids, expect)
    assert st["rows"] == 2000 and 0 < st["submits_sec1"] < st["submits"]
    # runs of one format, as an index would store them: dense pieces again
    keys2 = [k if (i // 500) % 2 == 0 else k[:33] for i, k in enumerate(make_rows(2000, 4, key_sizes=(64,))[0])]
    lists2 = make_rows(2000, 4, key_sizes=(64,))[1]
    exp2 = [i for i, (k, l) in enumerate(zip(keys2, lists2)) if int.from_bytes(k[1:9], "little", signed=True) in l]
    with A.query(piece=250, cap_rows=250, cap_outputs=4096, slots=4) as q:
        ids = A.arrow.scan_record_batch(q.h, batch_of(keys2, lists2), _lib=A.lib)
        assert q.stats()["submits"] == 8 and q.stats()["submits_sec1"] == 4
    assert ids.tolist() == exp2


def test_sliced_arrays_and_columns_entry(A):
    keys, lists, _ = make_rows(500, 5, p_null_list=0.1, p_null_child=0.2)
    b = batch_of(keys, lists, pa.binary(64))
    lo, n = 37, 301
    sl = b.slice(lo, n)
    want = [i - lo for i in range(lo, lo + n)
            if lists[i] is not None and int.from_bytes(keys[i][1:9], "little", signed=True) in [o for o in lists[i] if o is not None]]
    with A.query() as q:
        assert A.arrow.scan_record_batch(q.h, sl, _lib=A.lib).tolist() == want
        assert A.arrow.scan_columns(q.h, sl.column("tweak_key"), sl.column("outputs"), _lib=A.lib).tolist() == want
        # a struct array that is itself sliced (parent offset applies to unsliced children) and has NULL struct rows
        mask = np.zeros(500, bool)
        mask[::7] = True
        st = pa.StructArray.from_arrays([b.column("tweak_key"), b.column("outputs")], names=["tweak_key", "outputs"],
                                        mask=pa.array(mask))
        got = A.arrow.scan_record_batch(q.h, st.slice(lo, n), _lib=A.lib).tolist()
        assert got == [i for i in want if not mask[i + lo]] and len(got) < len(want)
        with pytest.raises(Exception, match="differ in length"):
            A.arrow.scan_columns(q.h, b.column("tweak_key"), sl.column("outputs"), _lib=A.lib)


def test_stream_of_batches_global_row_numbers(A):
    parts, expect, base = [], [], 0
    for s, n in enumerate([700, 1, 0, 1300, 64]):
        keys, lists, e = make_rows(n, 10 + s, p_null_key=0.05 if s % 2 else 0.0)
        parts.append(batch_of(keys, lists))
        expect.extend((e + base).tolist())
        base += n
    tbl = pa.Table.from_batches(parts)
    with A.query(piece=128, cap_rows=128, cap_outputs=1024, slots=3) as q:
        ids = A.arrow.scan_table(q.h, tbl, _lib=A.lib)
        assert ids.tolist() == expect
        reader = pa.RecordBatchReader.from_batches(tbl.schema, tbl.to_batches(max_chunksize=333))
        ids2, seen = A.arrow.scan_stream(q.h, reader, _lib=A.lib)
        assert ids2.tolist() == expect and seen == base
        # pieces of consecutive batches share the slot queue: nothing drains between batches
        assert q.stats()["max_in_flight"] == 3


def test_edge_cases_and_errors(A):
    with A.query() as q:
        empty = batch_of([], [])
        assert A.arrow.scan_record_batch(q.h, empty, _lib=A.lib).tolist() == []
        allnull = batch_of([None] * 10, [[1]] * 10)
        assert A.arrow.scan_record_batch(q.h, allnull, _lib=A.lib).tolist() == []
        assert q.stats()["submits"] == 0                         # an all-NULL batch costs no slot
        one = bytes(range(64))
        want = int.from_bytes(one[1:9], "little", signed=True)
        big = batch_of([one], [[7] * 5000 + [want]])             # one row with more outputs than the slot holds
        assert A.arrow.scan_record_batch(q.h, big, _lib=A.lib).tolist() == [0]
        bad = pa.record_batch({"tweak_key": pa.array([one], pa.binary()), "outputs": pa.array([[1]], pa.list_(pa.int32()))})
        with pytest.raises(Exception, match="int64"):
            A.arrow.scan_record_batch(q.h, bad, _lib=A.lib)
        bad2 = pa.record_batch({"tweak_key": pa.array(["x"], pa.string()), "outputs": pa.array([[1]], pa.list_(pa.int64()))})
        with pytest.raises(Exception, match="tweak_key must be"):
            A.arrow.scan_record_batch(q.h, bad2, _lib=A.lib)
        wrong_width = batch_of([b"\x01" * 32] * 5, [[1]] * 5, pa.binary(32))
        assert A.arrow.scan_record_batch(q.h, wrong_width, _lib=A.lib).tolist() == []


def test_first_pieces_ramp_up(A):
    """A scan starts with small pieces (1/8, 1/4, 1/2 of a full one) so that the device starts early."""
    n = 300000
    keys = np.zeros((n, 64), np.uint8)
    keys[:, 1:9] = np.arange(n, dtype=np.int64).view(np.uint8).reshape(n, 8)
    k = pa.FixedSizeBinaryArray.from_buffers(pa.binary(64), n, [None, pa.py_buffer(keys.tobytes())])
    vals = np.arange(n, dtype=np.int64)
    vals[1::2] = -1                                              # every even row matches the double's rule
    outs = pa.ListArray.from_arrays(pa.array(np.arange(n + 1, dtype=np.int32)), pa.array(vals))
    b = pa.record_batch({"tweak_key": k, "outputs": outs})
    with A.query(piece=65536, cap_rows=65536, cap_outputs=65536, slots=3) as q:
        ids = A.arrow.scan_record_batch(q.h, b, _lib=A.lib)
        st = q.stats()
    assert np.array_equal(ids, np.arange(0, n, 2, dtype=np.uint64))
    # 8192 + 16384 + 32768 + 65536 * 3 + 46040 = 300000 -> 7 pieces instead of ceil(300000 / 65536) = 5
    assert st["submits"] == 7 and st["rows"] == n


def test_small_batches_fill_whole_pieces(A):
    """Rows that a batch leaves over stay in an open piece which the next batch goes on filling: a stream of small record
    batches reaches the device in whole pieces, and a piece made of several batches (dense ones and ones with NULLs, which go
    row by row) still maps its matches back to global row numbers."""
    parts, expect, base = [], [], 0
    sizes = [100, 37, 1, 0, 250, 64, 90, 5, 333, 128, 19]
    for s, n in enumerate(sizes):
        keys, lists, e = make_rows(n, 40 + s, p_match=0.3, p_null_key=0.1 if s % 3 == 1 else 0.0,
                                   p_null_child=0.2 if s % 3 == 2 else 0.0)
        parts.append(batch_of(keys, lists))
        expect.extend((e + base).tolist())
        base += n
    staged = sum(sum(k is not None for k in p.column("tweak_key").to_pylist()) for p in parts)
    with A.query(piece=128, cap_rows=128, cap_outputs=4096, slots=3) as q:
        reader = pa.RecordBatchReader.from_batches(parts[0].schema, parts)
        ids, seen = A.arrow.scan_stream(q.h, reader, _lib=A.lib)
        st = q.stats()
    assert ids.tolist() == expect and seen == base
    assert st["rows"] == staged
    # whole pieces of 128 and one tail (a piece this small does not ramp) -- not one under-filled piece per batch
    assert st["submits"] == -(-staged // 128) < len([n for n in sizes if n])


def test_open_piece_is_cut_when_outputs_do_not_fit(A):
    """An open piece with no room left for the next rows' outputs is submitted as it is; the rows start a new piece."""
    n = 40
    keys = [bytes([i]) * 64 for i in range(n)]
    want = [int.from_bytes(k[1:9], "little", signed=True) for k in keys]
    lists = [[want[i]] + [7] * 9 for i in range(n)]              # 10 outputs per row, every row matches
    batches = [batch_of(keys[i:i + 8], lists[i:i + 8]) for i in range(0, n, 8)]
    with A.query(piece=64, cap_rows=64, cap_outputs=128, slots=2) as q:
        reader = pa.RecordBatchReader.from_batches(batches[0].schema, batches)
        ids, seen = A.arrow.scan_stream(q.h, reader, _lib=A.lib)
        st = q.stats()
    assert ids.tolist() == list(range(n)) and seen == n
    assert st["rows"] == n and st["outputs"] == 10 * n and st["reserves"] == 0
    # 128 outputs hold 12 rows; the dense path does not split a batch of 8 to fill the last 4: 8 rows a piece
    assert st["submits"] == 5


def test_open_pieces_random_streams(A):
    """Randomised streams (seeded): batch sizes from 0 to a few pieces, piece sizes, output capacities small enough to cut pieces
    early and to force an empty piece's staging to grow, mixed key lengths (format switches inside and between batches), NULL
    keys / lists / children -- the match set is always the closed form, every staged row is staged once, no slot leaks."""
    rng = np.random.default_rng(20251105)
    for case in range(60):
        piece = int(rng.choice([16, 64, 128]))
        cap_outputs = int(rng.choice([8, 40, 400, 4096]))
        slots = int(rng.integers(1, 4))
        n_batches = int(rng.integers(1, 9))
        mixed = rng.random() < 0.4
        nulls = rng.random() < 0.5
        parts, expect, base, staged = [], [], 0, 0
        for b in range(n_batches):
            n = int(rng.choice([0, 1, 5, int(rng.integers(1, 3 * piece + 5))]))
            keys, lists, e = make_rows(n, 1000 * case + b, p_match=0.3, key_sizes=(64, 33) if mixed else (64,),
                                       p_null_key=0.1 if nulls else 0.0, p_null_list=0.1 if nulls else 0.0,
                                       p_null_child=0.15 if nulls else 0.0, p_bad_len=0.05 if mixed else 0.0)
            parts.append(batch_of(keys, lists))
            expect.extend((e + base).tolist())
            staged += sum(k is not None and len(k) in (33, 64) for k in keys)
            base += n
        with A.query(piece=piece, cap_rows=piece, cap_outputs=cap_outputs, slots=slots) as q:
            reader = pa.RecordBatchReader.from_batches(parts[0].schema, parts)
            ids, seen = A.arrow.scan_stream(q.h, reader, _lib=A.lib)
            st = q.stats()
        assert ids.tolist() == expect and seen == base, (case, piece, cap_outputs, slots)
        assert st["rows"] == staged, (case, st, staged)
        if not mixed and not nulls and cap_outputs == 4096:
            assert st["submits"] == -(-staged // piece), (case, st, staged, piece)     # whole pieces across batches
