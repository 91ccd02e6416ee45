"""Mutation fuzzing of the Parquet front-end (csrc/parquet_ingest.cpp) against the staging double: valid files written with every
writer setting of tests/test_parquet_ingest.py are damaged (byte flips in the footer, near page headers, anywhere; truncations;
runs of extreme bytes; a wrong footer length) and scanned.  Every outcome is fine -- a result, CUDASP_ERR_IO, CUDASP_ERR_UNSUPPORTED
-- except a crash or an out-of-bounds access, which is why tests/test_parquet_fuzz.py runs this in a process of its own (and why it
is worth running under -fsanitize=address,undefined now and then: see the recipe there).
usage: parquet_fuzz.py <libparquetcheck.so> <iterations> <seed> <scratch dir>"""
import ctypes as C, os, sys, numpy as np, pyarrow as pa, pyarrow.parquet as pq
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import test_parquet_ingest as T
so = sys.argv[1]; n_iter = int(sys.argv[2]); seed = int(sys.argv[3])
L = C.CDLL(so)
L.mock_query_create.restype = C.c_void_p
L.mock_query_create.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32]
L.mock_query_destroy.argtypes = [C.c_void_p]
L.cudasp_scan_parquet.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.c_int, C.POINTER(C.POINTER(T.Result))]
L.cudasp_parquet_result_free.argtypes = [C.POINTER(T.Result)]
rng = np.random.default_rng(seed)
d = sys.argv[4]
bases = []
for si, st in enumerate(T.WRITER_SETTINGS):
    for nulls in (0.0, 0.1):
        rows, _ = T.make_rows(300, 7 + si, p_null=nulls, key_sizes=(64, 64, 33, 12))
        p = os.path.join(d, 'base_%d_%d.parquet' % (si, int(nulls * 10)))
        pq.write_table(T.table_of(rows), p, row_group_size=120, **st)
        bases.append(open(p, 'rb').read())
    rows, _ = T.make_rows(300, 90 + si)
    p = os.path.join(d, 'basef_%d.parquet' % si)
    pq.write_table(T.table_of(rows, key_type=pa.binary(64), txid_type=pa.binary(32)), p, row_group_size=120, **st)
    bases.append(open(p, 'rb').read())
stats = {}
mut = os.path.join(d, 'mut_%d.parquet' % seed)
for it in range(n_iter):
    b = bytearray(bases[int(rng.integers(len(bases)))])
    kind = int(rng.integers(6))
    n = len(b)
    if kind == 0:      # flips in the footer
        flen = int.from_bytes(b[-8:-4], 'little')
        lo = max(0, n - 8 - flen)
        for _ in range(int(rng.integers(1, 4))):
            b[int(rng.integers(lo, n - 4))] = int(rng.integers(256))
    elif kind == 1:    # flips anywhere
        for _ in range(int(rng.integers(1, 6))):
            b[int(rng.integers(n))] = int(rng.integers(256))
    elif kind == 2:    # truncate the middle
        a = int(rng.integers(4, n - 8)); c = int(rng.integers(a, n - 8))
        b = b[:a] + b[c:]
    elif kind == 3:    # extreme bytes in a short run
        a = int(rng.integers(4, n - 16))
        v = int(rng.choice([0, 0xff, 0x80, 0x7f]))
        for k in range(int(rng.integers(1, 12))):
            b[a + k] = v
    elif kind == 4:    # footer length field
        b[-8:-4] = int(rng.integers(0, 1 << 32)).to_bytes(4, 'little')
    else:              # flips in the first 2 KB after a random page-ish offset (headers + levels live at chunk starts)
        a = int(rng.integers(4, n - 8))
        for _ in range(int(rng.integers(1, 5))):
            b[min(n - 9, a + int(rng.integers(0, 64)))] = int(rng.integers(256))
    open(mut, 'wb').write(bytes(b))
    q = L.mock_query_create(97, 128, 1024, 3)
    arr = (C.c_char_p * 1)(mut.encode())
    out = C.POINTER(T.Result)()
    rc = L.cudasp_scan_parquet(q, arr, 1, 1, C.byref(out))
    stats[rc] = stats.get(rc, 0) + 1
    if rc == 0:
        L.cudasp_parquet_result_free(out)
    L.mock_query_destroy(q)
print('ok', stats)
