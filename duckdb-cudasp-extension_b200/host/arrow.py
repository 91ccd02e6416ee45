"""Columnar front door of the hot path: pyarrow RecordBatch / Table / RecordBatchReader -> include/cudasp_arrow.h.

The Arrow buffers are handed to the C ABI through the Arrow C data interface (no copies on the Python side); the
library stages them into pinned memory piece by piece and returns the indices of the matching rows, which are then
gathered with `take` -- txid / height never travel (compare /root/reference/src/cudasp_extension.cu:154-224, where
every column of every row is copied twice before the GPU sees it).

    ids = scan_record_batch(q, batch)                      # np.uint64 row indices
    ids = scan_table(q, table)                             # streams all chunks, one drain at the end
    out = cudasp_scan_arrow(table, sk, spend, labels)      # pyarrow.Table (txid, height, tweak_key) of the matches
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from .binding import CudaspError, Query, lib as _default_lib

vp = C.c_void_p


class ArrowSchema(C.Structure):
    pass


class ArrowArray(C.Structure):
    pass


class ArrowArrayStream(C.Structure):
    pass


ArrowSchema._fields_ = [("format", C.c_char_p), ("name", C.c_char_p), ("metadata", vp), ("flags", C.c_int64),
                        ("n_children", C.c_int64), ("children", C.POINTER(C.POINTER(ArrowSchema))), ("dictionary", vp),
                        ("release", C.CFUNCTYPE(None, C.POINTER(ArrowSchema))), ("private_data", vp)]
ArrowArray._fields_ = [("length", C.c_int64), ("null_count", C.c_int64), ("offset", C.c_int64), ("n_buffers", C.c_int64),
                       ("n_children", C.c_int64), ("buffers", C.POINTER(vp)), ("children", C.POINTER(C.POINTER(ArrowArray))),
                       ("dictionary", vp), ("release", C.CFUNCTYPE(None, C.POINTER(ArrowArray))), ("private_data", vp)]
ArrowArrayStream._fields_ = [("get_schema", vp), ("get_next", vp), ("get_last_error", vp),
                             ("release", C.CFUNCTYPE(None, C.POINTER(ArrowArrayStream))), ("private_data", vp)]

ARROW_EXPORTS = ["cudasp_scan_arrow_columns", "cudasp_scan_arrow", "cudasp_scan_arrow_stream", "cudasp_free"]
_typed = set()


def _typed_lib(L=None):
    L = L or _default_lib()
    if id(L) not in _typed:
        u64p = C.POINTER(C.c_uint64)
        L.cudasp_scan_arrow_columns.argtypes = [vp, C.POINTER(ArrowSchema), C.POINTER(ArrowArray), C.POINTER(ArrowSchema),
                                                C.POINTER(ArrowArray), vp, C.c_uint64, u64p]
        L.cudasp_scan_arrow.argtypes = [vp, C.POINTER(ArrowSchema), C.POINTER(ArrowArray), C.c_int, C.c_int, vp, C.c_uint64, u64p]
        L.cudasp_scan_arrow_stream.argtypes = [vp, C.POINTER(ArrowArrayStream), C.c_int, C.c_int, C.POINTER(u64p), u64p, u64p]
        L.cudasp_free.argtypes = [vp]
        L.cudasp_free.restype = None
        L.cudasp_last_error.restype = C.c_char_p
        L.cudasp_last_error.argtypes = [vp]
        _typed.add(id(L))
    return L


def _release(struct):
    if struct.release:
        struct.release(C.byref(struct))


def _raise(L, handle, rc):
    msg = L.cudasp_last_error(handle)
    raise CudaspError(rc, msg.decode() if msg else "error %d" % rc)


def _handle(q):
    return q.handle if isinstance(q, Query) else q


def _col_index(schema, col):
    return schema.get_field_index(col) if isinstance(col, str) else int(col)


def scan_record_batch(q, batch, tweak_col="tweak_key", outputs_col="outputs", _lib=None) -> np.ndarray:
    """One pyarrow.RecordBatch (or StructArray; a NULL struct row is dropped) -> ascending uint64 indices of its
    matching rows."""
    L = _typed_lib(_lib)
    schema = batch.schema if hasattr(batch, "schema") else batch.type
    ti, oi = _col_index(schema, tweak_col), _col_index(schema, outputs_col)
    if ti < 0 or oi < 0:
        raise KeyError("column not found")
    rows = len(batch)
    sc, ar = ArrowSchema(), ArrowArray()
    batch._export_to_c(C.addressof(ar), C.addressof(sc))
    try:
        ids = np.zeros(max(rows, 1), np.uint64)
        n = C.c_uint64(0)
        rc = L.cudasp_scan_arrow(_handle(q), C.byref(sc), C.byref(ar), ti, oi, ids.ctypes.data, rows, C.byref(n))
        if rc:
            _raise(L, _handle(q), rc)
        return ids[: n.value].copy()
    finally:
        _release(ar)
        _release(sc)


def scan_columns(q, tweak_key, outputs, _lib=None) -> np.ndarray:
    """Two pyarrow Arrays of equal length (any slicing) -> ascending uint64 indices of the matching rows."""
    L = _typed_lib(_lib)
    ks, ka, os_, oa = ArrowSchema(), ArrowArray(), ArrowSchema(), ArrowArray()
    tweak_key._export_to_c(C.addressof(ka), C.addressof(ks))
    outputs._export_to_c(C.addressof(oa), C.addressof(os_))
    try:
        ids = np.zeros(max(len(tweak_key), 1), np.uint64)
        n = C.c_uint64(0)
        rc = L.cudasp_scan_arrow_columns(_handle(q), C.byref(ks), C.byref(ka), C.byref(os_), C.byref(oa), ids.ctypes.data,
                                         len(tweak_key), C.byref(n))
        if rc:
            _raise(L, _handle(q), rc)
        return ids[: n.value].copy()
    finally:
        for s in (ka, ks, oa, os_):
            _release(s)


def scan_stream(q, reader, tweak_col="tweak_key", outputs_col="outputs", _lib=None):
    """pyarrow.RecordBatchReader (Parquet / IPC / dataset scanner ...) -> (uint64 row indices over the stream, rows seen)."""
    L = _typed_lib(_lib)
    ti, oi = _col_index(reader.schema, tweak_col), _col_index(reader.schema, outputs_col)
    if ti < 0 or oi < 0:
        raise KeyError("column not found")
    st = ArrowArrayStream()
    reader._export_to_c(C.addressof(st))
    try:
        out = C.POINTER(C.c_uint64)()
        n, seen = C.c_uint64(0), C.c_uint64(0)
        rc = L.cudasp_scan_arrow_stream(_handle(q), C.byref(st), ti, oi, C.byref(out), C.byref(n), C.byref(seen))
        if rc:
            _raise(L, _handle(q), rc)
        ids = np.ctypeslib.as_array(out, shape=(n.value,)).copy() if n.value else np.zeros(0, np.uint64)
        L.cudasp_free(out)
        return ids.astype(np.uint64), int(seen.value)
    finally:
        _release(st)


def scan_table(q, table, tweak_col="tweak_key", outputs_col="outputs", _lib=None) -> np.ndarray:
    """pyarrow.Table (any chunking) -> uint64 row indices; the chunks are streamed, the device queue drains once."""
    import pyarrow as pa
    reader = pa.RecordBatchReader.from_batches(table.schema, table.to_batches())
    ids, seen = scan_stream(q, reader, tweak_col, outputs_col, _lib)
    assert seen == table.num_rows
    return ids


def parquet_batches(path, columns=("tweak_key", "outputs"), threads: int = 4, lookahead: int = 8):
    """(schema, generator of RecordBatches) over a Parquet file, row groups decoded by `threads` workers ahead of the
    consumer (pyarrow releases the GIL while decoding), delivered in file order."""
    import collections
    from concurrent.futures import ThreadPoolExecutor
    import pyarrow as pa
    import pyarrow.parquet as pq
    pf = pq.ParquetFile(path)
    cols = list(columns)
    schema = pf.schema_arrow
    schema = pa.schema([schema.field(c) for c in cols])

    def gen():
        with ThreadPoolExecutor(max_workers=max(1, threads)) as pool:
            def read(i):
                return pq.ParquetFile(path).read_row_group(i, columns=cols, use_threads=False)
            pending = collections.deque()
            nxt = 0
            while nxt < pf.num_row_groups or pending:
                while nxt < pf.num_row_groups and len(pending) < lookahead:
                    pending.append(pool.submit(read, nxt))
                    nxt += 1
                for b in pending.popleft().result().to_batches():
                    yield b
    return schema, gen()


def scan_parquet(q, path, tweak_col="tweak_key", outputs_col="outputs", threads: int = 4, _lib=None):
    """Parquet file -> (uint64 row indices over the file, rows seen): decode on `threads` workers, staging + GPU scan on
    the calling thread, all overlapped (the sibling `cudasp_scan_parquet(path, ...)` of SURVEY.md 8f)."""
    import pyarrow as pa
    schema, batches = parquet_batches(path, (tweak_col, outputs_col), threads)
    reader = pa.RecordBatchReader.from_batches(schema, batches)
    return scan_stream(q, reader, tweak_col, outputs_col, _lib)


def cudasp_scan_arrow(table, scan_private_key: bytes, spend_public_key: bytes, label_keys: Sequence[bytes] = (),
                      devices: Optional[Sequence[int]] = None, batch_size: int = 300000):
    """The table function over Arrow data: same arguments, checks and result columns as cudasp_scan
    (/root/reference/src/cudasp_extension.cu:420-509,592-623); `table` needs txid, height, tweak_key, outputs."""
    import pyarrow.compute as pc
    from .table_function import CudaspScanBind, IOException
    bind = CudaspScanBind([table, scan_private_key, spend_public_key, list(label_keys)], {"batch_size": batch_size})
    try:
        with Query(bind.scan_private_key_data, bind.spend_public_key_data, bind.label_keys_data, devices=devices,
                   batch_rows=bind.batch_size) as q:
            ids = scan_table(q, table)
    except CudaspError as e:
        raise IOException(str(e))
    out = table.select(["txid", "height", "tweak_key"]).take(ids)
    # rows with NULL txid / height are dropped by the reference before the scan (:195-197)
    keep = pc.and_(pc.is_valid(out["txid"]), pc.is_valid(out["height"]))
    return out.filter(keep)
