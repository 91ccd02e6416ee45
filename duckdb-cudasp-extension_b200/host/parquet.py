"""Parquet files straight into the C ABI: include/cudasp_parquet.h (csrc/parquet_ingest.cpp decodes the pages itself, one row group
per host thread in flight).  For files outside its subset (other codecs than none / Snappy, other shapes) the library answers
CUDASP_ERR_UNSUPPORTED and `scan_parquet_files(..., fallback=True)` goes through pyarrow's reader instead (host/arrow.py).

    rows = scan_parquet_files(q, ["a.parquet", "b.parquet"])     # [(file_index, row, txid, height, tweak_key)]
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence, Tuple

from .binding import CudaspError, Query, lib as _lib

ERR_UNSUPPORTED, ERR_IO = -5, -6
PARQUET_EXPORTS = ["cudasp_scan_parquet", "cudasp_parquet_result_free"]


class ParquetResult(C.Structure):      # include/cudasp_parquet.h cudasp_parquet_result
    _fields_ = [("n_matches", C.c_uint64), ("rows_scanned", C.c_uint64), ("file_index", C.POINTER(C.c_uint32)),
                ("row_in_file", C.POINTER(C.c_uint64)), ("heights", C.POINTER(C.c_int32)), ("txids", C.POINTER(C.c_uint8)),
                ("txid_offsets", C.POINTER(C.c_uint64)), ("tweak_keys", C.POINTER(C.c_uint8)), ("key_offsets", C.POINTER(C.c_uint64))]


def _bind(L):
    if not getattr(L, "_parquet_bound", False):
        L.cudasp_scan_parquet.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.c_int, C.POINTER(C.POINTER(ParquetResult))]
        L.cudasp_parquet_result_free.argtypes = [C.POINTER(ParquetResult)]
        L.cudasp_parquet_result_free.restype = None
        L._parquet_bound = True
    return L


def scan_parquet_files(q: Query, paths: Sequence[str], threads: int = 0, fallback: bool = False, with_rows: bool = False):
    """-> [(file_index, row_in_file, txid, height, tweak_key)] of the matching rows, file order then row order
    (and the number of rows scanned when with_rows)."""
    L = _bind(_lib())
    arr = (C.c_char_p * max(len(paths), 1))(*[p.encode() for p in paths])
    out = C.POINTER(ParquetResult)()
    rc = L.cudasp_scan_parquet(q.handle, arr, len(paths), int(threads), C.byref(out))
    if rc == ERR_UNSUPPORTED and fallback:
        import pyarrow.parquet as pq
        from .arrow import scan_table
        res: List[Tuple[int, int, bytes, int, bytes]] = []
        total = 0
        for fi, p in enumerate(paths):
            t = pq.read_table(p, columns=["txid", "height", "tweak_key", "outputs"])
            total += t.num_rows
            for i in scan_table(q, t).tolist():
                txid, height = t["txid"][int(i)].as_py(), t["height"][int(i)].as_py()
                if txid is not None and height is not None:        # the reference drops such rows before the scan
                    res.append((fi, int(i), txid, height, t["tweak_key"][int(i)].as_py()))
        return (res, total) if with_rows else res
    if rc:
        raise CudaspError(rc, (L.cudasp_last_error(q.handle) or b"").decode())
    r = out.contents
    try:
        res = []
        for i in range(r.n_matches):
            t0, t1, k0, k1 = r.txid_offsets[i], r.txid_offsets[i + 1], r.key_offsets[i], r.key_offsets[i + 1]
            res.append((int(r.file_index[i]), int(r.row_in_file[i]), bytes(r.txids[t0:t1]), int(r.heights[i]), bytes(r.tweak_keys[k0:k1])))
        return (res, int(r.rows_scanned)) if with_rows else res
    finally:
        L.cudasp_parquet_result_free(out)
