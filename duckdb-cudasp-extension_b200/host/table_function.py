"""Host-side mirror of the reference's DuckDB table function `cudasp_scan` above the C ABI.

Same names, argument meaning, NULL handling and error texts as
/root/reference/src/cudasp_extension.cu (Bind :420-509, AccumulateInput :154-224, ProcessBatch
:226-410, in-out / final :534-624), so the parity tests read like test/sql/cudasp.test.  The real
DuckDB shim (C++, same C ABI) is in ../duckdb_ext/; DuckDB itself is not available to Python in
this image, so this module is what the test-suite drives.

    cudasp_scan(rows, scan_private_key, spend_public_key, label_keys, batch_size=300000)
        rows: iterable of (txid BLOB, height INTEGER, tweak_key BLOB, outputs BIGINT[]) with None = NULL
        -> list of (txid, height, tweak_key) for matching rows
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

from .binding import KEYS_SEC1_33, KEYS_XY64, CudaspError, Query

TWEAK_KEY_SIZE = 64   # CudaspScanBindData::TWEAK_KEY_SIZE  (src/cudasp_extension.cu:90)
COMPRESSED_KEY_SIZE = 33   # extension: SEC1 keys as the index stores them, decompressed on the GPU
SCALAR_SIZE = 32      # CudaspScanBindData::SCALAR_SIZE     (:91)
STANDARD_VECTOR_SIZE = 2048


class InvalidInputException(ValueError):
    """duckdb::InvalidInputException stand-in."""


class IOException(RuntimeError):
    """Raised for GPU failures (the reference swallows them, src/cudasp_extension.cu:325,378)."""


class CudaspScanBindData:
    def __init__(self, scan_private_key: bytes, spend_public_key: bytes, label_keys: List[bytes], batch_size: int,
                 details: bool = False):
        self.scan_private_key_data = scan_private_key
        self.spend_public_key_data = spend_public_key
        self.label_keys_data = label_keys
        self.batch_size = batch_size
        self.details = details
        self.return_types = ["BLOB", "INTEGER", "BLOB"]
        self.names = ["txid", "height", "tweak_key"]
        if details:
            # extension (SURVEY.md 8f rank 3+4): one result row per found output instead of one per transaction
            self.return_types += ["UINTEGER", "UINTEGER", "INTEGER", "BLOB"]
            self.names += ["k", "output_index", "label_index", "t_k"]


def CudaspScanBind(inputs: Sequence, named_parameters: Optional[dict] = None) -> CudaspScanBindData:
    """src/cudasp_extension.cu:420-509 -- same checks, same messages."""
    named_parameters = named_parameters or {}
    if len(inputs) != 4:
        raise InvalidInputException("cudasp_scan requires 4 arguments: TABLE, scan_private_key BLOB, spend_public_key BLOB, and label_keys LIST[BLOB]")
    _, scan_private_key, spend_public_key, label_keys = inputs
    if not isinstance(scan_private_key, (bytes, bytearray)):
        raise InvalidInputException("Second argument must be a BLOB (32-byte scan_private_key)")
    if len(scan_private_key) != SCALAR_SIZE:
        raise InvalidInputException("scan_private_key must be exactly 32 bytes, got %d bytes" % len(scan_private_key))
    if not isinstance(spend_public_key, (bytes, bytearray)):
        raise InvalidInputException("Third argument must be a BLOB (64-byte spend_public_key)")
    if len(spend_public_key) != TWEAK_KEY_SIZE:
        raise InvalidInputException("spend_public_key must be exactly 64 bytes, got %d bytes" % len(spend_public_key))
    if not isinstance(label_keys, (list, tuple)):
        raise InvalidInputException("Fourth argument must be a LIST[BLOB] (label keys)")
    labels = []
    for k in label_keys:
        if not isinstance(k, (bytes, bytearray)):
            raise InvalidInputException("All elements in label_keys must be BLOBs")
        if len(k) != TWEAK_KEY_SIZE:
            raise InvalidInputException("Each label key must be exactly 64 bytes, got %d bytes" % len(k))
        labels.append(bytes(k))
    batch_size = 300000
    if "batch_size" in named_parameters:
        v = named_parameters["batch_size"]
        if isinstance(v, bool) or not isinstance(v, (int, np.integer)):
            raise InvalidInputException("batch_size parameter must be an INTEGER")
        if v <= 0:
            raise InvalidInputException("batch_size must be positive, got %d" % v)
        if v > 10000000:
            raise InvalidInputException("batch_size too large (max 10,000,000), got %d" % v)
        batch_size = int(v)
    details = False
    if "details" in named_parameters:
        if not isinstance(named_parameters["details"], (bool, np.bool_)):
            raise InvalidInputException("details parameter must be a BOOLEAN")
        details = bool(named_parameters["details"])
    return CudaspScanBindData(bytes(scan_private_key), bytes(spend_public_key), labels, batch_size, details)


class CudaspScanLocalState:
    """Per-thread accumulation (src/cudasp_extension.cu:106-120), columnar instead of std::string per row."""

    def __init__(self):
        self.accumulated_txids: List[bytes] = []
        self.accumulated_heights: List[int] = []
        self.accumulated_tweak_keys: List[bytes] = []
        self.accumulated_outputs: List[int] = []
        self.accumulated_output_offsets: List[int] = [0]
        self.finalized = False


class CudaspScanState:
    """Global state (src/cudasp_extension.cu:122-152): owns the C-ABI query handle and the result vectors."""

    def __init__(self, bind_data: CudaspScanBindData, devices: Optional[Sequence[int]] = None):
        try:
            self.query = Query(bind_data.scan_private_key_data, bind_data.spend_public_key_data, bind_data.label_keys_data,
                               devices=devices, batch_rows=bind_data.batch_size)
            if bind_data.details:
                self.query.enable_details()
        except CudaspError as e:
            raise IOException(str(e))
        self.output_details: List[Tuple] = []      # details := true: (k, output_index, label_index, t_k) per result row
        self.output_txids: List[bytes] = []
        self.output_heights: List[int] = []
        self.output_tweak_keys: List[bytes] = []
        self.output_position = 0

    def close(self):
        self.query.close()


def AccumulateInput(local_state: CudaspScanLocalState, chunk: Iterable[Tuple]) -> None:
    """src/cudasp_extension.cu:154-224: drop rows with NULL txid/height/tweak_key; NULL list -> length 0;
    NULL children skipped."""
    for txid, height, tweak_key, outputs in chunk:
        if txid is None or height is None or tweak_key is None:
            continue
        local_state.accumulated_txids.append(bytes(txid))
        local_state.accumulated_heights.append(int(height))
        local_state.accumulated_tweak_keys.append(bytes(tweak_key))
        if outputs is not None:
            local_state.accumulated_outputs.extend(int(o) for o in outputs if o is not None)
        local_state.accumulated_output_offsets.append(len(local_state.accumulated_outputs))


def ProcessBatch(local_state: CudaspScanLocalState, bind_data: CudaspScanBindData, global_state: CudaspScanState) -> None:
    """src/cudasp_extension.cu:226-410: one C-ABI call instead of LaunchBatchScan/RunBatchScanKernels/flag scan."""
    n = len(local_state.accumulated_txids)
    if n == 0:
        return
    # rows are scanned per key format: 64-byte keys the reference's way, 33-byte SEC1 keys decompressed on the
    # GPU (what scripts/decompress_tweak_key.py does on the CPU before the table is loaded); a tweak_key of any
    # other length can never match (the reference reads out of bounds there) and is skipped
    offs = np.array(local_state.accumulated_output_offsets, dtype=np.int64)
    outs = np.array(local_state.accumulated_outputs, dtype=np.int64)
    lens = np.fromiter((len(t) for t in local_state.accumulated_tweak_keys), dtype=np.int64, count=n)
    matched: List[int] = []
    details = {}                                   # source row -> [(k, output_index, label_index, t_k)]
    for size, fmt in ((TWEAK_KEY_SIZE, KEYS_XY64), (COMPRESSED_KEY_SIZE, KEYS_SEC1_33)):
        sel = np.nonzero(lens == size)[0]
        if len(sel) == 0:
            continue
        tw = np.frombuffer(b"".join(local_state.accumulated_tweak_keys[i] for i in sel), np.uint8).reshape(len(sel), size)
        if len(sel) == n:
            sub_out, sub_off = outs, offs
        else:
            cnt = offs[sel + 1] - offs[sel]
            sub_off = np.concatenate([[0], np.cumsum(cnt)])
            sub_out = np.concatenate([outs[offs[i]: offs[i + 1]] for i in sel]) if len(sel) else outs[:0]
        try:
            if bind_data.details:
                ids, recs = global_state.query.scan_batch_details(tw, sub_out, sub_off.astype(np.uint32), key_format=fmt)
                for row, k, oi, li, t in recs:
                    details.setdefault(int(sel[row]), []).append((k, oi, li, t.to_bytes(32, "big")))
            else:
                ids = global_state.query.scan_batch(tw, sub_out, sub_off.astype(np.uint32), key_format=fmt)
        except CudaspError as e:
            raise IOException(str(e))
        matched.extend(int(sel[i]) for i in ids.tolist())
    for i in sorted(matched):
        for rec in (details.get(i, []) if bind_data.details else [None]):
            global_state.output_txids.append(local_state.accumulated_txids[i])
            global_state.output_heights.append(local_state.accumulated_heights[i])
            global_state.output_tweak_keys.append(local_state.accumulated_tweak_keys[i])
            if rec is not None:
                global_state.output_details.append(rec)
    local_state.__init__()


def CudaspScanFunction(bind_data, global_state, local_state, chunk) -> str:
    """in_out_function (src/cudasp_extension.cu:534-552): accumulate, cut a batch when full, emit nothing."""
    AccumulateInput(local_state, chunk)
    if len(local_state.accumulated_txids) >= bind_data.batch_size:
        ProcessBatch(local_state, bind_data, global_state)
    return "NEED_MORE_INPUT"


def CudaspScanFinalFunction(bind_data, global_state, local_state):
    """in_out_function_final (src/cudasp_extension.cu:554-624): flush, then stream the results out in
    STANDARD_VECTOR_SIZE chunks."""
    if not local_state.finalized:
        ProcessBatch(local_state, bind_data, global_state)
        local_state.finalized = True
    while global_state.output_position < len(global_state.output_txids):
        a = global_state.output_position
        b = min(a + STANDARD_VECTOR_SIZE, len(global_state.output_txids))
        global_state.output_position = b
        base = list(zip(global_state.output_txids[a:b], global_state.output_heights[a:b], global_state.output_tweak_keys[a:b]))
        if bind_data.details:
            base = [r + d for r, d in zip(base, global_state.output_details[a:b])]
        yield base


def cudasp_scan(rows: Iterable[Tuple], scan_private_key, spend_public_key, label_keys, devices=None, **named_parameters):
    """SELECT * FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM rows), sk, spend, labels, batch_size := N)"""
    bind_data = CudaspScanBind([rows, scan_private_key, spend_public_key, label_keys], named_parameters)
    global_state = CudaspScanState(bind_data, devices)
    local_state = CudaspScanLocalState()
    try:
        chunk = []
        for r in rows:
            chunk.append(r)
            if len(chunk) == STANDARD_VECTOR_SIZE:
                CudaspScanFunction(bind_data, global_state, local_state, chunk)
                chunk = []
        if chunk:
            CudaspScanFunction(bind_data, global_state, local_state, chunk)
        out = []
        for part in CudaspScanFinalFunction(bind_data, global_state, local_state):
            out.extend(part)
        return out
    finally:
        global_state.close()
