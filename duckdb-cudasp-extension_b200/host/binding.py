"""ctypes binding of libcudasp_b200.so (include/cudasp.h).  No torch types cross this boundary.

The library is the product; this module only marshals pointers.  If the shared object is missing
or no CUDA device is present, calls fail loudly (RuntimeError) -- there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# CUDASP_B200_LIB: a differently-built copy of the same library (tuning builds, tools/build_variants.sh)
LIB_PATH = os.environ.get("CUDASP_B200_LIB") or os.path.join(os.path.dirname(_HERE), "libcudasp_b200.so")
_lib = None

OK, ERR_ARG, ERR_CUDA, ERR_STATE, ERR_KEY = 0, -1, -2, -3, -4

u8p, i64p, u32p = C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.POINTER(C.c_uint32)


class Slot(C.Structure):
    _fields_ = [("slot_id", C.c_uint32), ("device", C.c_int32), ("tweak_keys", u8p), ("outputs", i64p),
                ("out_offsets", u32p), ("cap_rows", C.c_uint32), ("cap_outputs", C.c_uint64)]


class Result(C.Structure):
    _fields_ = [("tag", C.c_uint64), ("slot_id", C.c_uint32), ("rows", C.c_uint32), ("n_matches", C.c_uint32),
                ("flags", C.c_uint32), ("match_ids", u32p)]


class MatchDetail(C.Structure):      # include/cudasp.h cudasp_match_detail
    _fields_ = [("row", C.c_uint32), ("k", C.c_uint32), ("output_index", C.c_uint32), ("label_index", C.c_int32),
                ("t_k", C.c_uint8 * 32)]


DETAIL_DTYPE = np.dtype([("row", "<u4"), ("k", "<u4"), ("output_index", "<u4"), ("label_index", "<i4"), ("t_k", "V32")])
KEYS_XY64, KEYS_SEC1_33 = 0, 1          # include/cudasp.h CUDASP_KEYS_*
KEY_BYTES = {KEYS_XY64: 64, KEYS_SEC1_33: 33}

EXPORTS = ["cudasp_device_count", "cudasp_last_error", "cudasp_query_create", "cudasp_query_destroy", "cudasp_scan_batch",
           "cudasp_acquire_slot", "cudasp_submit_slot", "cudasp_poll", "cudasp_wait", "cudasp_release_result",
           "cudasp_scan_device", "cudasp_profile_stages", "cudasp_scan_candidates", "cudasp_scan_intermediates", "cudasp_test_field",
           "cudasp_launch_count", "cudasp_query_info", "cudasp_version", "cudasp_query_create_ex", "cudasp_try_acquire_slot", "cudasp_release_cached_memory",
           "cudasp_scan_batch_fmt", "cudasp_submit_slot_fmt", "cudasp_scan_candidates_fmt",
           "cudasp_query_enable_details", "cudasp_result_details", "cudasp_scan_batch_details", "cudasp_debug_gtab",
           "cudasp_device_sm_count", "cudasp_scan_batch_mt", "cudasp_transfer_bytes", "cudasp_batch_inverse_static"]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libcudasp_b200.so is not built (run __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.cudasp_device_count.restype = C.c_int
        L.cudasp_last_error.restype = C.c_char_p
        L.cudasp_last_error.argtypes = [vp]
        L.cudasp_version.restype = C.c_char_p
        L.cudasp_query_create.argtypes = [u8p, u8p, u8p, C.c_uint32, C.POINTER(C.c_int), C.c_int, C.c_uint32, C.c_uint64, C.POINTER(vp)]
        L.cudasp_query_destroy.argtypes = [vp]
        L.cudasp_query_destroy.restype = None
        L.cudasp_scan_batch.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, u32p]
        L.cudasp_acquire_slot.argtypes = [vp, C.POINTER(Slot)]
        L.cudasp_submit_slot.argtypes = [vp, C.POINTER(Slot), C.c_uint32, C.c_uint64, C.c_uint64]
        L.cudasp_poll.argtypes = [vp, C.POINTER(Result)]
        L.cudasp_wait.argtypes = [vp, C.POINTER(Result)]
        L.cudasp_release_result.argtypes = [vp, C.POINTER(Result)]
        L.cudasp_scan_device.argtypes = [vp, C.c_int, vp, vp, vp, C.c_uint32, vp, vp, vp]
        L.cudasp_profile_stages.argtypes = [vp, C.c_int, vp, vp, vp, C.c_uint32, vp, vp, C.POINTER(C.c_float)]
        L.cudasp_scan_candidates.argtypes = [vp, vp, C.c_uint32, vp]
        L.cudasp_scan_batch_fmt.argtypes = [vp, vp, C.c_int, vp, vp, C.c_uint32, vp, u32p]
        L.cudasp_submit_slot_fmt.argtypes = [vp, C.POINTER(Slot), C.c_uint32, C.c_uint64, C.c_uint64, C.c_int]
        L.cudasp_scan_candidates_fmt.argtypes = [vp, vp, C.c_int, C.c_uint32, vp, vp]
        L.cudasp_query_enable_details.argtypes = [vp, C.c_uint32]
        L.cudasp_result_details.argtypes = [vp, C.POINTER(Result), C.POINTER(C.POINTER(MatchDetail)), u32p]
        L.cudasp_scan_batch_details.argtypes = [vp, vp, C.c_int, vp, vp, C.c_uint32, vp, u32p, C.POINTER(MatchDetail), C.c_uint32, u32p]
        L.cudasp_scan_intermediates.argtypes = [vp, vp, C.c_uint32, vp, vp, vp]
        L.cudasp_test_field.argtypes = [C.c_int, C.c_int, vp, vp, vp, C.c_uint32]
        L.cudasp_debug_gtab.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, vp]
        L.cudasp_device_sm_count.argtypes = [C.c_int]
        L.cudasp_scan_batch_mt.argtypes = [vp, vp, C.c_int, vp, vp, C.c_uint32, vp, u32p, C.c_int]
        L.cudasp_transfer_bytes.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.cudasp_transfer_bytes.restype = None
        L.cudasp_batch_inverse_static.argtypes = [C.c_int, vp, vp, C.c_uint32, vp]
        L.cudasp_launch_count.argtypes = [vp]
        L.cudasp_launch_count.restype = C.c_uint64
        L.cudasp_query_info.argtypes = [vp] + [C.POINTER(C.c_int)] * 5
        _lib = L
    return _lib


class CudaspError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("cudasp error %d: %s" % (code, msg))
        self.code = code


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Query:
    """Owns one cudasp_query handle (keys recoded once, staging allocated once)."""

    def __init__(self, scan_sk: bytes, spend_pk: bytes, label_keys: Sequence[bytes] = (), devices: Optional[Sequence[int]] = None,
                 batch_rows: int = 300000, max_outputs_per_batch: int = 0):
        L = lib()
        self._h = C.c_void_p()
        labels = b"".join(label_keys)
        lab = (C.c_uint8 * max(len(labels), 1)).from_buffer_copy(labels or b"\0")
        sk = (C.c_uint8 * len(scan_sk)).from_buffer_copy(scan_sk)
        sp = (C.c_uint8 * len(spend_pk)).from_buffer_copy(spend_pk)
        if len(scan_sk) != 32 or len(spend_pk) != 64 or any(len(k) != 64 for k in label_keys):
            raise CudaspError(ERR_ARG, "key sizes: scan 32 B, spend 64 B, labels 64 B each")
        dv = None
        nd = 0
        if devices is not None:
            dv = (C.c_int * len(devices))(*devices)
            nd = len(devices)
        rc = L.cudasp_query_create(sk, sp, lab, len(label_keys), dv, nd, batch_rows, max_outputs_per_batch, C.byref(self._h))
        if rc:
            raise CudaspError(rc, (L.cudasp_last_error(None) or b"").decode())
        self.n_labels = len(label_keys)
        self.batch_rows = batch_rows

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().cudasp_query_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc < 0:
            raise CudaspError(rc, (lib().cudasp_last_error(self._h) or b"").decode())
        return rc

    @property
    def handle(self):
        return self._h

    def info(self):
        v = [C.c_int() for _ in range(5)]
        self._check(lib().cudasp_query_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(["n_devices", "ladder_dbl", "ladder_add", "fixed_window_bits", "n_labels"], [x.value for x in v]))

    def launch_count(self) -> int:
        return int(lib().cudasp_launch_count(self._h))

    def transfer_bytes(self):
        """(host->device, device->host) bytes this query has copied so far."""
        a, b = C.c_uint64(0), C.c_uint64(0)
        lib().cudasp_transfer_bytes(self._h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    def scan_batch(self, tweaks: np.ndarray, outputs: Optional[np.ndarray], out_offsets: Optional[np.ndarray],
                   key_format: int = KEYS_XY64, host_threads: Optional[int] = None) -> np.ndarray:
        """tweaks uint8 [rows,64] (or [rows,33] with key_format=KEYS_SEC1_33); outputs int64 [n]; out_offsets uint32
        [rows+1] -> ascending matching row ids."""
        tweaks = np.ascontiguousarray(tweaks, dtype=np.uint8).reshape(-1, KEY_BYTES[key_format])
        rows = tweaks.shape[0]
        if out_offsets is not None:
            out_offsets = np.ascontiguousarray(out_offsets, dtype=np.uint32)
            assert out_offsets.shape[0] == rows + 1
            outputs = np.ascontiguousarray(outputs if outputs is not None else np.zeros(0, np.int64), dtype=np.int64)
        ids = np.zeros(max(rows, 1), np.uint32)
        n = C.c_uint32(0)
        outp = _ptr(outputs) if out_offsets is not None else None
        if host_threads is not None:
            self._check(lib().cudasp_scan_batch_mt(self._h, _ptr(tweaks), key_format, outp, _ptr(out_offsets), rows, _ptr(ids),
                                                   C.byref(n), int(host_threads)))
        elif key_format == KEYS_XY64:     # the reference-shaped entry point
            self._check(lib().cudasp_scan_batch(self._h, _ptr(tweaks), outp, _ptr(out_offsets), rows, _ptr(ids), C.byref(n)))
        else:
            self._check(lib().cudasp_scan_batch_fmt(self._h, _ptr(tweaks), key_format, outp, _ptr(out_offsets), rows, _ptr(ids),
                                                    C.byref(n)))
        return ids[: n.value].copy()

    def enable_details(self, max_details_per_batch: int = 0):
        """Opt in to the BIP-352 k-loop on matching rows (include/cudasp.h, match details); before the first batch."""
        self._check(lib().cudasp_query_enable_details(self._h, max_details_per_batch))
        self.details_enabled = True

    def scan_batch_details(self, tweaks: np.ndarray, outputs: np.ndarray, out_offsets: np.ndarray, key_format: int = KEYS_XY64,
                           capacity: Optional[int] = None):
        """-> (ascending matching row ids, [(row, k, output_index, label_index, t_k int)] sorted by (row, k))."""
        tweaks = np.ascontiguousarray(tweaks, dtype=np.uint8).reshape(-1, KEY_BYTES[key_format])
        rows = tweaks.shape[0]
        out_offsets = np.ascontiguousarray(out_offsets, dtype=np.uint32)
        assert out_offsets.shape[0] == rows + 1
        outputs = np.ascontiguousarray(outputs if outputs is not None else np.zeros(0, np.int64), dtype=np.int64)
        ids = np.zeros(max(rows, 1), np.uint32)
        cap = capacity if capacity is not None else min(max(rows, 1), 4096)
        n, nd = C.c_uint32(0), C.c_uint32(0)
        while True:
            det = (MatchDetail * max(cap, 1))()
            rc = lib().cudasp_scan_batch_details(self._h, _ptr(tweaks), key_format, _ptr(outputs), _ptr(out_offsets), rows, _ptr(ids),
                                                 C.byref(n), det, cap, C.byref(nd))
            if rc == ERR_ARG and capacity is None and nd.value > cap:      # more records than guessed: n_details says how many
                cap = nd.value
                continue
            self._check(rc)
            break
        a = np.frombuffer(det, dtype=DETAIL_DTYPE, count=nd.value)
        recs = list(zip(a["row"].tolist(), a["k"].tolist(), a["output_index"].tolist(), a["label_index"].tolist(),
                        [int.from_bytes(t.tobytes(), "big") for t in a["t_k"]]))
        return ids[: n.value].copy(), recs

    def candidates(self, tweaks: np.ndarray, key_format: int = KEYS_XY64, with_valid: bool = False):
        tweaks = np.ascontiguousarray(tweaks, dtype=np.uint8).reshape(-1, KEY_BYTES[key_format])
        rows = tweaks.shape[0]
        out = np.zeros((max(rows, 1), 1 + self.n_labels), np.int64)
        if key_format == KEYS_XY64 and not with_valid:
            self._check(lib().cudasp_scan_candidates(self._h, _ptr(tweaks), rows, _ptr(out)))
            return out[:rows]
        valid = np.zeros(max(rows, 1), np.uint8)
        self._check(lib().cudasp_scan_candidates_fmt(self._h, _ptr(tweaks), key_format, rows, _ptr(out), _ptr(valid)))
        return (out[:rows], valid[:rows].astype(bool)) if with_valid else out[:rows]

    def intermediates(self, tweaks: np.ndarray):
        tweaks = np.ascontiguousarray(tweaks, dtype=np.uint8).reshape(-1, 64)
        rows = tweaks.shape[0]
        S, t, P0 = np.zeros((max(rows, 1), 64), np.uint8), np.zeros((max(rows, 1), 32), np.uint8), np.zeros((max(rows, 1), 64), np.uint8)
        self._check(lib().cudasp_scan_intermediates(self._h, _ptr(tweaks), rows, _ptr(S), _ptr(t), _ptr(P0)))
        return S[:rows], t[:rows], P0[:rows]

    def scan_device(self, device: int, d_tweaks: int, d_outputs: int, d_offsets: int, rows: int, d_ids: int, d_count: int, stream: int = 0):
        """Raw device pointers (ints), e.g. torch tensor .data_ptr(); asynchronous on `stream`."""
        self._check(lib().cudasp_scan_device(self._h, device, d_tweaks, d_outputs, d_offsets, rows, d_ids, d_count, stream))


STAGES = ["ladder", "inv_z1", "hash_fixedbase", "inv_z2", "final_base", "inv_labels", "final_labels"]


def profile_stages(q: "Query", device: int, d_tweaks: int, d_outputs: int, d_offsets: int, rows: int, d_ids: int, d_count: int):
    ms = (C.c_float * 7)()
    q._check(lib().cudasp_profile_stages(q._h, device, d_tweaks, d_outputs, d_offsets, rows, d_ids, d_count, ms))
    return dict(zip(STAGES, [float(x) for x in ms]))


def debug_gtab(window: int, first_digit: int, n: int, device: int = 0):
    """n entries of the device-built fixed-base table as (x, y) integer pairs (include/cudasp.h cudasp_debug_gtab)."""
    out = np.zeros(64 * max(n, 1), np.uint8)
    rc = lib().cudasp_debug_gtab(device, window, first_digit, n, _ptr(out))
    if rc:
        raise CudaspError(rc, (lib().cudasp_last_error(None) or b"").decode())
    return [(int.from_bytes(out[64 * i:64 * i + 32].tobytes(), "little"), int.from_bytes(out[64 * i + 32:64 * i + 64].tobytes(), "little"))
            for i in range(n)]


def test_field(op: int, a: Sequence[int], b: Sequence[int], device: int = 0):
    n = len(a)
    A = np.frombuffer(b"".join(x.to_bytes(32, "little") for x in a), np.uint8).copy()
    B = np.frombuffer(b"".join(x.to_bytes(32, "little") for x in b), np.uint8).copy()
    out = np.zeros(32 * max(n, 1), np.uint8)
    rc = lib().cudasp_test_field(device, op, _ptr(A), _ptr(B), _ptr(out), n)
    if rc:
        raise CudaspError(rc, (lib().cudasp_last_error(None) or b"").decode())
    return [int.from_bytes(out[32 * i:32 * i + 32].tobytes(), "little") for i in range(n)]
