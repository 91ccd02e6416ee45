"""ctypes binding of oracle/libcudasp_oracle.so -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this.  `load()` builds the shared object with gcc on first use.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libcudasp_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "cudasp_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B" if force else "-s"])
    return _SO


def load():
    global _lib
    if _lib is None:
        build()
        lib = C.CDLL(_SO)
        u8p, i64p, u32p = C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.POINTER(C.c_uint32)
        lib.oracle_row.argtypes = [u8p, u8p, u8p, C.c_uint32, u8p, i64p, u8p, u8p, u8p]
        lib.oracle_row.restype = C.c_int
        lib.oracle_scan.argtypes = [u8p, u8p, u8p, C.c_uint32, u8p, i64p, u32p, C.c_uint32, u8p, i64p, C.c_int]
        lib.oracle_scan.restype = C.c_int
        lib.oracle_pt_mul.argtypes = [u8p, u8p, u8p]
        lib.oracle_pt_mul.restype = C.c_int
        lib.oracle_pt_add.argtypes = [u8p, u8p, u8p]
        lib.oracle_pt_add.restype = C.c_int
        lib.oracle_fe_op.argtypes = [C.c_int, u8p, u8p, u8p]
        lib.oracle_fe_op.restype = C.c_int
        lib.oracle_gen_chain.argtypes = [u8p, u8p, C.c_uint32, u8p]
        lib.oracle_gen_chain.restype = C.c_int
        lib.oracle_tagged_hash.argtypes = [u8p, C.c_uint32, u8p]
        lib.oracle_tagged_hash.restype = None
        _lib = lib
    return _lib


def _p8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _buf(b: bytes) -> np.ndarray:
    return np.frombuffer(bytes(b), dtype=np.uint8).copy()


def row(sk: bytes, spend: bytes, labels: Sequence[bytes], tweak: bytes):
    """-> dict(cands=[int64...], S=(x,y), t=int, P0=(x,y))"""
    lib = load()
    L = len(labels)
    lab = _buf(b"".join(labels)) if L else np.zeros(1, np.uint8)
    cands = np.zeros(1 + L, np.int64)
    S, t, P0 = np.zeros(64, np.uint8), np.zeros(32, np.uint8), np.zeros(64, np.uint8)
    lib.oracle_row(_p8(_buf(sk)), _p8(_buf(spend)), _p8(lab), L, _p8(_buf(tweak)),
                   cands.ctypes.data_as(C.POINTER(C.c_int64)), _p8(S), _p8(t), _p8(P0))
    be = lambda a: int.from_bytes(a.tobytes(), "big")
    return {"cands": [int(c) for c in cands], "S": (be(S[:32]), be(S[32:])), "t": be(t),
            "P0": (be(P0[:32]), be(P0[32:]))}


def scan(sk: bytes, spend: bytes, labels: Sequence[bytes], tweaks: np.ndarray,
         outputs: Optional[np.ndarray], out_off: Optional[np.ndarray], threads: int = 1,
         want_cands: bool = False):
    """tweaks: uint8 [rows,64]; outputs int64 [n]; out_off uint32 [rows+1].
    -> (flags uint8[rows], cands int64[rows,1+L] or None)"""
    lib = load()
    tweaks = np.ascontiguousarray(tweaks, dtype=np.uint8).reshape(-1, 64)
    rows = tweaks.shape[0]
    L = len(labels)
    lab = _buf(b"".join(labels)) if L else np.zeros(1, np.uint8)
    flags = np.zeros(max(rows, 1), np.uint8)
    cands = np.zeros((max(rows, 1), 1 + L), np.int64) if want_cands else None
    if outputs is None:
        outputs = np.zeros(1, np.int64)
        out_off = np.zeros(rows + 1, np.uint32)
    outputs = np.ascontiguousarray(outputs, dtype=np.int64)
    if outputs.size == 0:
        outputs = np.zeros(1, np.int64)
    out_off = np.ascontiguousarray(out_off, dtype=np.uint32)
    assert out_off.shape[0] == rows + 1
    lib.oracle_scan(_p8(_buf(sk)), _p8(_buf(spend)), _p8(lab), L, _p8(tweaks),
                    outputs.ctypes.data_as(C.POINTER(C.c_int64)),
                    out_off.ctypes.data_as(C.POINTER(C.c_uint32)), rows, _p8(flags),
                    cands.ctypes.data_as(C.POINTER(C.c_int64)) if want_cands else None, int(threads))
    return flags[:rows], (cands[:rows] if want_cands else None)


def pt_mul(k_le: bytes, p: bytes):
    lib = load()
    out = np.zeros(64, np.uint8)
    inf = lib.oracle_pt_mul(_p8(_buf(k_le)), _p8(_buf(p)), _p8(out))
    return out.tobytes(), bool(inf)


def pt_add(a: bytes, b: bytes):
    lib = load()
    out = np.zeros(64, np.uint8)
    inf = lib.oracle_pt_add(_p8(_buf(a)), _p8(_buf(b)), _p8(out))
    return out.tobytes(), bool(inf)


def fe_op(op: int, a: int, b: int = 0) -> int:
    lib = load()
    out = np.zeros(32, np.uint8)
    lib.oracle_fe_op(op, _p8(_buf(a.to_bytes(32, "little"))), _p8(_buf(b.to_bytes(32, "little"))), _p8(out))
    return int.from_bytes(out.tobytes(), "little")


def tagged_hash(msg: bytes) -> bytes:
    lib = load()
    out = np.zeros(32, np.uint8)
    lib.oracle_tagged_hash(_p8(_buf(msg)), len(msg), _p8(out))
    return out.tobytes()


def gen_chain(p0: bytes, q: bytes, n: int) -> np.ndarray:
    """uint8 [n,64]: p0 + i*q for i < n (valid curve points)."""
    lib = load()
    out = np.zeros((max(n, 1), 64), np.uint8)
    rc = lib.oracle_gen_chain(_p8(_buf(p0)), _p8(_buf(q)), n, _p8(out))
    assert rc == 0
    return out[:n]
