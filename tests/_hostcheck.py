"""Loader for tests/hostcheck/libhostcheck.so (TEST INFRASTRUCTURE): the product's host/device row
functions compiled with g++ so their logic can be checked on the CPU against the oracle."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "hostcheck", "host_check.cpp")
# HOSTCHECK_DEFS="-DCSP_DBL_2M5S -DCSP_GTAB_WBITS=16": check a tuning build of the row functions (see csrc/Makefile VARIANT_DEFS)
EXTRA_DEFS = os.environ.get("HOSTCHECK_DEFS", "").split()
# the CPU logic tests build the whole fixed-base table on the host: they use 13-bit windows (160 K entries) whatever
# width the product is compiled with (the window code is parametric; the device-built table has its own GPU test)
WBITS_DEF = [] if any("CSP_GTAB_WBITS" in d for d in EXTRA_DEFS) else ["-DCSP_GTAB_WBITS=13"]
_SUFFIX = ("_" + "".join(c if c.isalnum() else "_" for c in "".join(EXTRA_DEFS))) if EXTRA_DEFS else ""
SO = os.path.join(ROOT, "tests", "hostcheck", "libhostcheck%s.so" % _SUFFIX)
SO_EMUL = os.path.join(ROOT, "tests", "hostcheck", "libhostcheck_emul%s.so" % _SUFFIX)   # -DCSP_FE_EMUL: the device carry-chain code, emulated
CSRC = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "csrc")
_lib = None


def build(force=False):
    deps = [SRC] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".hpp"))]
    newest = max(os.path.getmtime(d) for d in deps)
    for so, extra in ((SO, []), (SO_EMUL, ["-DCSP_FE_EMUL"])):
        if force or not os.path.exists(so) or os.path.getmtime(so) < newest:
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", "-I", CSRC,
                                   "-Wno-unknown-pragmas", "-pthread"] + extra + WBITS_DEF + EXTRA_DEFS + ["-o", so, SRC])
    return SO


_libs = {}
EMUL = False   # module-wide switch used by the parametrised tests: True -> the emulated carry-chain build


def load():
    if EMUL not in _libs:
        build()
        _libs[EMUL] = C.CDLL(SO_EMUL if EMUL else SO)
    return _libs[EMUL]


def _p(a, t=C.c_uint8):
    return a.ctypes.data_as(C.POINTER(t))


def _b(b):
    return np.frombuffer(bytes(b), np.uint8).copy()


def field(op, a_list, b_list):
    lib = load()
    n = len(a_list)
    a = _b(b"".join(x.to_bytes(32, "little") for x in a_list))
    b = _b(b"".join(x.to_bytes(32, "little") for x in b_list))
    out = np.zeros(32 * n, np.uint8)
    lib.hc_field(op, _p(a), _p(b), _p(out), n)
    return [int.from_bytes(out[32 * i:32 * i + 32].tobytes(), "little") for i in range(n)]


def glv(k):
    lib = load()
    k1, k2 = np.zeros(32, np.uint8), np.zeros(32, np.uint8)
    n1, n2 = C.c_int(), C.c_int()
    lib.hc_glv(_p(_b(k.to_bytes(32, "little"))), _p(k1), C.byref(n1), _p(k2), C.byref(n2))
    a, b = int.from_bytes(k1.tobytes(), "little"), int.from_bytes(k2.tobytes(), "little")
    return (-a if n1.value else a), (-b if n2.value else b)


def plan(sk_le):
    lib = load()
    ops = np.zeros(256, np.uint8)
    v = [C.c_int() for _ in range(5)]
    lib.hc_plan(_p(_b(sk_le)), _p(ops), *[C.byref(x) for x in v])
    nops, n_dbl, n_add, zero, regular = [x.value for x in v]
    return dict(ops=ops[:nops].tolist(), n_dbl=n_dbl, n_add=n_add, zero=bool(zero), regular=bool(regular))


def plan_variant(sk_le, i, j):
    lib = load()
    ops = np.zeros(256, np.uint8)
    v = [C.c_int() for _ in range(4)]
    lib.hc_plan_variant(_p(_b(sk_le)), i, j, _p(ops), *[C.byref(x) for x in v])
    nops, n_dbl, n_add, regular = [x.value for x in v]
    if nops < 0:
        return None
    return dict(ops=ops[:nops].tolist(), n_dbl=n_dbl, n_add=n_add, regular=bool(regular))


def plan_chosen(sk_le):
    lib = load()
    ops = np.zeros(256, np.uint8)
    v = [C.c_int() for _ in range(3)]
    rc = lib.hc_plan_chosen(_p(_b(sk_le)), _p(ops), *[C.byref(x) for x in v])
    assert rc == 0
    nops, n_dbl, n_add = [x.value for x in v]
    return dict(ops=ops[:nops].tolist(), n_dbl=n_dbl, n_add=n_add)


def midstate():
    lib = load()
    out = np.zeros(8, np.uint32)
    lib.hc_midstate(_p(out, C.c_uint32))
    return out.tolist()


def ladder(sk_le, points):
    lib = load()
    pts = _b(b"".join(points))
    out = np.zeros(64 * len(points), np.uint8)
    rc = lib.hc_ladder(_p(_b(sk_le)), _p(pts), len(points), _p(out))
    assert rc == 0, rc
    return [out[64 * i:64 * i + 64].tobytes() for i in range(len(points))]


def scan_candidates(sk, spend, labels, tweaks, T=7, want_inter=False):
    lib = load()
    L = len(labels)
    tw = _b(b"".join(tweaks))
    rows = len(tweaks)
    lab = _b(b"".join(labels)) if L else np.zeros(1, np.uint8)
    cands = np.zeros((rows, 1 + L), np.int64)
    S, t, P0 = np.zeros(64 * rows, np.uint8), np.zeros(32 * rows, np.uint8), np.zeros(64 * rows, np.uint8)
    rc = lib.hc_scan_candidates(_p(_b(sk)), _p(_b(spend)), _p(lab), L, _p(tw), rows, T, _p(cands, C.c_int64),
                                _p(S) if want_inter else None, _p(t) if want_inter else None, _p(P0) if want_inter else None)
    if rc:
        raise RuntimeError("hc_scan_candidates rc=%d" % rc)
    if want_inter:
        return cands, S.reshape(rows, 64), t.reshape(rows, 32), P0.reshape(rows, 64)
    return cands


def gtab_entry(w, d):
    lib = load()
    out = np.zeros(64, np.uint8)
    lib.hc_gtab_entry(w, d, _p(out))
    return out.tobytes()


GTAB_EPT = 16


def gtab_chunk(w, d0):
    """rowfn.cuh gtab_chunk (the device table builder's row function) run on the CPU -> list of 64-byte entries"""
    lib = load()
    out = np.zeros(64 * GTAB_EPT, np.uint8)
    n = lib.hc_gtab_chunk(w, C.c_uint32(d0), _p(out))
    assert n == GTAB_EPT, "gtab_chunk reported an exceptional addition"
    return [out[64 * i:64 * i + 64].tobytes() for i in range(n)]


def decompress(sec1: bytes):
    lib = load()
    out = np.zeros(64, np.uint8)
    ok = lib.hc_decompress(_p(_b(sec1)), _p(out))
    return bool(ok), out.tobytes()


def details_row(sk, spend, labels, S_le, outputs, cap=64):
    """-> [(k, output_index, label_index, t_k int)] from the product's details_row (K8 body) on the CPU."""
    lib = load()
    lab = _b(b"".join(labels)) if labels else np.zeros(1, np.uint8)
    outs = np.ascontiguousarray(outputs, dtype=np.int64) if len(outputs) else np.zeros(1, np.int64)
    out = np.zeros(44 * cap, np.uint8)
    n = lib.hc_details_row(_p(_b(sk)), _p(_b(spend)), _p(lab), len(labels), _p(_b(S_le)), outs.ctypes.data_as(C.POINTER(C.c_int64)),
                           len(outputs), _p(out), cap)
    assert n >= 0, n
    res = []
    for i in range(n):
        rec = out[44 * i: 44 * i + 44]
        k, oi, li = np.frombuffer(rec[:12].tobytes(), np.int32).tolist()
        res.append((k, oi, li, int.from_bytes(rec[12:].tobytes(), "big")))
    return res


def tweak_in_range(t: int) -> bool:
    return bool(load().hc_tweak_in_range(_p(_b(t.to_bytes(32, "little")))))


def gtab_shape():
    lib = load()
    a, b = C.c_int(), C.c_int()
    lib.hc_gtab_shape(C.byref(a), C.byref(b))
    return a.value, b.value


def gtab_digit(t, w):
    lib = load()
    lib.hc_gtab_digit.restype = C.c_uint32
    return int(lib.hc_gtab_digit(_p(_b(t.to_bytes(32, "little"))), w))


def gtab_offset(w):
    lib = load()
    out = np.zeros(32, np.uint8)
    lib.hc_gtab_offset(w, _p(out))
    return int.from_bytes(out.tobytes(), "little")


def piece_plan(rows, piece, n_devices):
    """starts of the pieces cudasp_scan_batch cuts `rows` rows into (csrc/piece_plan.hpp)."""
    lib = load()
    cap = rows // max(min(piece, 8192), 1) + 64 * n_devices + 8
    out = np.zeros(cap, np.uint32)
    n = lib.hc_piece_plan(rows, piece, n_devices, out.ctypes.data_as(C.POINTER(C.c_uint32)), cap)
    assert 0 <= n < cap
    return out[: n + 1].tolist()
