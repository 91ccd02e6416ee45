"""ctypes binding of baseline/libcpu_scan_openssl.so -- BENCH / TEST INFRASTRUCTURE, NOT PRODUCT CODE (see the .cpp header)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libcpu_scan_openssl.so")
_lib = None


def available() -> bool:
    try:
        load()
        return True
    except Exception:      # noqa: BLE001  (no OpenSSL headers / library: the baseline is optional)
        return False


def load():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "cpu_scan_openssl.cpp")
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
            subprocess.check_call(["make", "-s", "-C", _HERE])
        lib = C.CDLL(_SO)
        vp = C.c_void_p
        lib.cpu_scan_openssl.argtypes = [vp, vp, vp, C.c_uint32, vp, vp, vp, C.c_uint32, vp, vp, C.c_int]
        lib.cpu_scan_openssl.restype = C.c_int
        _lib = lib
    return _lib


def scan(sk, spend, labels, tweaks, outputs, out_off, threads=1, want_cands=False):
    """Same contract as oracle.c_oracle.scan: -> (flags uint8[rows], cands int64[rows, 1+L] or None)."""
    lib = load()
    tweaks = np.ascontiguousarray(tweaks, dtype=np.uint8).reshape(-1, 64)
    rows, L = tweaks.shape[0], len(labels)
    lab = np.frombuffer(b"".join(labels) or b"\0", np.uint8).copy()
    flags = np.zeros(max(rows, 1), np.uint8)
    cands = np.zeros((max(rows, 1), 1 + L), np.int64) if want_cands else None
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)      # noqa: E731
    if out_off is not None:
        outputs = np.ascontiguousarray(outputs if outputs is not None and len(outputs) else np.zeros(1, np.int64), dtype=np.int64)
        out_off = np.ascontiguousarray(out_off, dtype=np.uint32)
    lib.cpu_scan_openssl(p(np.frombuffer(sk, np.uint8).copy()), p(np.frombuffer(spend, np.uint8).copy()), p(lab), L, p(tweaks),
                         p(outputs) if out_off is not None else None, p(out_off), rows, p(flags), p(cands), int(threads))
    return flags[:rows], (cands[:rows] if want_cands else None)
