"""ctypes harness around the UNMODIFIED reference GPU pipeline -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

oracle/_ref/libcudasp_gecc_ref.so is the reference's own src/cudasp_gpu.cu (+ its gECC headers and the
headers its build-time generator emits) compiled for sm_100 by `make -C oracle ref` from the sources
where they lie under /root/reference (see oracle/Makefile).  It exports the reference's internal FFI
    LaunchBatchScan / RunBatchScanKernels / FreeBatchScanState     (src/cudasp_extension.cu:21-52)
and this module drives it exactly the way the reference's ProcessBatch does
(src/cudasp_extension.cu:226-410): column-major points written into the managed buffers the callee
allocated, the three calls, the flag scan, then cudaFree x3 + FreeBatchScanState.

Only tests/ (as the checker) and bench.py's reference/baseline legs may import this.  It needs a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libcudasp_gecc_ref.so")
_lib = None
_rt = None


def available() -> bool:
    return os.path.exists(SO)


def load():
    global _lib, _rt
    if _lib is None:
        lib = C.CDLL(SO, mode=C.RTLD_GLOBAL)
        u32p, u32pp = C.POINTER(C.c_uint32), C.POINTER(C.POINTER(C.c_uint32))
        u8pp = C.POINTER(C.POINTER(C.c_uint8))
        lib.LaunchBatchScan.argtypes = [u32pp, u32pp, u32p, u32p, u32p, u32p, u32p, C.c_uint32, C.POINTER(C.c_int64), u32p, u32p,
                                        u8pp, C.c_uint32, C.c_size_t]
        lib.LaunchBatchScan.restype = C.c_void_p
        lib.RunBatchScanKernels.argtypes = [C.c_void_p, u32p, u32p, u32p, u32p, u32p, u32p, u32p, C.c_uint32, C.POINTER(C.c_uint8),
                                            C.c_uint32]
        lib.RunBatchScanKernels.restype = C.c_int
        lib.FreeBatchScanState.argtypes = [C.c_void_p]
        lib.FreeBatchScanState.restype = None
        # the library is linked with --cudart shared, so this is the same runtime instance
        rt = None
        for name in ("libcudart.so.12", "libcudart.so"):
            try:
                rt = C.CDLL(name)
                break
            except OSError:
                continue
        if rt is None:
            rt = C.CDLL("/usr/local/cuda/lib64/libcudart.so.12")
        rt.cudaFree.argtypes = [C.c_void_p]
        rt.cudaFree.restype = C.c_int
        _lib, _rt = lib, rt
    return _lib


def _limbs(b: bytes) -> np.ndarray:
    """32 or 64 B little-endian -> u32 limbs (ConvertScalarToU32 / ConvertTweakKeyToU32, :59-85)."""
    return np.frombuffer(bytes(b), dtype="<u4").copy()


def _u32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def scan_batch(sk: bytes, spend: bytes, labels, tweaks: np.ndarray, outputs: np.ndarray, offsets: np.ndarray):
    """One reference batch.  tweaks (n,64) u8; outputs int64; offsets (n+1) CSR.  -> (flags u8[n], seconds)."""
    lib = load()
    n = int(tweaks.shape[0])
    L = len(labels)
    scalar = _limbs(sk)
    sp = _limbs(spend)
    spx, spy = sp[:8].copy(), sp[8:].copy()
    lx = np.zeros(max(L, 1) * 8, dtype=np.uint32)
    ly = np.zeros(max(L, 1) * 8, dtype=np.uint32)
    for i, lab in enumerate(labels):
        v = _limbs(lab)
        lx[i * 8:(i + 1) * 8] = v[:8]
        ly[i * 8:(i + 1) * 8] = v[8:]
    off = np.ascontiguousarray(offsets[:n].astype(np.uint32) - np.uint32(offsets[0]))
    ln = np.ascontiguousarray(np.diff(offsets[: n + 1]).astype(np.uint32))
    o0, o1 = int(offsets[0]), int(offsets[n])
    outs = np.ascontiguousarray(outputs[o0:o1].astype(np.int64)) if o1 > o0 else np.zeros(1, dtype=np.int64)
    n_out = o1 - o0
    w = np.ascontiguousarray(tweaks).view("<u4").reshape(n, 16)
    t0 = time.perf_counter()
    mx, my = C.POINTER(C.c_uint32)(), C.POINTER(C.c_uint32)()
    mf = C.POINTER(C.c_uint8)()
    h = lib.LaunchBatchScan(C.byref(mx), C.byref(my), _u32p(scalar), _u32p(spx), _u32p(spy), _u32p(lx), _u32p(ly), L,
                            outs.ctypes.data_as(C.POINTER(C.c_int64)), _u32p(off), _u32p(ln), C.byref(mf), n, max(n_out, 1) if n_out else 0)
    if not h:
        raise RuntimeError("reference LaunchBatchScan returned NULL")
    try:
        # column-major fill of the managed buffers (src/cudasp_extension.cu:327-337)
        ax = np.ctypeslib.as_array(mx, shape=(8, n))
        ay = np.ctypeslib.as_array(my, shape=(8, n))
        ax[:, :] = w[:, :8].T
        ay[:, :] = w[:, 8:].T
        rc = lib.RunBatchScanKernels(h, mx, my, _u32p(scalar), _u32p(spx), _u32p(spy), _u32p(lx), _u32p(ly), L, mf, n)
        if rc != 0:
            raise RuntimeError("reference RunBatchScanKernels returned %d" % rc)
        flags = np.ctypeslib.as_array(mf, shape=(n,)).copy()
    finally:
        _rt.cudaFree(C.cast(mx, C.c_void_p))
        _rt.cudaFree(C.cast(my, C.c_void_p))
        _rt.cudaFree(C.cast(mf, C.c_void_p))
        lib.FreeBatchScanState(h)
    return flags, time.perf_counter() - t0


def scan(sk, spend, labels, tweaks, outputs, offsets, batch_rows=300000):
    """Whole table in reference-sized batches (default batch_size 300000, :473-489). -> (flags, seconds)."""
    n = int(tweaks.shape[0])
    flags = np.zeros(n, dtype=np.uint8)
    total = 0.0
    for base in range(0, n, batch_rows):
        cnt = min(batch_rows, n - base)
        f, dt = scan_batch(sk, spend, labels, tweaks[base: base + cnt], outputs, offsets[base: base + cnt + 1])
        flags[base: base + cnt] = f
        total += dt
    return flags, total
