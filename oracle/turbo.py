"""Binding of oracle/_ref/libref_turbo.so: the reference's own CPU JPEG path
(libjpeg_turbo::Uncompress from extensions/libjpeg_turbo/jpeg_mem.cpp, compiled from the reference
tree by oracle/build_ref.sh) over the libjpeg-turbo binary that ships with Pillow.

TEST INFRASTRUCTURE ONLY: pins the oracle (incl. fancy_upsampling=0 and cropped decodes) and is the
CPU baseline / reference arm bench.py times.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libref_turbo.so")
_lib = None


def available():
    return os.path.exists(_SO)


def engine():
    p = os.path.join(_HERE, "_ref", "libref_turbo.engine")
    return open(p).read().strip() if os.path.exists(p) else "unknown"


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(_SO)
        L.ref_turbo_decode.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int)]
        L.ref_turbo_decode.restype = ctypes.c_int
        L.ref_turbo_time_batch.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int), ctypes.c_int,
                                           ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.ref_turbo_time_batch.restype = ctypes.c_double
        _lib = L
    return _lib


def decode(data, fmt="I_RGB", fancy=True, roi=None):
    """roi = (y0, x0, y1, x1).  Returns HxWx3 (I_RGB / I_BGR) or HxW (I_Y)."""
    a = np.frombuffer(bytes(data), dtype=np.uint8)
    w, h = ctypes.c_int(), ctypes.c_int()
    code = {"I_RGB": 0, "I_BGR": 1, "I_Y": 4, "I_YUV": 5}[fmt]
    crop = (0, 0, 0, 0) if roi is None else (roi[1], roi[0], roi[3] - roi[1], roi[2] - roi[0])
    rc = lib().ref_turbo_decode(a.ctypes.data, a.size, int(bool(fancy)), code, *crop, None, ctypes.byref(w), ctypes.byref(h))
    if rc:
        raise RuntimeError(f"reference libjpeg_turbo path failed ({rc})")
    ch = 1 if fmt == "I_Y" else 3
    out = np.empty((h.value, w.value, ch), dtype=np.uint8)
    rc = lib().ref_turbo_decode(a.ctypes.data, a.size, int(bool(fancy)), code, *crop, out.ctypes.data, ctypes.byref(w), ctypes.byref(h))
    if rc:
        raise RuntimeError(f"reference libjpeg_turbo path failed ({rc})")
    return out[..., 0] if ch == 1 else out


def time_batch(jpegs, threads, passes, fancy=True):
    """Seconds to decode every stream `passes` times on `threads` host threads, plus a description."""
    arrs = [np.frombuffer(bytes(j), dtype=np.uint8) for j in jpegs]
    n = len(arrs)
    ptrs = (ctypes.c_void_p * n)(*[a.ctypes.data for a in arrs])
    lens = (ctypes.c_int * n)(*[a.size for a in arrs])
    dt = lib().ref_turbo_time_batch(n, ptrs, lens, int(bool(fancy)), 0, int(threads), int(passes))
    if dt < 0:
        raise RuntimeError("reference libjpeg_turbo path failed on a stream")
    return dt, ("reference libjpeg_turbo::Uncompress (extensions/libjpeg_turbo/jpeg_mem.cpp) over "
                + os.path.basename(engine()) + ", I_RGB, ISLOW, fancy_upsampling=%d" % int(bool(fancy)))
