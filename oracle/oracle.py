"""ctypes binding of oracle/jpeg_oracle.c (the checker, never the product).

Mirrors what the reference's CPU plugin produces for one image
(extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:295-420): tight uint8 output,
I_RGB / I_BGR / P_RGB / P_BGR / Y, optional 2-D ROI == crop of the full decode.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_jpeg.so")

FMT_I_RGB, FMT_I_BGR, FMT_P_RGB, FMT_P_BGR, FMT_Y, FMT_I_YUV, FMT_P_YUV = 0, 1, 2, 3, 4, 5, 6
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "jpeg_oracle.c")
    if force or not os.path.exists(_SO) or (os.path.exists(src) and os.path.getmtime(_SO) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle_jpeg.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.orc_decode.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t]
        L.orc_decode.restype = ctypes.c_int
        L.orc_info.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
        L.orc_info.restype = ctypes.c_int
        L.orc_coefficients.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
        L.orc_coefficients.restype = ctypes.c_int
        _lib = L
    return _lib


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__({1: "not a JPEG", 2: "unsupported JPEG", 3: "corrupt JPEG", 4: "bad argument"}.get(code, str(code)))
        self.code = code


def _buf(data):
    a = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data, dtype=np.uint8)
    return a


def info(data):
    a = _buf(data)
    out = np.zeros(16, dtype=np.int32)
    rc = lib().orc_info(a.ctypes.data, a.size, out.ctypes.data)
    if rc:
        raise OracleError(rc)
    ncomp = int(out[2])
    return dict(width=int(out[0]), height=int(out[1]), ncomp=ncomp, hmax=int(out[3]), vmax=int(out[4]), mcus_x=int(out[5]),
                mcus_y=int(out[6]), restart_interval=int(out[7]), blocks_per_mcu=int(out[8]),
                sampling=[(int(out[9 + 2 * c]), int(out[10 + 2 * c])) for c in range(ncomp)], scan_offset=int(out[15]))


def decode(data, fmt=FMT_I_RGB, fancy=True, roi=None):
    """roi = (y0, x0, y1, x1) like nvimgcodecRegion_t start/end ([y,x] order)."""
    a = _buf(data)
    inf = info(a)
    if roi is None:
        y0, x0, y1, x1 = 0, 0, inf["height"], inf["width"]
    else:
        y0, x0, y1, x1 = roi
    h, w = y1 - y0, x1 - x0
    if fmt == FMT_Y:
        out = np.empty((h, w), dtype=np.uint8)
    elif fmt in (FMT_P_RGB, FMT_P_BGR, FMT_P_YUV):
        out = np.empty((3, h, w), dtype=np.uint8)
    else:
        out = np.empty((h, w, 3), dtype=np.uint8)
    rc = lib().orc_decode(a.ctypes.data, a.size, int(bool(fancy)), fmt, x0, y0, x1, y1, out.ctypes.data, 0)
    if rc:
        raise OracleError(rc)
    return out


def apply_orientation(img, code, planar=False):
    """EXIF orientation 1..8 applied to an HxW[xC] (or CxHxW when planar) array, as OpenCV's imread and
    Pillow's ImageOps.exif_transpose do (tests/test_oracle.py checks both)."""
    if code in (0, 1):
        return img
    a = np.moveaxis(img, 0, -1) if planar else img
    t = (lambda x: x.swapaxes(0, 1))
    out = {2: lambda x: x[:, ::-1], 3: lambda x: x[::-1, ::-1], 4: lambda x: x[::-1], 5: t, 6: lambda x: t(x)[:, ::-1],
           7: lambda x: t(x)[::-1, ::-1], 8: lambda x: t(x)[::-1]}[code](a)
    out = np.ascontiguousarray(out)
    return np.ascontiguousarray(np.moveaxis(out, -1, 0)) if planar else out


def coefficients(data):
    """int16 [n_mcu, blocks_per_mcu, 64], zig-zag order, MCU decode order, absolute DC."""
    a = _buf(data)
    inf = info(a)
    out = np.empty((inf["mcus_x"] * inf["mcus_y"], inf["blocks_per_mcu"], 64), dtype=np.int16)
    rc = lib().orc_coefficients(a.ctypes.data, a.size, out.ctypes.data)
    if rc:
        raise OracleError(rc)
    return out
