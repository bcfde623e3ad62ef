"""Host-only stream probing through the C ABI (b200jpegProbe); no GPU needed."""
import ctypes

import numpy as np

from . import _lib


def probe(data):
    """Marker walk of one JPEG: returns the filled StreamInfo (see include/b200jpeg.h)."""
    L = _lib.load()
    buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    info = _lib.StreamInfo()
    L.b200jpegProbe(ctypes.c_void_p(buf.ctypes.data), buf.size, ctypes.byref(info))
    return info
