"""ctypes loader for libb200jpeg_ext.so (the C ABI in include/b200jpeg.h).

There is no CPU fallback: if the shared object is missing the import fails loudly,
and creating a decoder without a CUDA device raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200JPEG_LIB: another build of the same library (kernel experiments: variants are compared in one GPU session)
LIB_PATH = os.environ.get("B200JPEG_LIB") or os.path.join(_HERE, "lib", "libb200jpeg_ext.so")

OK, ERR_ARG, ERR_NOT_JPEG, ERR_UNSUPPORTED, ERR_CORRUPT, ERR_CUDA, ERR_STATE, ERR_TOO_LARGE = range(8)
STATUS_NAMES = ["OK", "ERR_ARG", "ERR_NOT_JPEG", "ERR_UNSUPPORTED", "ERR_CORRUPT", "ERR_CUDA", "ERR_STATE", "ERR_TOO_LARGE"]
FMT = {"I_RGB": 0, "I_BGR": 1, "P_RGB": 2, "P_BGR": 3, "P_Y": 4, "I_Y": 4, "Y": 4, "I_YUV": 5, "P_YUV": 6}

NOALLOC_MAX_SEGMENT_BYTES = 8 << 20

(STAT_KERNEL_LAUNCHES, STAT_ALGO_BYTES, STAT_COMPRESSED_BYTES, STAT_NUM_COEFS, STAT_OUTPUT_BYTES, STAT_RESYNC_IMAGES,
 STAT_DEVICE_BYTES, STAT_NUM_SUBSEQ) = range(8)


class StreamInfo(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int), ("height", ctypes.c_int), ("num_components", ctypes.c_int),
                ("h", ctypes.c_int * 4), ("v", ctypes.c_int * 4), ("sof_marker", ctypes.c_int), ("precision", ctypes.c_int),
                ("restart_interval", ctypes.c_int), ("num_scans_seen", ctypes.c_int), ("supported", ctypes.c_int),
                ("reason", ctypes.c_int), ("scan_offset", ctypes.c_uint64), ("scan_length", ctypes.c_uint64),
                ("mcus_x", ctypes.c_int), ("mcus_y", ctypes.c_int), ("blocks_per_mcu", ctypes.c_int)]


class ScanResult(ctypes.Structure):
    _fields_ = [("unstuffed_bytes", ctypes.c_uint32), ("num_restart_markers", ctypes.c_uint32),
                ("num_subsequences", ctypes.c_uint32), ("flags", ctypes.c_uint32)]


ALLOC_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_void_p)
FREE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p)


class Allocators(ctypes.Structure):
    _fields_ = [("device_ctx", ctypes.c_void_p), ("device_malloc", ALLOC_FN), ("device_free", FREE_FN),
                ("pinned_ctx", ctypes.c_void_p), ("pinned_malloc", ALLOC_FN), ("pinned_free", FREE_FN)]


class Output(ctypes.Structure):
    _fields_ = [("buffer", ctypes.c_void_p), ("format", ctypes.c_int), ("buffer_is_host", ctypes.c_int),
                ("row_stride", ctypes.c_size_t), ("roi_x0", ctypes.c_int), ("roi_y0", ctypes.c_int),
                ("roi_x1", ctypes.c_int), ("roi_y1", ctypes.c_int), ("orientation", ctypes.c_int), ("plane_stride", ctypes.c_size_t)]


EXPORTS = ["b200jpegProbe", "b200jpegCreate", "b200jpegCreateEx", "b200jpegDestroy", "b200jpegDecodeBatch", "b200jpegPrepareBatch",
           "b200jpegExecuteBatch", "b200jpegFinishBatch", "b200jpegGetStat", "b200jpegSetProfiling", "b200jpegGetStageTimes",
           "b200jpegDebugCopyCoefficients", "b200jpegGetLastError", "b200jpegVersion", "b200jpegScanSegment",
           "b200jpegScanSegmentNoAlloc", "b200jpegMultiCreate", "b200jpegMultiDestroy", "b200jpegMultiPlan", "b200jpegMultiDecodeBatch"]

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvimagecodec_b200 has no CPU fallback)")
    L = ctypes.CDLL(LIB_PATH)
    vp, sz, ci = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int
    L.b200jpegProbe.argtypes = [vp, sz, ctypes.POINTER(StreamInfo)]
    L.b200jpegProbe.restype = ci
    L.b200jpegCreate.argtypes = [ctypes.POINTER(vp), ci, ctypes.c_char_p]
    L.b200jpegCreate.restype = ci
    L.b200jpegCreateEx.argtypes = [ctypes.POINTER(vp), ci, ctypes.c_char_p, ctypes.POINTER(Allocators)]
    L.b200jpegCreateEx.restype = ci
    L.b200jpegDestroy.argtypes = [vp]
    L.b200jpegDestroy.restype = ci
    batch = [vp, ci, ctypes.POINTER(vp), ctypes.POINTER(sz), ctypes.POINTER(Output)]
    L.b200jpegDecodeBatch.argtypes = batch + [ctypes.POINTER(ci), vp]
    L.b200jpegDecodeBatch.restype = ci
    L.b200jpegPrepareBatch.argtypes = batch + [vp]
    L.b200jpegPrepareBatch.restype = ci
    L.b200jpegExecuteBatch.argtypes = [vp, vp]
    L.b200jpegExecuteBatch.restype = ci
    L.b200jpegFinishBatch.argtypes = [vp, ctypes.POINTER(ci), vp]
    L.b200jpegFinishBatch.restype = ci
    L.b200jpegGetStat.argtypes = [vp, ci]
    L.b200jpegGetStat.restype = ctypes.c_int64
    L.b200jpegSetProfiling.argtypes = [vp, ci]
    L.b200jpegSetProfiling.restype = ci
    L.b200jpegGetStageTimes.argtypes = [vp, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_char_p), ci]
    L.b200jpegGetStageTimes.restype = ci
    L.b200jpegDebugCopyCoefficients.argtypes = [vp, ci, vp, sz]
    L.b200jpegDebugCopyCoefficients.restype = ci
    L.b200jpegScanSegment.argtypes = [vp, vp, sz, ctypes.c_uint32, vp, vp, ctypes.c_uint32, vp, vp]
    L.b200jpegScanSegment.restype = ci
    L.b200jpegScanSegmentNoAlloc.argtypes = [vp, sz, ctypes.c_uint32, vp, vp, ctypes.c_uint32, vp, vp, ctypes.c_char_p, sz]
    L.b200jpegScanSegmentNoAlloc.restype = ci
    L.b200jpegMultiCreate.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(ci), ci, ctypes.c_char_p]
    L.b200jpegMultiCreate.restype = ci
    L.b200jpegMultiDestroy.argtypes = [vp]
    L.b200jpegMultiDestroy.restype = ci
    L.b200jpegMultiPlan.argtypes = [vp, ci, ctypes.POINTER(sz), ctypes.POINTER(ci)]
    L.b200jpegMultiPlan.restype = ci
    L.b200jpegMultiDecodeBatch.argtypes = [vp, ci, ctypes.POINTER(vp), ctypes.POINTER(sz), ctypes.POINTER(Output), ctypes.POINTER(ci),
                                           ctypes.POINTER(ci)]
    L.b200jpegMultiDecodeBatch.restype = ci
    L.b200jpegGetLastError.argtypes = [vp]
    L.b200jpegGetLastError.restype = ctypes.c_char_p
    L.b200jpegVersion.argtypes = []
    L.b200jpegVersion.restype = ctypes.c_char_p
    _lib = L
    return L
