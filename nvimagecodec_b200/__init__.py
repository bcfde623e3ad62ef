"""nvimagecodec_b200 -- B200-native batched baseline-JPEG decode path for nvImageCodec.

Only what the hot path needs: csrc/ (CUDA kernels, C ABI, nvImageCodec extension adapter),
a ctypes loader and the Python mirror of the reference's decoder object.
"""
from . import _lib  # noqa: F401

__all__ = ["Decoder", "MultiGpuDecoder", "DecodeError"]


def __getattr__(name):
    if name in ("Decoder", "MultiGpuDecoder", "DecodeError"):
        from . import decoder

        return getattr(decoder, name)
    raise AttributeError(name)
