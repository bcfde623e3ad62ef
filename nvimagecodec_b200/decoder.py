"""Host-side mirror of the reference's Python decoder object for the JPEG path.

``nvidia.nvimgcodec.Decoder(device_id=..., options=":fancy_upsampling=0")`` /
``Decoder.decode(srcs)`` (python/decoder.cpp:39-72, 94-341, default options :492)
decode a list of encoded streams into uint8 device images (I_RGB / I_Y, tight pitch).
This class does the same for baseline JPEG through the C ABI in include/b200jpeg.h;
torch is used only to own device memory and the CUDA stream.
"""
import ctypes

import numpy as np
import torch

from . import _lib


class DecodeError(RuntimeError):
    pass


def _as_u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8)
    return np.frombuffer(data, dtype=np.uint8)


class Batch:
    """One prepared batch: keeps the ctypes argument arrays and the output storage alive."""

    def __init__(self, n):
        self.n = n
        self.keep = []
        self.ptrs = (ctypes.c_void_p * max(n, 1))()
        self.lens = (ctypes.c_size_t * max(n, 1))()
        self.outs = (_lib.Output * max(n, 1))()
        self.status = (ctypes.c_int * max(n, 1))()
        self.images = [None] * n
        self.h2d_bytes = 0
        self.d2h_bytes = 0


class Decoder:
    # Same default as the reference's Python binding (python/decoder.cpp:492).
    def __init__(self, device_id=None, options=":fancy_upsampling=0", allocators=None):
        """allocators: an optional ``_lib.Allocators`` (the caller's device / pinned allocation callbacks, the C-ABI
        form of nvimgcodecExecutionParams_t::device_allocator / pinned_allocator); kept alive with the decoder."""
        self._L = _lib.load()
        if not torch.cuda.is_available():
            raise DecodeError("nvimagecodec_b200 needs a CUDA device (there is no CPU fallback)")
        self.device_id = torch.cuda.current_device() if device_id is None else int(device_id)
        self.device = torch.device("cuda", self.device_id)
        h = ctypes.c_void_p()
        self._allocators = allocators
        if allocators is None:
            rc = self._L.b200jpegCreate(ctypes.byref(h), self.device_id, (options or "").encode())
        else:
            rc = self._L.b200jpegCreateEx(ctypes.byref(h), self.device_id, (options or "").encode(), ctypes.byref(allocators))
        if rc != _lib.OK:
            raise DecodeError(f"b200jpegCreate failed: {_lib.STATUS_NAMES[rc]}")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.b200jpegDestroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ batch construction
    def make_batch(self, srcs, fmt="I_RGB", rois=None, to_host=False, out_storage=None, row_align=1, orientations=None, size_only=False):
        """srcs: list of bytes / uint8 arrays.  fmt: one name or a list of names per image.
        rois: None or list of None | (y0, x0, y1, x1) (nvimgcodecRegion_t order).
        to_host: write into pinned host tensors (decoder bounces through HBM + D2H)."""
        n = len(srcs)
        b = Batch(n)
        fmts = [fmt] * n if isinstance(fmt, str) else list(fmt)
        rois = [None] * n if rois is None else list(rois)
        orientations = [0] * n if orientations is None else list(orientations)  # EXIF codes; 5..8 swap width and height
        infos = []
        total = 0
        shapes = []
        for i, s in enumerate(srcs):
            a = _as_u8(s)
            b.keep.append(a)
            b.ptrs[i] = a.ctypes.data
            b.lens[i] = a.size
            b.h2d_bytes += a.size
            info = _lib.StreamInfo()
            self._L.b200jpegProbe(ctypes.c_void_p(a.ctypes.data), a.size, ctypes.byref(info))
            infos.append(info)
            if rois[i] is None:
                hgt, wid = info.height, info.width
            else:
                y0, x0, y1, x1 = rois[i]
                hgt, wid = y1 - y0, x1 - x0
            if orientations[i] >= 5:
                hgt, wid = wid, hgt
            f = fmts[i]
            if f in ("I_RGB", "I_BGR", "I_YUV"):
                shape, stride = (hgt, wid, 3), wid * 3
            elif f in ("P_RGB", "P_BGR", "P_YUV"):
                shape, stride = (3, hgt, wid), wid
            else:
                shape, stride = (hgt, wid), wid
            if row_align > 1:
                stride = (stride + row_align - 1) // row_align * row_align
            nbytes = stride * hgt * (3 if f.startswith("P_") and f != "P_Y" else 1)
            shapes.append((shape, stride, total, nbytes, f))
            total += (max(nbytes, 1) + 255) // 256 * 256
        if size_only:  # how many bytes of output storage the batch needs
            return total
        if out_storage is None:
            if to_host:
                out_storage = torch.empty(max(total, 1), dtype=torch.uint8, pin_memory=True)
            else:
                out_storage = torch.empty(max(total, 1), dtype=torch.uint8, device=self.device)
        assert out_storage.numel() >= total
        b.storage = out_storage
        base = out_storage.data_ptr()
        for i, (shape, stride, off, nbytes, f) in enumerate(shapes):
            o = b.outs[i]
            o.buffer = base + off
            o.format = _lib.FMT[f]
            o.buffer_is_host = 1 if to_host else 0
            o.row_stride = stride
            o.orientation = int(orientations[i])
            if rois[i] is not None:
                y0, x0, y1, x1 = rois[i]
                o.roi_x0, o.roi_y0, o.roi_x1, o.roi_y1 = x0, y0, x1, y1
            if infos[i].supported and nbytes > 0:
                flat = out_storage[off:off + nbytes]
                if f in ("I_RGB", "I_BGR", "I_YUV"):
                    img = flat.view(shape[0], stride)[:, :shape[1] * 3].view(shape[0], shape[1], 3) if stride == shape[1] * 3 else \
                        flat.view(shape[0], stride)[:, :shape[1] * 3].unflatten(1, (shape[1], 3))
                elif f in ("P_RGB", "P_BGR", "P_YUV"):
                    img = flat.view(3, shape[1], stride)[:, :, :shape[2]]
                else:
                    img = flat.view(shape[0], stride)[:, :shape[1]]
                b.images[i] = img
                if to_host:
                    b.d2h_bytes += nbytes
        b.infos = infos
        return b

    def _stream(self, stream):
        if stream is None:
            stream = torch.cuda.current_stream(self.device)
        return ctypes.c_void_p(stream.cuda_stream if hasattr(stream, "cuda_stream") else int(stream))

    def _check(self, rc, what):
        if rc != _lib.OK:
            msg = self._L.b200jpegGetLastError(self._h)
            raise DecodeError(f"{what}: {_lib.STATUS_NAMES[rc]} ({msg.decode() if msg else ''})")

    def prepare(self, batch, stream=None):
        self._check(self._L.b200jpegPrepareBatch(self._h, batch.n, batch.ptrs, batch.lens, batch.outs, self._stream(stream)), "prepare")

    def execute(self, stream=None):
        self._check(self._L.b200jpegExecuteBatch(self._h, self._stream(stream)), "execute")

    def finish(self, batch, stream=None):
        self._check(self._L.b200jpegFinishBatch(self._h, batch.status, self._stream(stream)), "finish")
        return [batch.status[i] for i in range(batch.n)]

    def run(self, batch, stream=None):
        self._check(self._L.b200jpegDecodeBatch(self._h, batch.n, batch.ptrs, batch.lens, batch.outs, batch.status, self._stream(stream)),
                    "decode")
        return [batch.status[i] for i in range(batch.n)]

    # ------------------------------------------------------------------ the user-facing call
    def decode(self, srcs, fmt="I_RGB", rois=None, to_host=False, stream=None, orientations=None):
        """Decode a batch.  Returns a list with one uint8 tensor per source (None where the
        stream was rejected; the per-image status codes are in ``self.last_status``)."""
        single = isinstance(srcs, (bytes, bytearray, np.ndarray))
        if single:
            srcs = [srcs]
            rois = None if rois is None else [rois]
        with torch.cuda.device(self.device):
            batch = self.make_batch(srcs, fmt, rois, to_host, orientations=orientations)
            self.last_status = self.run(batch, stream)
        out = [img if st == _lib.OK else None for img, st in zip(batch.images, self.last_status)]
        return out[0] if single else out

    # ------------------------------------------------------------------ introspection
    def stat(self, which):
        return int(self._L.b200jpegGetStat(self._h, which))

    def set_profiling(self, on=True):
        self._L.b200jpegSetProfiling(self._h, 1 if on else 0)

    def stage_times(self):
        ms = (ctypes.c_float * 16)()
        names = (ctypes.c_char_p * 16)()
        n = self._L.b200jpegGetStageTimes(self._h, ms, names, 16)
        return {names[i].decode(): float(ms[i]) for i in range(n)}

    def coefficients(self, index, n_elems):
        out = np.empty(n_elems, dtype=np.int16)
        self._check(self._L.b200jpegDebugCopyCoefficients(self._h, index, ctypes.c_void_p(out.ctypes.data), out.size), "coefficients")
        return out


class MultiGpuDecoder:
    """One batch across several GPUs of one box: LPT partition by compressed bytes, one pipeline, stream and
    host thread per device, no collective (b200jpegMulti* in include/b200jpeg.h)."""

    def __init__(self, device_ids, options=":fancy_upsampling=0"):
        self._L = _lib.load()
        if not torch.cuda.is_available():
            raise DecodeError("nvimagecodec_b200 needs a CUDA device (there is no CPU fallback)")
        self.device_ids = [int(d) for d in device_ids]
        ids = (ctypes.c_int * len(self.device_ids))(*self.device_ids)
        h = ctypes.c_void_p()
        rc = self._L.b200jpegMultiCreate(ctypes.byref(h), ids, len(self.device_ids), (options or "").encode())
        if rc != _lib.OK:
            raise DecodeError(f"b200jpegMultiCreate failed: {_lib.STATUS_NAMES[rc]}")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.b200jpegMultiDestroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def plan(self, srcs):
        arrs = [_as_u8(s) for s in srcs]
        n = len(arrs)
        lens = (ctypes.c_size_t * max(n, 1))(*[a.size for a in arrs])
        dev = (ctypes.c_int * max(n, 1))()
        rc = self._L.b200jpegMultiPlan(self._h, n, lens, dev)
        if rc != _lib.OK:
            raise DecodeError(f"b200jpegMultiPlan failed: {_lib.STATUS_NAMES[rc]}")
        return arrs, lens, dev

    def decode(self, srcs, fmt="I_RGB"):
        """Returns (images, statuses, device index per image); every image lives on the device that decoded it."""
        arrs, lens, dev = self.plan(srcs)
        n = len(arrs)
        ptrs = (ctypes.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrs])
        outs = (_lib.Output * max(n, 1))()
        status = (ctypes.c_int * max(n, 1))()
        images = []
        for i, a in enumerate(arrs):
            info = _lib.StreamInfo()
            self._L.b200jpegProbe(ctypes.c_void_p(a.ctypes.data), a.size, ctypes.byref(info))
            shape = (info.height, info.width, 3) if fmt in ("I_RGB", "I_BGR") else ((3, info.height, info.width) if fmt.startswith("P_R") or fmt.startswith("P_B") else (info.height, info.width))
            t = torch.empty(shape if info.supported else (1,), dtype=torch.uint8, device=torch.device("cuda", self.device_ids[dev[i]]))
            images.append(t)
            outs[i].buffer = t.data_ptr()
            outs[i].format = _lib.FMT[fmt]
            outs[i].buffer_is_host = 0
            outs[i].row_stride = 0
        for d in set(self.device_ids):
            torch.cuda.synchronize(d)  # the outputs were allocated on torch's streams
        rc = self._L.b200jpegMultiDecodeBatch(self._h, n, ptrs, lens, outs, dev, status)
        if rc != _lib.OK:
            raise DecodeError(f"b200jpegMultiDecodeBatch failed: {_lib.STATUS_NAMES[rc]}")
        st = [status[i] for i in range(n)]
        return [im if s == _lib.OK else None for im, s in zip(images, st)], st, [dev[i] for i in range(n)]
