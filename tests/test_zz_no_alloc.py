"""GPU test of b200jpegScanSegmentNoAlloc, the scan entry point for callers that cannot allocate.

Some users of this library run where device memory can no longer be requested: everything was
handed to a framework's pool or a fixed budget at start-up, or the call sits on a path where the
device-wide stall of cudaMalloc / cudaFree is not acceptable.  For them the library carries a small
fixed scratch array and this entry point, which allocates nothing.  The test calls it through the C
ABI twice in a row (a large segment, then a smaller one, so stale scratch from the first call would
show) and requires exactly the bytes b200jpegScanSegment -- the allocating entry point -- produces.
It lives in a file pytest collects last; it also passes when run alone in a fresh process.
"""
import ctypes

import numpy as np
import pytest

import oracle
from tests.scan_ref import unstuff

pytestmark = pytest.mark.gpu


def _segment(h, w, samp, rst, q, seed):
    from nvimagecodec_b200 import corpus

    data = corpus.encode(corpus.synth(h, w, seed), q, samp, rst)
    return data[oracle.info(data)["scan_offset"]:]


def _run(fn, handle, seg, expected, message=None):
    import torch

    dev = torch.device("cuda", 0)
    pad = (len(seg) + 15) // 16 * 16
    raw = torch.zeros(pad + 16, dtype=torch.uint8, device=dev)
    raw[:len(seg)] = torch.frombuffer(bytearray(seg), dtype=torch.uint8).to(dev)
    words = torch.zeros(pad + 32, dtype=torch.uint8, device=dev)
    bnd = torch.full((expected + 1,), -1, dtype=torch.int32, device=dev)
    res = torch.full((4,), -1, dtype=torch.int32, device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    args = [ctypes.c_void_p(raw.data_ptr()), len(seg), expected, ctypes.c_void_p(words.data_ptr()), ctypes.c_void_p(bnd.data_ptr()),
            expected + 1, ctypes.c_void_p(res.data_ptr()), stream]
    if handle is not None:
        rc = fn(handle, *args)
    else:
        rc = fn(*args, message, 0 if message is None else len(message))
    torch.cuda.synchronize(dev)
    return rc, words.cpu().numpy(), bnd.cpu().numpy(), res.cpu().numpy()


def test_no_alloc_scan_matches_allocating_entry_point(decoder_cls):
    import torch
    from nvimagecodec_b200 import _lib

    dec = decoder_cls(0)
    L = dec._L
    segs = [_segment(1080, 1920, "420", 60, 93, 901),   # ~0.6 MB, 135 restart markers
            _segment(120, 200, "444", 3, 80, 902)]      # small: must not see anything the first call left behind
    for seg in segs:  # two calls in a row
        n_rst = len(unstuff(seg)[1])
        assert len(seg) <= _lib.NOALLOC_MAX_SEGMENT_BYTES
        rc_a, words_a, bnd_a, res_a = _run(L.b200jpegScanSegment, dec._h, seg, n_rst)
        msg = ctypes.create_string_buffer(256)
        torch.cuda.synchronize()
        rc_b, words_b, bnd_b, res_b = _run(L.b200jpegScanSegmentNoAlloc, None, seg, n_rst, msg)
        assert rc_a == _lib.OK and rc_b == _lib.OK, msg.value
        assert msg.value == b""
        # bit-exact, the tolerance every test of this integer path uses
        assert np.array_equal(res_a, res_b)
        assert int(res_b[3]) == 0 and int(res_b[1]) == n_rst
        nbytes = int(res_b[0])
        n4 = (nbytes + 3) // 4 * 4
        assert np.array_equal(words_a[:n4], words_b[:n4])
        assert np.array_equal(bnd_a, bnd_b)
        assert int(bnd_b[n_rst]) == nbytes
    # beyond the fixed scratch: refused with a clear error, nothing launched, outputs untouched
    dev = torch.device("cuda", 0)
    big = _lib.NOALLOC_MAX_SEGMENT_BYTES + 1
    raw = torch.zeros(big + 32, dtype=torch.uint8, device=dev)
    words = torch.full((big + 64,), 0x5A, dtype=torch.uint8, device=dev)
    bnd = torch.full((1,), 77, dtype=torch.int32, device=dev)
    res = torch.full((4,), 99, dtype=torch.int32, device=dev)
    msg = ctypes.create_string_buffer(256)
    rc = L.b200jpegScanSegmentNoAlloc(ctypes.c_void_p(raw.data_ptr()), big, 0, ctypes.c_void_p(words.data_ptr()), ctypes.c_void_p(bnd.data_ptr()),
                                      1, ctypes.c_void_p(res.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream), msg, 256)
    torch.cuda.synchronize()
    assert rc == _lib.ERR_TOO_LARGE
    assert b"8 MiB" in msg.value
    assert int(res.cpu()[0]) == 99 and int(words[:16].cpu().max()) == 0x5A
    dec.close()
