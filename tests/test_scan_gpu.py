"""GPU test of the K0 scan kernel on its own (b200jpegScanSegment) against a byte-level restatement."""
import ctypes

import numpy as np
import pytest

import oracle
from tests.scan_ref import unstuff, words_to_bytes

pytestmark = pytest.mark.gpu


def _segments():
    from nvimagecodec_b200 import corpus

    segs = []
    for i, (h, w, samp, rst, q) in enumerate([(300, 400, "420", 7, 95), (64, 64, "444", 0, 50), (1080, 1920, "420", 120, 92),
                                              (33, 500, "422", 1, 99)]):
        data = corpus.encode(corpus.synth(h, w, 300 + i), q, samp, rst)
        off = oracle.info(data)["scan_offset"]
        segs.append(data[off:])  # includes the EOI marker and, deliberately, nothing after it
    # hand-made: fill bytes before a marker, stuffed FF runs, trailing garbage after EOI
    segs.append(bytes([0x12, 0xFF, 0x00, 0xFF, 0x00, 0x34, 0xFF, 0xFF, 0xD0, 0x56, 0xFF, 0x00, 0xFF, 0xD1, 0x78, 0xFF, 0xD9, 0xAA, 0xFF, 0xD2]))
    return segs


def run_scan(L, fn, handle, seg, device, expected=None):
    import torch

    ref_bytes, ref_bnd, ref_marks = unstuff(seg)
    exp = len(ref_bnd) if expected is None else expected
    raw = torch.zeros((len(seg) + 15) // 16 * 16 + 16, dtype=torch.uint8, device=device)
    raw[:len(seg)] = torch.frombuffer(bytearray(seg), dtype=torch.uint8).to(device)
    words = torch.full(((len(seg) + 15) // 16 * 16 + 32,), 0xEE, dtype=torch.uint8, device=device)
    bnd = torch.full((exp + 1,), 0x7FFFFFFF, dtype=torch.int32, device=device)
    res = torch.zeros(4, dtype=torch.int32, device=device)
    stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    args = [ctypes.c_void_p(raw.data_ptr()), len(seg), exp, ctypes.c_void_p(words.data_ptr()), ctypes.c_void_p(bnd.data_ptr()), exp + 1,
            ctypes.c_void_p(res.data_ptr()), stream]
    rc = fn(*(([handle] if handle is not None else []) + args))
    torch.cuda.synchronize(device)
    return rc, words.cpu().numpy(), bnd.cpu().numpy(), res.cpu().numpy(), (ref_bytes, ref_bnd, ref_marks)


def check_against_reference(words, bnd, res, ref):
    ref_bytes, ref_bnd, ref_marks = ref
    in_sequence = all(m == 0xD0 + (i & 7) for i, m in enumerate(ref_marks))
    assert int(res[0]) == len(ref_bytes)
    assert int(res[1]) == len(ref_bnd)
    assert (int(res[3]) == 0) == in_sequence
    assert words_to_bytes(words, len(ref_bytes)) == ref_bytes
    assert [int(x) for x in bnd[:len(ref_bnd)]] == ref_bnd
    assert int(bnd[len(ref_bnd)]) == len(ref_bytes)
    if in_sequence:
        assert int(res[2]) == (len(ref_bytes) * 8 + 1023) // 1024


def test_scan_segment_matches_byte_level_restatement(decoder_cls):
    import torch

    dec = decoder_cls(0)
    L = dec._L
    dev = torch.device("cuda", 0)
    for seg in _segments():
        rc, words, bnd, res, ref = run_scan(L, L.b200jpegScanSegment, dec._h, seg, dev)
        assert rc == 0
        check_against_reference(words, bnd, res, ref)
    # wrong expected restart count is flagged, not fatal
    seg = _segments()[0]
    rc, words, bnd, res, ref = run_scan(L, L.b200jpegScanSegment, dec._h, seg, dev, expected=len(unstuff(seg)[1]) + 3)
    assert rc == 0 and int(res[3]) & 1
    dec.close()
