"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle and the
committed libjpeg-turbo golden vectors.  Bar: bit-exact (integer / byte work)."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
OFMT = {"I_RGB": oracle.FMT_I_RGB, "I_BGR": oracle.FMT_I_BGR, "P_RGB": oracle.FMT_P_RGB, "P_BGR": oracle.FMT_P_BGR,
        "I_Y": oracle.FMT_Y, "P_Y": oracle.FMT_Y}


def _gold():
    with open(os.path.join(GOLD, "golden.json")) as f:
        entries = json.load(f)["entries"]
    return entries, [open(os.path.join(GOLD, "jpeg", e["name"] + ".jpg"), "rb").read() for e in entries]


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def dec_fancy(decoder_cls):
    d = decoder_cls(0, options=":fancy_upsampling=1")
    yield d
    d.close()


@pytest.fixture(scope="module")
def dec_plain(decoder_cls):
    d = decoder_cls(0)  # binding default ":fancy_upsampling=0"
    yield d
    d.close()


def test_native_library_is_what_runs(dec_fancy):
    from nvimagecodec_b200 import _lib

    assert os.path.exists(_lib.LIB_PATH)
    with open("/proc/self/maps") as f:
        assert "libb200jpeg_ext.so" in f.read()


def test_golden_vectors_bit_exact(dec_fancy):
    """Every committed libjpeg-turbo vector (config #1 substitutes + edge cases), one batch."""
    from nvimagecodec_b200 import _lib

    entries, datas = _gold()
    outs = dec_fancy.decode(datas, "I_RGB")
    assert dec_fancy.last_status == [_lib.OK] * len(datas)
    assert dec_fancy.stat(_lib.STAT_KERNEL_LAUNCHES) >= 8
    for e, o in zip(entries, outs):
        got = o.cpu().numpy()
        assert got.shape == (e["height"], e["width"], 3), e["name"]
        assert _sha(got) == e["sha256_rgb_fancy"], e["name"]
    outs = dec_fancy.decode(datas, "I_Y")
    for e, o in zip(entries, outs):
        assert _sha(o.cpu().numpy()) == e["sha256_y"], e["name"]


def test_golden_vectors_box_upsampling_and_roi_bit_exact(dec_plain):
    """fancy_upsampling=0 (the Python binding / nvimtrans default) and cropped decodes against the
    vectors produced by the reference's own libjpeg_turbo::Uncompress."""
    from nvimagecodec_b200 import _lib

    entries, datas = _gold()
    outs = dec_plain.decode(datas, "I_RGB")
    assert dec_plain.last_status == [_lib.OK] * len(datas)
    for e, o in zip(entries, outs):
        assert _sha(o.cpu().numpy()) == e["sha256_rgb_box"], e["name"]
    outs = dec_plain.decode(datas, "I_RGB", [tuple(e["roi"]) for e in entries])
    for e, o in zip(entries, outs):
        assert _sha(o.cpu().numpy()) == e["sha256_roi_box"], e["name"]


@pytest.mark.parametrize("fmt", ["I_RGB", "I_BGR", "P_RGB", "P_BGR", "I_Y"])
@pytest.mark.parametrize("fancy", [True, False])
def test_formats_and_upsampling_modes_vs_oracle(decoder_cls, dec_fancy, dec_plain, fmt, fancy):
    from nvimagecodec_b200 import _lib

    entries, datas = _gold()
    pick = [i for i, e in enumerate(entries) if e["name"].startswith("syn_") or "padlock_4" in e["name"]]
    srcs = [datas[i] for i in pick]
    dec = dec_fancy if fancy else dec_plain
    outs = dec.decode(srcs, fmt)
    assert dec.last_status == [_lib.OK] * len(srcs)
    for d, o in zip(srcs, outs):
        ref = oracle.decode(d, OFMT[fmt], fancy=fancy)
        assert np.array_equal(o.cpu().numpy(), ref)


def test_coefficients_match_oracle(dec_fancy):
    from nvimagecodec_b200 import corpus

    srcs = [corpus.encode(corpus.synth(120, 200, 11), 92, "420", rst_interval=5),
            corpus.encode(corpus.synth(96, 80, 12), 60, "444", optimize=True),
            corpus.encode(corpus.synth(70, 70, 13), 85, gray=True)]
    dec_fancy.decode(srcs, "I_RGB")
    for i, d in enumerate(srcs):
        ref = oracle.coefficients(d)
        got = dec_fancy.coefficients(i, ref.size).reshape(ref.shape)
        assert np.array_equal(got, ref)


def test_roi_equals_crop_and_mixed_requests(dec_plain):
    """ROI == crop of the full decode (test/python/test_decode.py:381), mixed formats in one batch."""
    from nvimagecodec_b200 import _lib, corpus

    rng = np.random.default_rng(3)
    srcs, fmts, rois = [], [], []
    for i in range(24):
        h, w = int(rng.integers(20, 300)), int(rng.integers(20, 300))
        samp = ["444", "422", "420", "440", "411"][i % 5]
        srcs.append(corpus.encode(corpus.synth(h, w, 40 + i), 85, samp, rst_interval=int(rng.integers(0, 9))))
        y0, x0 = int(rng.integers(0, h)), int(rng.integers(0, w))
        y1, x1 = int(rng.integers(y0 + 1, h + 1)), int(rng.integers(x0 + 1, w + 1))
        rois.append((y0, x0, y1, x1) if i % 4 else None)
        fmts.append(["I_RGB", "I_Y", "P_RGB", "I_BGR"][i % 4])
    outs = dec_plain.decode(srcs, fmts, rois)
    assert dec_plain.last_status == [_lib.OK] * len(srcs)
    for d, f, r, o in zip(srcs, fmts, rois, outs):
        assert np.array_equal(o.cpu().numpy(), oracle.decode(d, OFMT[f], fancy=False, roi=r))


def test_host_output_and_padded_stride(dec_fancy):
    from nvimagecodec_b200 import _lib, corpus
    import torch

    srcs = [corpus.encode(corpus.synth(61, 45, 5), 90, "420"), corpus.encode(corpus.synth(33, 130, 6), 90, "422", 2)]
    outs = dec_fancy.decode(srcs, "I_RGB", to_host=True)
    for d, o in zip(srcs, outs):
        assert not o.is_cuda
        assert np.array_equal(o.numpy(), oracle.decode(d, oracle.FMT_I_RGB, True))
    with torch.cuda.device(0):
        b = dec_fancy.make_batch(srcs, ["P_RGB", "I_RGB"], row_align=64)
        b.storage.fill_(7)
        assert dec_fancy.run(b) == [_lib.OK, _lib.OK]
    assert np.array_equal(b.images[0].cpu().numpy(), oracle.decode(srcs[0], oracle.FMT_P_RGB, True))
    assert np.array_equal(b.images[1].cpu().numpy(), oracle.decode(srcs[1], oracle.FMT_I_RGB, True))


def test_bad_and_unsupported_streams_report_per_image_status(dec_fancy):
    cv2 = pytest.importorskip("cv2")
    from nvimagecodec_b200 import _lib, corpus

    img = corpus.synth(90, 120, 9)
    good = corpus.encode(img, 90, "420", rst_interval=3)
    ok, prog = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])
    prog = prog.tobytes()
    prog = prog[:prog.rfind(b"\xff\xda")] + b"\xff\xd9"  # a progression that stops early: left to libjpeg (it smooths such blocks)
    info = oracle.info(good)
    cut = bytearray(good)
    mid = info["scan_offset"] + (len(good) - info["scan_offset"]) // 2
    trunc = bytes(cut[:mid]) + b"\xff\xd9"
    # damage the entropy data: the decoder must flag it, never crash or write out of bounds
    noisy = bytearray(good)
    rng = np.random.default_rng(1)
    for p in rng.integers(info["scan_offset"] + 10, len(good) - 10, size=40):
        if noisy[p] != 0xFF and noisy[p - 1] != 0xFF:
            noisy[p] = (noisy[p] ^ 0x5A) if (noisy[p] ^ 0x5A) != 0xFF else 0x11
    badrst = good.replace(b"\xff\xd1", b"\xff\xd5", 1)
    srcs = [good, prog, b"garbage", trunc, bytes(noisy), badrst, good]
    outs = dec_fancy.decode(srcs, "I_RGB")
    st = dec_fancy.last_status
    assert st[0] == _lib.OK and st[6] == _lib.OK
    assert st[1] == _lib.ERR_UNSUPPORTED
    assert st[2] == _lib.ERR_NOT_JPEG
    assert st[3] == _lib.ERR_CORRUPT
    assert st[5] == _lib.ERR_CORRUPT
    assert st[4] in (_lib.ERR_CORRUPT, _lib.OK)  # random damage may still parse; it must not take the batch down
    ref = oracle.decode(good, oracle.FMT_I_RGB, True)
    assert np.array_equal(outs[0].cpu().numpy(), ref) and np.array_equal(outs[6].cpu().numpy(), ref)


def test_resync_path_and_no_rst_large_stream(decoder_cls):
    """Config #4 shape (4:4:4, q95, no RST, large scan): relies on self-synchronisation only.  The
    debug_no_warm hook starts every sequence from a bare guess, which forces the fix-up path."""
    from nvimagecodec_b200 import _lib, corpus

    big = corpus.encode(corpus.synth(1080, 1920, 21), 95, "444")
    flat = np.zeros((1200, 1600, 3), np.uint8)
    flat[:, :] = (12, 200, 90)
    flat[500:520, 300:900] = 255
    srcs = [big, corpus.encode(flat, 90, "420"), corpus.encode(corpus.synth(600, 800, 22), 90, "420", optimize=True)]
    refs = [oracle.decode(d, oracle.FMT_I_RGB, True) for d in srcs]
    for opts, expect_resync in [(":fancy_upsampling=1", False), (":fancy_upsampling=1 b200_jpeg_decoder:debug_no_warm=1", True)]:
        dec = decoder_cls(0, options=opts)
        outs = dec.decode(srcs, "I_RGB")
        assert dec.last_status == [_lib.OK] * len(srcs)
        if expect_resync:
            assert dec.stat(_lib.STAT_RESYNC_IMAGES) >= 1
        for o, r in zip(outs, refs):
            assert np.array_equal(o.cpu().numpy(), r)
        dec.close()


@pytest.mark.parametrize("bits", [512, 1024, 2048, 4096])
def test_every_subsequence_length_is_bit_exact(decoder_cls, bits):
    """The host picks 1024 / 2048 / 4096-bit subsequences per batch; force each, with and without the
    warm-up / probe heuristics, on streams with and without restart markers."""
    from nvimagecodec_b200 import _lib, corpus

    srcs = [corpus.encode(corpus.synth(700, 900, 31), 93, "420"), corpus.encode(corpus.synth(500, 640, 32), 96, "444"),
            corpus.encode(corpus.synth(480, 360, 33), 80, "422", rst_interval=7), corpus.encode(corpus.synth(90, 90, 34), 90, gray=True)]
    refs = [oracle.decode(d, oracle.FMT_I_RGB, True) for d in srcs]
    for extra in ("", " b200_jpeg_decoder:debug_no_warm=1"):
        dec = decoder_cls(0, options=f":fancy_upsampling=1 b200_jpeg_decoder:subseq_bits={bits}" + extra)
        outs = dec.decode(srcs, "I_RGB")
        assert dec.last_status == [_lib.OK] * len(srcs)
        for o, r in zip(outs, refs):
            assert np.array_equal(o.cpu().numpy(), r)
        dec.close()


def test_deterministic_and_reusable(dec_plain):
    """Sequential reuse of one decoder gives byte-identical results (the reference's
    concurrent_decode_test.cpp demands equality with a sequential decode)."""
    from nvimagecodec_b200 import corpus

    a = [corpus.encode(corpus.synth(200, 300, 70 + i), 80, ["420", "444", "422"][i % 3], i) for i in range(9)]
    b = [corpus.encode(corpus.synth(50, 60, 80 + i), 90, "420") for i in range(3)]
    first = [o.cpu().numpy().copy() for o in dec_plain.decode(a, "I_RGB")]
    dec_plain.decode(b, "I_RGB")
    dec_plain.decode([], "I_RGB")
    second = [o.cpu().numpy() for o in dec_plain.decode(a, "I_RGB")]
    for x, y in zip(first, second):
        assert np.array_equal(x, y)


def test_full_size_properties_c2_shape(dec_plain, dec_fancy):
    """BASELINE config #2 shape (1920x1080 4:2:0 q90): size-independent properties plus the oracle
    on a couple of images (the scalar oracle needs ~0.1 s per image at this size)."""
    from nvimagecodec_b200 import _lib, corpus

    srcs = [corpus.encode(corpus.synth(1080, 1920, s), 90, "420") for s in range(4)]
    rgb = [o.cpu().numpy() for o in dec_plain.decode(srcs, "I_RGB")]
    assert dec_plain.last_status == [_lib.OK] * 4
    prgb = [o.cpu().numpy() for o in dec_plain.decode(srcs, "P_RGB")]
    bgr = [o.cpu().numpy() for o in dec_plain.decode(srcs, "I_BGR")]
    roi = (100, 37, 1001, 1900)
    crop = [o.cpu().numpy() for o in dec_plain.decode(srcs, "I_RGB", [roi] * 4)]
    for i in range(4):
        assert np.array_equal(prgb[i], rgb[i].transpose(2, 0, 1))
        assert np.array_equal(bgr[i], rgb[i][..., ::-1])
        assert np.array_equal(crop[i], rgb[i][roi[0]:roi[2], roi[1]:roi[3]])
    for i in range(2):
        assert np.array_equal(rgb[i], oracle.decode(srcs[i], oracle.FMT_I_RGB, fancy=False))
    f = dec_fancy.decode(srcs[:1], "I_RGB")[0].cpu().numpy()
    assert np.array_equal(f, oracle.decode(srcs[0], oracle.FMT_I_RGB, fancy=True))
    algo = dec_plain.stat(_lib.STAT_ALGO_BYTES)
    assert algo > 0


def test_concurrent_decoders_equal_sequential(decoder_cls):
    """Several decoders driven from several host threads at once (one decoder per thread, like the reference's
    test/python/test_decode_concurrent.py) must reproduce the sequential result byte for byte
    (test/api/concurrent_decode_test.cpp demands the same of the reference's decoders)."""
    import threading

    import torch
    from nvimagecodec_b200 import _lib, corpus

    sets = []
    for s in range(4):
        sets.append([corpus.encode(corpus.synth(150 + 40 * s + 7 * i, 210 + 13 * i, 2000 + 10 * s + i), 70 + 5 * s,
                                   ["420", "444", "422"][(s + i) % 3], (s + i) % 6) for i in range(10)])
    seq = decoder_cls(0, options=":fancy_upsampling=1")
    want = [[o.cpu().numpy().copy() for o in seq.decode(srcs, "I_RGB")] for srcs in sets]
    seq.close()
    results, errors = [None] * len(sets), []

    def work(j):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                dec = decoder_cls(0, options=":fancy_upsampling=1")
                for _ in range(3):
                    outs = dec.decode(sets[j], "I_RGB")
                    assert dec.last_status == [_lib.OK] * len(sets[j])
                results[j] = [o.cpu().numpy() for o in outs]
                dec.close()
        except Exception as e:  # surfaced below: an assert in a thread would otherwise vanish
            errors.append(e)

    threads = [threading.Thread(target=work, args=(j,)) for j in range(len(sets))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for got, ref in zip(results, want):
        for a, b in zip(got, ref):
            assert np.array_equal(a, b)


def _add_fill_bytes(data):
    """Insert optional 0xFF fill bytes before every RSTn marker and before EOI (T.81 B.1.1.2 allows them;
    the reference's parser tests exercise the same in headers, test/parsers/jpeg_test.cpp:361)."""
    off = oracle.info(data)["scan_offset"]
    head, scan = bytearray(data[:off]), data[off:]
    out = bytearray()
    i = 0
    while i < len(scan):
        if scan[i] == 0xFF and i + 1 < len(scan) and (0xD0 <= scan[i + 1] <= 0xD7 or scan[i + 1] == 0xD9):
            out += b"\xff" * (1 + (i % 3))
        out.append(scan[i])
        i += 1
    return bytes(head + out)


def test_awkward_but_valid_streams(dec_fancy, dec_plain):
    """Restart interval of one MCU (a boundary every few bytes), fill bytes before markers, gray with RST,
    very wide / very tall strips, 4:1:1 and 4:4:0 through the generic and the specialised colour paths."""
    from nvimagecodec_b200 import _lib, corpus

    srcs = [corpus.encode(corpus.synth(300, 400, 51), 92, "420", rst_interval=1),
            corpus.encode(corpus.synth(240, 320, 52), 85, "444", rst_interval=1),
            _add_fill_bytes(corpus.encode(corpus.synth(200, 300, 53), 90, "422", rst_interval=4)),
            _add_fill_bytes(corpus.encode(corpus.synth(64, 64, 54), 75, "420", rst_interval=2)),
            corpus.encode(corpus.synth(8, 16000, 55), 90, "420"),
            corpus.encode(corpus.synth(12000, 9, 56), 90, "422", rst_interval=3),
            corpus.encode(corpus.synth(333, 777, 57), 88, "411", rst_interval=5),
            corpus.encode(corpus.synth(333, 777, 58), 88, "440")]
    import cv2

    ok, g = cv2.imencode(".jpg", cv2.cvtColor(corpus.synth(150, 222, 59), cv2.COLOR_RGB2GRAY), [cv2.IMWRITE_JPEG_QUALITY, 90, cv2.IMWRITE_JPEG_RST_INTERVAL, 3])
    srcs.append(g.tobytes())
    for dec, fancy in ((dec_fancy, True), (dec_plain, False)):
        for fmt in ("I_RGB", "P_BGR", "I_Y"):
            outs = dec.decode(srcs, fmt)
            assert dec.last_status == [_lib.OK] * len(srcs), dec.last_status
            for d, o in zip(srcs, outs):
                assert np.array_equal(o.cpu().numpy(), oracle.decode(d, OFMT[fmt], fancy=fancy))


def _as_sof1(data):
    """Relabel a baseline stream as extended-sequential Huffman (SOF1): every baseline stream is a valid SOF1 stream."""
    i = data.index(b"\xff\xc0")
    return data[:i] + b"\xff\xc1" + data[i + 2:]


def test_ycbcr_output_and_extended_sequential(dec_fancy, dec_plain):
    """I_YUV / P_YUV (libjpeg's out_color_space = JCS_YCbCr: upsampled samples, no colour transform;
    the reference allows it for 4:4:4 streams, libjpeg_turbo_decoder.cpp:180-189) and SOF1 streams."""
    from nvimagecodec_b200 import _lib, corpus

    srcs = [corpus.encode(corpus.synth(120, 200, 91), 90, "444"), corpus.encode(corpus.synth(97, 131, 92), 80, "444", 4),
            corpus.encode(corpus.synth(64, 80, 93), 90, "420"), corpus.encode(corpus.synth(50, 70, 94), 90, "422", 2)]
    for dec, fancy in ((dec_fancy, True), (dec_plain, False)):
        for fmt, ofmt in (("I_YUV", oracle.FMT_I_YUV), ("P_YUV", oracle.FMT_P_YUV)):
            outs = dec.decode(srcs, fmt)
            assert dec.last_status == [_lib.OK] * len(srcs)
            for d, o in zip(srcs, outs):
                assert np.array_equal(o.cpu().numpy(), oracle.decode(d, ofmt, fancy=fancy))
    sof1 = [_as_sof1(d) for d in srcs]
    outs = dec_fancy.decode(sof1, "I_RGB")
    assert dec_fancy.last_status == [_lib.OK] * len(sof1)
    for d, s1, o in zip(srcs, sof1, outs):
        ref = oracle.decode(d, oracle.FMT_I_RGB, fancy=True)
        assert np.array_equal(oracle.decode(s1, oracle.FMT_I_RGB, fancy=True), ref)
        assert np.array_equal(o.cpu().numpy(), ref)


def test_exif_orientations_applied_while_storing(dec_fancy):
    """All eight EXIF orientations, every output layout, with and without a crop: the stored image must be the
    oriented crop of the plain decode (semantics pinned against OpenCV / Pillow in tests/test_oracle.py)."""
    from nvimagecodec_b200 import _lib, corpus

    srcs = [corpus.encode(corpus.synth(75, 131, 301), 90, "420"), corpus.encode(corpus.synth(64, 40, 302), 85, "444", 3),
            corpus.encode(corpus.synth(33, 90, 303), 90, "422"), corpus.encode(corpus.synth(50, 50, 304), 90, "411")]
    rois = [None, (5, 3, 60, 31), None, (10, 0, 41, 50)]
    for fmt in ("I_RGB", "P_BGR", "I_Y", "P_YUV"):
        for code in range(1, 9):
            outs = dec_fancy.decode(srcs, fmt, rois, orientations=[code] * len(srcs))
            assert dec_fancy.last_status == [_lib.OK] * len(srcs)
            for d, r, o in zip(srcs, rois, outs):
                ref = oracle.apply_orientation(oracle.decode(d, OFMT.get(fmt, oracle.FMT_P_YUV), fancy=True, roi=r), code, planar=fmt.startswith("P_"))
                assert np.array_equal(o.cpu().numpy(), ref), (fmt, code)


@pytest.mark.parametrize("fmt", ["I_RGB", "P_RGB", "I_Y"])
def test_tight_rows_at_every_byte_alignment(dec_fancy, fmt):
    """Tightly pitched outputs of odd widths start their rows at every address mod 8; the 8-pixel units of the
    colour kernel then store through the funnel-shifted word path.  Also with a ROI whose left edge is odd."""
    from nvimagecodec_b200 import _lib, corpus

    srcs, rois = [], []
    for i, w in enumerate(range(65, 81)):  # 16 widths: 3*w and w cover all residues mod 8 over the rows
        samp = ["420", "422", "444"][i % 3]
        srcs.append(corpus.encode(corpus.synth(40 + i, w, 900 + i), 85, samp))
        rois.append(None if i % 2 == 0 else (3, 5 + i % 7, 35, w - 2))
    outs = dec_fancy.decode(srcs, fmt, rois=rois)
    assert dec_fancy.last_status == [_lib.OK] * len(srcs)
    for d, roi, o in zip(srcs, rois, outs):
        assert np.array_equal(o.cpu().numpy(), oracle.decode(d, OFMT[fmt], fancy=True, roi=roi))


@pytest.mark.parametrize("extra", ["", " b200_jpeg_decoder:debug_no_warm=1", " b200_jpeg_decoder:subseq_bits=512"])
def test_fuzzed_entropy_data_never_takes_the_batch_down(decoder_cls, extra):
    """A few hundred damaged variants (byte flips anywhere in the scan incl. fresh FF xx pairs, truncations, stray and
    missing RSTn, tail garbage) share batches with pristine images: the call returns, every pristine image stays
    bit-exact (no out-of-bounds write reached a neighbour's coefficients, planes or pixels), and the decoder is reusable."""
    from nvimagecodec_b200 import _lib, corpus

    rng = np.random.default_rng(20261107)
    base = [corpus.encode(corpus.synth(96, 128, 70), 90, "420"), corpus.encode(corpus.synth(120, 88, 71), 75, "444", rst_interval=2),
            corpus.encode(corpus.synth(64, 200, 72), 95, "422", rst_interval=5), corpus.encode(corpus.synth(72, 72, 73), 85, gray=True),
            corpus.encode(corpus.synth(300, 420, 74), 92, "420")]
    refs = [oracle.decode(d, oracle.FMT_I_RGB, True) for d in base]
    dec = decoder_cls(0, options=":fancy_upsampling=1" + extra)
    for rnd in range(4):
        srcs, clean = [], []
        for k in range(60):
            b = int(rng.integers(0, len(base)))
            d = bytearray(base[b])
            so = oracle.info(base[b])["scan_offset"]
            kind = int(rng.integers(0, 6))
            if kind == 0:  # flips, any value (may create markers, stuffing, fill bytes)
                for p in rng.integers(so, len(d) - 2, size=int(rng.integers(1, 30))):
                    d[p] = int(rng.integers(0, 256))
            elif kind == 1:  # truncation, with or without an EOI
                d = d[:int(rng.integers(so + 1, len(d)))] + (bytearray(b"\xff\xd9") if rng.integers(0, 2) else bytearray())
            elif kind == 2:  # stray restart markers
                for p in sorted(rng.integers(so, len(d) - 2, size=3), reverse=True):
                    d[p:p] = bytes([0xFF, 0xD0 + int(rng.integers(0, 8))])
            elif kind == 3:  # a run of FF / 00
                p = int(rng.integers(so, len(d) - 40))
                d[p:p + 32] = bytes([0xFF if rng.integers(0, 2) else 0x00]) * 32
            elif kind == 4:  # garbage behind the data, no EOI
                d = d[:-2] + bytearray(rng.integers(0, 256, size=64, dtype=np.uint8).tobytes())
            else:  # a chunk cut out of the middle
                p = int(rng.integers(so, len(d) - 50))
                del d[p:p + int(rng.integers(1, 40))]
            srcs.append(bytes(d))
            if k % 4 == 0:
                clean.append((len(srcs), b))
                srcs.append(base[b])
        outs = dec.decode(srcs, "I_RGB")
        st = dec.last_status
        assert len(st) == len(srcs)
        for pos, b in clean:
            assert st[pos] == _lib.OK
            assert np.array_equal(outs[pos].cpu().numpy(), refs[b])
    outs = dec.decode(base, "I_RGB")
    assert dec.last_status == [_lib.OK] * len(base)
    for o, r in zip(outs, refs):
        assert np.array_equal(o.cpu().numpy(), r)
    dec.close()


def test_three_scans_of_one_component_each(decoder_cls, dec_fancy, dec_plain):
    """Non-interleaved sequential JPEG (one scan per component, T.81 A.2.2): each scan is decoded as a one-component
    image of the component's size, the colour stage reads the three planes.  tests/remux.py derives such files from
    interleaved ones without touching a coefficient, so the reference result is the oracle's decode of the original."""
    from tests import remux
    from nvimagecodec_b200 import _lib, corpus

    srcs, origs, fmts, rois = [], [], [], []
    cases = [(90, 120, "420", 0, (0, 1, 2)), (50, 70, "444", 5, (0, 1, 2)), (64, 48, "422", 3, (2, 0, 1)), (33, 17, "420", 0, (1, 2, 0)),
             (40, 100, "440", 7, (0, 1, 2)), (48, 64, "411", 0, (0, 1, 2)), (300, 420, "420", 0, (0, 1, 2)), (131, 259, "420", 64, (0, 2, 1))]
    for k, (h, w, samp, rb, order) in enumerate(cases):
        d = corpus.encode(corpus.synth(h, w, 50 + k), 88, samp)
        origs.append(d)
        srcs.append(remux.to_three_scans(d, rb, order))
        fmts.append(["I_RGB", "I_Y", "P_RGB", "I_BGR"][k % 4])
        rois.append(None if k % 3 else (h // 4, w // 5, h - 3, w - 2))
        if k % 2:  # interleaved neighbours in the same batch
            origs.append(d)
            srcs.append(d)
            fmts.append("I_RGB")
            rois.append(None)
    for dec, fancy in ((dec_fancy, True), (dec_plain, False)):
        outs = dec.decode(srcs, fmts, rois)
        assert dec.last_status == [_lib.OK] * len(srcs)
        for d, f, r, o in zip(origs, fmts, rois, outs):
            assert np.array_equal(o.cpu().numpy(), oracle.decode(d, OFMT[f], fancy=fancy, roi=r))
    # the re-synchronisation path (forced: no warm-up, no probe) works on the one-component images and redoes the colour stage
    dec = decoder_cls(0, options=":fancy_upsampling=1 b200_jpeg_decoder:debug_no_warm=1 b200_jpeg_decoder:subseq_bits=512")
    outs = dec.decode(srcs, fmts, rois)
    assert dec.last_status == [_lib.OK] * len(srcs) and dec.stat(_lib.STAT_RESYNC_IMAGES) > 0
    for d, f, r, o in zip(origs, fmts, rois, outs):
        assert np.array_equal(o.cpu().numpy(), oracle.decode(d, OFMT[f], fancy=True, roi=r))
    dec.close()
    # damage in one component's scan is reported for the image, the neighbours are untouched
    bad = bytearray(srcs[0])
    j = bad.index(b"\xff\xda", bad.index(b"\xff\xda") + 2)
    bad[j + 14:j + 30] = bytes(16)
    outs = dec_fancy.decode([bytes(bad), srcs[1], origs[0]], "I_RGB")
    assert dec_fancy.last_status[1:] == [_lib.OK, _lib.OK]
    assert np.array_equal(outs[2].cpu().numpy(), oracle.decode(origs[0], oracle.FMT_I_RGB, True))


def test_huffman_set_cache_crossing_its_limit(decoder_cls):
    """Optimised-Huffman streams carry their own table sets.  More distinct sets than the decoder's cache holds (4096), in
    batches that cross the limit: the sets of a batch stay valid while it is being built (eviction only between batches),
    and equal-looking sets are told apart by content, not by a hash."""
    from nvimagecodec_b200 import _lib, corpus

    rng = np.random.default_rng(77)
    dec = decoder_cls(0, options=":fancy_upsampling=1")
    seen = set()
    for rnd in range(4):
        srcs = []
        for k in range(1300):
            img = rng.integers(0, 256, size=(16, 24, 3), dtype=np.uint8)
            img[:, : int(rng.integers(1, 24))] //= int(rng.integers(1, 9))
            srcs.append(corpus.encode(img, int(rng.integers(40, 99)), ["420", "444", "422"][k % 3], optimize=True))
        for d in srcs:
            i = d.index(b"\xff\xc4")
            seen.add(d[i:d.index(b"\xff\xda")])
        outs = dec.decode(srcs, "I_RGB")
        assert dec.last_status == [_lib.OK] * len(srcs)
        for j in list(range(0, len(srcs), 53)) + [len(srcs) - 1]:
            assert np.array_equal(outs[j].cpu().numpy(), oracle.decode(srcs[j], oracle.FMT_I_RGB, True)), (rnd, j)
    assert len(seen) > 4096 + 600, len(seen)  # the cache limit was crossed inside a batch
    dec.close()


@pytest.mark.parametrize("fancy", [1, 0])
def test_fused_pixel_kernel_bit_exact(decoder_cls, fancy):
    """Option fused_pixels=1: IDCT, upsampling and colour in one kernel (component tiles in shared memory, one tile per warp,
    sliding window for the vertical triangle filter).  Every golden vector, cropped and planar / BGR / Y requests, and a
    full-size image, against the same references as the default pixel kernels."""
    from nvimagecodec_b200 import _lib, corpus

    entries, datas = _gold()
    dec = decoder_cls(0, options=f":fancy_upsampling={fancy} b200_jpeg_decoder:fused_pixels=1")
    outs = dec.decode(datas, "I_RGB")
    assert dec.last_status == [_lib.OK] * len(datas)
    for e, o in zip(entries, outs):
        assert _sha(o.cpu().numpy()) == e["sha256_rgb_fancy" if fancy else "sha256_rgb_box"], e["name"]
    if not fancy:
        outs = dec.decode(datas, "I_RGB", [tuple(e["roi"]) for e in entries])
        for e, o in zip(entries, outs):
            assert _sha(o.cpu().numpy()) == e["sha256_roi_box"], e["name"]
    rng = np.random.default_rng(11)
    srcs, fmts, rois = [], [], []
    for i in range(30):
        h, w = int(rng.integers(17, 700)), int(rng.integers(17, 900))
        samp = ["444", "422", "420", "440"][i % 4]
        srcs.append(corpus.encode(corpus.synth(h, w, 7000 + i), 88, samp, rst_interval=int(rng.integers(0, 9))))
        y0, x0 = int(rng.integers(0, h)), int(rng.integers(0, w))
        rois.append((y0, x0, int(rng.integers(y0 + 1, h + 1)), int(rng.integers(x0 + 1, w + 1))) if i % 3 else None)
        fmts.append(["I_RGB", "I_Y", "P_RGB", "I_BGR", "P_BGR"][i % 5])
    srcs.append(corpus.encode(corpus.synth(1080, 1920, 7100), 90, "420"))
    fmts.append("I_RGB")
    rois.append(None)
    srcs.append(corpus.encode(corpus.synth(90, 70, 7101), 90, gray=True))
    fmts.append("I_RGB")
    rois.append(None)
    outs = dec.decode(srcs, fmts, rois)
    assert dec.last_status == [_lib.OK] * len(srcs)
    for d, f, r, o in zip(srcs, fmts, rois, outs):
        assert np.array_equal(o.cpu().numpy(), oracle.decode(d, OFMT[f], fancy=bool(fancy), roi=r)), (f, r)
    dec.close()
