"""Progressive (SOF2) streams through the CUDA path (reference: libjpeg-turbo's jdphuff.c behind
extensions/libjpeg_turbo/jpeg_mem.cpp:154-219, pinned by test/extensions/libjpeg_turbo_ext_decoder_test.cpp:115-130).
Bar: bit-exact against the oracle (itself pinned against Pillow's libjpeg-turbo in tests/test_oracle.py)."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec(decoder_cls):
    d = decoder_cls(0, options=":fancy_upsampling=1")
    yield d
    d.close()


def _cases():
    from nvimagecodec_b200 import corpus

    out = []
    seed = 100
    for (h, w) in [(213, 200), (64, 64), (375, 500), (17, 9), (8, 8), (600, 801)]:
        for sampling in ("420", "444", "422", "440"):
            for rst in (0, 3):
                seed += 1
                out.append(corpus.encode(corpus.synth(h, w, seed), 60 + (seed * 7) % 38, sampling, rst_interval=rst, progressive=True))
        out.append(corpus.encode(corpus.synth(h, w, seed), 85, gray=True, progressive=True))
    return out


def test_progressive_bit_exact(dec):
    from nvimagecodec_b200 import _lib

    srcs = _cases()
    outs = dec.decode(srcs, "I_RGB")
    assert dec.last_status == [_lib.OK] * len(srcs)
    for i, (s, o) in enumerate(zip(srcs, outs)):
        assert np.array_equal(o.cpu().numpy(), oracle.decode(s, oracle.FMT_I_RGB, fancy=True)), i
    outs = dec.decode(srcs, "I_Y")
    for i, (s, o) in enumerate(zip(srcs, outs)):
        assert np.array_equal(o.cpu().numpy(), oracle.decode(s, oracle.FMT_Y, fancy=True)), i


def test_progressive_coefficients_match_oracle(dec):
    from nvimagecodec_b200 import corpus

    srcs = [corpus.encode(corpus.synth(120, 200, 11), 92, "420", rst_interval=5, progressive=True),
            corpus.encode(corpus.synth(96, 80, 12), 60, "444", progressive=True),
            corpus.encode(corpus.synth(70, 70, 13), 85, gray=True, progressive=True),
            corpus.encode(corpus.synth(1080, 1920, 14), 90, "420", progressive=True)]
    dec.decode(srcs, "I_RGB")
    for i, d in enumerate(srcs):
        ref = oracle.coefficients(d)
        got = dec.coefficients(i, ref.size).reshape(ref.shape)
        assert np.array_equal(got, ref), i


def test_progressive_mixed_with_sequential_roi_and_formats(dec):
    """Progressive and baseline streams in one batch, regions of interest, BGR / planar output."""
    from nvimagecodec_b200 import _lib, corpus

    srcs, rois = [], []
    for k in range(24):
        h, w = 90 + 37 * k, 120 + 53 * k
        srcs.append(corpus.encode(corpus.synth(h, w, 300 + k), 70 + k, ("420", "444", "422")[k % 3], rst_interval=(0, 7)[k % 2],
                                  progressive=(k % 2 == 0)))
        rois.append((h // 5, w // 7, h - h // 6, w - w // 9) if k % 3 else None)
    for fmt, ofmt in (("I_BGR", oracle.FMT_I_BGR), ("P_RGB", oracle.FMT_P_RGB)):
        outs = dec.decode(srcs, fmt, rois)
        assert dec.last_status == [_lib.OK] * len(srcs)
        for i, (s, o) in enumerate(zip(srcs, outs)):
            assert np.array_equal(o.cpu().numpy(), oracle.decode(s, ofmt, fancy=True, roi=rois[i])), (fmt, i)


def test_progressive_damaged_scans_are_reported_and_neighbours_stay_exact(dec):
    from nvimagecodec_b200 import _lib, corpus

    good = corpus.encode(corpus.synth(200, 300, 501), 88, "420", progressive=True)
    base = corpus.encode(corpus.synth(160, 240, 502), 88, "420")
    rng = np.random.default_rng(5)
    srcs = [good, base]
    for k in range(40):
        b = bytearray(good)
        for _ in range(1 + k % 4):
            pos = int(rng.integers(len(b) // 3, len(b) - 2))
            b[pos] = int(rng.integers(0, 255))
        srcs.append(bytes(b[:len(b) - (k % 5) * 37]) if k % 5 else bytes(b))
    outs = dec.decode(srcs, "I_RGB")
    assert dec.last_status[0] == _lib.OK and dec.last_status[1] == _lib.OK
    assert np.array_equal(outs[0].cpu().numpy(), oracle.decode(good))
    assert np.array_equal(outs[1].cpu().numpy(), oracle.decode(base))
    # and the decoder is usable afterwards
    again = dec.decode([good], "I_RGB")
    assert np.array_equal(again[0].cpu().numpy(), oracle.decode(good))


def test_progressive_can_be_switched_off(decoder_cls):
    from nvimagecodec_b200 import _lib, corpus

    d = decoder_cls(0, options=":progressive=0")
    try:
        s = corpus.encode(corpus.synth(64, 64, 9), 80, "420", progressive=True)
        d.decode([s], "I_RGB")
        assert d.last_status == [_lib.ERR_UNSUPPORTED]
    finally:
        d.close()


def test_sixteen_bit_quantisation_tables(decoder_cls):
    """Quantisers above 255 (DQT with Pq = 1: extended-sequential and progressive streams only) take the IDCT stage's plain
    dequantisation path instead of the packed dot products -- from the entry stream, from dense groups and from progressive
    blocks."""
    import io

    PIL = pytest.importorskip("PIL.Image")
    from nvimagecodec_b200 import _lib, corpus

    srcs = []
    for k, (h, w) in enumerate([(120, 168), (64, 64), (250, 333)]):
        img = corpus.synth(h, w, 40 + k)
        qt = [min(1000, 16 + (12 + 3 * k) * i) for i in range(64)]
        for prog in (False, True):
            for sub in (0, 2):
                buf = io.BytesIO()
                PIL.fromarray(img).save(buf, format="JPEG", qtables=[qt, qt], subsampling=sub, progressive=prog)
                srcs.append(buf.getvalue())
    assert any(s[s.find(b"\xff\xdb") + 4] >> 4 == 1 for s in srcs)  # 16-bit tables are really in there
    for opts in (":fancy_upsampling=1", ":fancy_upsampling=1 :entry_stream=0"):
        d = decoder_cls(0, options=opts)
        try:
            outs = d.decode(srcs, "I_RGB")
            assert d.last_status == [_lib.OK] * len(srcs)
            for i, (s, o) in enumerate(zip(srcs, outs)):
                assert np.array_equal(o.cpu().numpy(), oracle.decode(s, oracle.FMT_I_RGB, fancy=True)), (opts, i)
        finally:
            d.close()


def test_progressive_random_shapes_and_qualities(dec):
    """Forty random progressive streams (sizes 1..700, quality 15..98, every sampling, restart intervals, gray) in one batch,
    box and triangle upsampling."""
    from nvimagecodec_b200 import _lib, corpus

    rng = np.random.default_rng(2024)
    srcs = []
    for k in range(40):
        h, w = int(rng.integers(1, 700)), int(rng.integers(1, 700))
        srcs.append(corpus.encode(corpus.synth(h, w, 4000 + k), int(rng.integers(15, 99)), ("420", "444", "422", "440", "411")[k % 5],
                                  rst_interval=int(rng.integers(0, 6)) if k % 3 == 0 else 0, gray=(k % 11 == 10), progressive=True))
    outs = dec.decode(srcs, "I_RGB")
    assert dec.last_status == [_lib.OK] * len(srcs)
    for i, (s, o) in enumerate(zip(srcs, outs)):
        assert np.array_equal(o.cpu().numpy(), oracle.decode(s, oracle.FMT_I_RGB, fancy=True)), i


def test_more_chains_than_the_gpu_holds(dec):
    """Restart intervals cut every scan into chains; 400 small images make far more one-warp CTAs than are resident at once.
    The launch must finish -- a chain's predecessor always has the lower CTA index -- and the pixels must match."""
    from nvimagecodec_b200 import _lib, corpus

    base = [corpus.encode(corpus.synth(96 + 8 * (k % 5), 128 + 16 * (k % 3), 7000 + k), 85, ("420", "444")[k % 2], rst_interval=1 + k % 3,
                          progressive=True) for k in range(16)]
    srcs = [base[i % len(base)] for i in range(400)]
    outs = dec.decode(srcs, "I_RGB")
    assert dec.last_status == [_lib.OK] * len(srcs)
    refs = [oracle.decode(b) for b in base]
    for i in range(0, len(srcs), 5):
        assert np.array_equal(outs[i].cpu().numpy(), refs[i % len(base)]), i


def test_progressive_damaged_headers_tables_and_scans(dec):
    """400 damaged progressive streams (bytes flipped anywhere incl. SOS parameters and Huffman tables, truncation, dropped
    slices) in batches between pristine ones: whatever the verdicts, nothing crashes or hangs and the neighbours stay exact."""
    from nvimagecodec_b200 import corpus

    rng = np.random.default_rng(99)
    seeds = [corpus.encode(corpus.synth(72 + 16 * k, 88 + 24 * k, 60 + k), 82, ("420", "444", "422")[k % 3], rst_interval=(0, 2, 5)[k % 3],
                           progressive=True) for k in range(6)]
    good = seeds[0]
    ref = oracle.decode(good)
    for b0 in range(0, 400, 100):
        srcs = [good]
        for it in range(b0, b0 + 100):
            b = bytearray(seeds[it % len(seeds)])
            kind = it % 4
            if kind == 0:
                for _ in range(1 + it % 5):
                    b[int(rng.integers(2, len(b)))] = int(rng.integers(0, 256))
            elif kind == 1:
                marks = [i for i in range(len(b) - 1) if b[i] == 0xFF and b[i + 1] in (0xC2, 0xC4, 0xDA, 0xDB, 0xDD)]
                m = marks[int(rng.integers(0, len(marks)))]
                b[min(len(b) - 1, m + int(rng.integers(2, 24)))] = int(rng.integers(0, 256))
            elif kind == 2:
                b = b[:int(rng.integers(4, len(b)))]
            else:
                a = int(rng.integers(2, len(b) - 2))
                del b[a:a + int(rng.integers(1, 64))]
            srcs.append(bytes(b))
        srcs.append(good)
        outs = dec.decode(srcs, "I_RGB")
        assert dec.last_status[0] == 0 and dec.last_status[-1] == 0
        assert np.array_equal(outs[0].cpu().numpy(), ref) and np.array_equal(outs[-1].cpu().numpy(), ref)
