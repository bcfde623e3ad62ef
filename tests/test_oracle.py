"""CPU tests: the oracle against the committed libjpeg-turbo golden vectors (no GPU)."""
import hashlib
import io
import json
import os

import numpy as np
import pytest

import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _entries():
    with open(os.path.join(GOLD, "golden.json")) as f:
        return json.load(f)["entries"]


def _data(e):
    with open(os.path.join(GOLD, "jpeg", e["name"] + ".jpg"), "rb") as f:
        return f.read()


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


ENTRIES = _entries()


def test_golden_has_config1_and_edge_cases():
    names = [e["name"] for e in ENTRIES]
    assert sum(n.startswith("c1_") for n in names) >= 10
    assert any(e["width"] == 1 and e["height"] == 1 for e in ENTRIES)


@pytest.mark.parametrize("e", ENTRIES, ids=[e["name"] for e in ENTRIES])
def test_oracle_matches_libjpeg_turbo_golden(e):
    data = _data(e)
    rgb = oracle.decode(data, oracle.FMT_I_RGB, fancy=True)
    assert rgb.shape == (e["height"], e["width"], 3)
    assert _sha(rgb) == e["sha256_rgb_fancy"]
    if "rgb_fancy" in e:
        assert np.array_equal(rgb.reshape(-1), np.asarray(e["rgb_fancy"], dtype=np.uint8))
    y = oracle.decode(data, oracle.FMT_Y, fancy=True)
    assert _sha(y) == e["sha256_y"]
    # fancy_upsampling=0 and a cropped decode, as produced by the reference's own jpeg_mem.cpp
    assert _sha(oracle.decode(data, oracle.FMT_I_RGB, fancy=False)) == e["sha256_rgb_box"]
    assert _sha(oracle.decode(data, oracle.FMT_I_RGB, fancy=False, roi=tuple(e["roi"]))) == e["sha256_roi_box"]


def test_formats_are_permutations_of_rgb():
    data = _data(next(e for e in ENTRIES if e["name"] == "syn_37x53_420_r1"))
    rgb = oracle.decode(data, oracle.FMT_I_RGB)
    assert np.array_equal(oracle.decode(data, oracle.FMT_I_BGR), rgb[..., ::-1])
    assert np.array_equal(oracle.decode(data, oracle.FMT_P_RGB), rgb.transpose(2, 0, 1))
    assert np.array_equal(oracle.decode(data, oracle.FMT_P_BGR), rgb[..., ::-1].transpose(2, 0, 1))


def test_roi_is_crop_of_full_decode():
    # the reference demands roi == full[y0:y1, x0:x1] exactly (test/python/test_decode.py:381)
    data = _data(next(e for e in ENTRIES if e["name"] == "c1_padlock_422"))
    full = oracle.decode(data, oracle.FMT_I_RGB)
    for roi in [(0, 0, 426, 640), (10, 20, 200, 333), (425, 639, 426, 640), (101, 1, 102, 640)]:
        y0, x0, y1, x1 = roi
        assert np.array_equal(oracle.decode(data, oracle.FMT_I_RGB, roi=roi), full[y0:y1, x0:x1])


def test_fancy_off_is_box_replication():
    data = _data(next(e for e in ENTRIES if e["name"] == "syn_48x64_420_r0"))
    a = oracle.decode(data, oracle.FMT_I_RGB, fancy=False)
    b = oracle.decode(data, oracle.FMT_I_RGB, fancy=True)
    assert a.shape == b.shape and not np.array_equal(a, b)
    # luma is untouched by the chroma upsampling mode
    assert np.array_equal(oracle.decode(data, oracle.FMT_Y, fancy=False), oracle.decode(data, oracle.FMT_Y, fancy=True))


def test_coefficients_shape_and_dc_prediction():
    e = next(e for e in ENTRIES if e["name"] == "syn_37x53_420_r1")
    data = _data(e)
    inf = oracle.info(data)
    co = oracle.coefficients(data)
    assert co.shape == (inf["mcus_x"] * inf["mcus_y"], 6, 64)
    assert inf["restart_interval"] == 1
    assert np.abs(co[:, :, 0]).max() > 0


def test_rejects_bad_streams():
    data = _data(ENTRIES[0])
    with pytest.raises(oracle.OracleError):
        oracle.decode(b"not a jpeg at all", oracle.FMT_I_RGB)
    with pytest.raises(oracle.OracleError):
        oracle.decode(data[:200], oracle.FMT_I_RGB)


def test_live_against_pillow_when_available():
    PIL = pytest.importorskip("PIL.Image")
    pytest.importorskip("cv2")
    from nvimagecodec_b200 import corpus

    rng = np.random.default_rng(5)
    for i in range(12):
        h, w = int(rng.integers(1, 90)), int(rng.integers(1, 90))
        samp = ["444", "422", "420", "440", "411"][i % 5]
        data = corpus.encode(corpus.synth(h, w, 900 + i), int(rng.integers(30, 100)), samp, int(rng.integers(0, 5)), bool(i & 1))
        ref = np.asarray(PIL.open(io.BytesIO(data)).convert("RGB"))
        assert np.array_equal(oracle.decode(data, oracle.FMT_I_RGB, fancy=True), ref)


def test_progressive_live_against_pillow_when_available():
    """SOF2 streams (spectral selection + successive approximation, ten scans from libjpeg's simple progression; jdphuff.c)
    against Pillow's libjpeg-turbo, and the incomplete progression that is left to libjpeg."""
    PIL = pytest.importorskip("PIL.Image")
    pytest.importorskip("cv2")
    from nvimagecodec_b200 import corpus

    rng = np.random.default_rng(7)
    for i in range(16):
        h, w = int(rng.integers(1, 260)), int(rng.integers(1, 260))
        samp = ["444", "422", "420", "440"][i % 4]
        data = corpus.encode(corpus.synth(h, w, 700 + i), int(rng.integers(25, 100)), samp, int(rng.integers(0, 5)) if i & 1 else 0,
                             gray=(i % 7 == 6), progressive=True)
        ref = np.asarray(PIL.open(io.BytesIO(data)).convert("RGB"))
        assert np.array_equal(oracle.decode(data, oracle.FMT_I_RGB, fancy=True), ref), i
    cut = data[:data.rfind(b"\xff\xda")] + b"\xff\xd9"
    with pytest.raises(oracle.OracleError) as ei:
        oracle.decode(cut)
    assert ei.value.code == 2


def test_live_against_reference_glue_when_built():
    """oracle/_ref/libref_turbo.so = the reference's libjpeg_turbo::Uncompress compiled from its own
    sources (oracle/build_ref.sh); it can switch fancy upsampling off and crop, which Pillow cannot."""
    from oracle import turbo

    if not turbo.available():
        pytest.skip("oracle/_ref/libref_turbo.so not built")
    pytest.importorskip("cv2")
    from nvimagecodec_b200 import corpus

    rng = np.random.default_rng(11)
    for i in range(10):
        h, w = int(rng.integers(9, 120)), int(rng.integers(9, 120))
        samp = ["444", "422", "420", "440", "411"][i % 5]
        data = corpus.encode(corpus.synth(h, w, 700 + i), int(rng.integers(30, 100)), samp, int(rng.integers(0, 5)), bool(i & 1))
        y0, x0 = int(rng.integers(0, h - 1)), int(rng.integers(0, w - 1))
        roi = (y0, x0, int(rng.integers(y0 + 1, h + 1)), int(rng.integers(x0 + 1, w + 1)))
        for fancy in (True, False):
            assert np.array_equal(oracle.decode(data, oracle.FMT_I_RGB, fancy), turbo.decode(data, "I_RGB", fancy))
            assert np.array_equal(oracle.decode(data, oracle.FMT_I_BGR, fancy), turbo.decode(data, "I_BGR", fancy))
            assert np.array_equal(oracle.decode(data, oracle.FMT_Y, fancy), turbo.decode(data, "I_Y", fancy))
            assert np.array_equal(oracle.decode(data, oracle.FMT_I_RGB, fancy, roi), turbo.decode(data, "I_RGB", fancy, roi))


def test_orientation_helper_matches_opencv_and_pillow():
    """oracle.apply_orientation is what the GPU path is compared with when EXIF orientation is applied."""
    cv2 = pytest.importorskip("cv2")
    from PIL import Image, ImageOps
    from nvimagecodec_b200 import corpus

    img = corpus.synth(40, 64, 5)
    for code in range(1, 9):
        ex = Image.Exif()
        ex[0x0112] = code
        buf = io.BytesIO()
        Image.fromarray(img).save(buf, "JPEG", quality=90, subsampling=2, exif=ex.tobytes())
        data = buf.getvalue()
        want = oracle.apply_orientation(oracle.decode(data, oracle.FMT_I_RGB, True), code)
        assert np.array_equal(want, cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)[..., ::-1])
        assert np.array_equal(want, np.asarray(ImageOps.exif_transpose(Image.open(io.BytesIO(data))).convert("RGB")))
        planar = oracle.apply_orientation(oracle.decode(data, oracle.FMT_P_RGB, True), code, planar=True)
        assert np.array_equal(planar, want.transpose(2, 0, 1))
