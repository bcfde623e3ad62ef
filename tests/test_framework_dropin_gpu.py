"""Drop-in test: the UNMODIFIED nvImageCodec core (oracle/_ref/libnvimgcodec.so, compiled from the
reference's own sources by oracle/build_ref.sh) discovers libb200jpeg_ext.so in an extension
directory and decodes a batch through nvimgcodecDecoderDecode -> decodeBatch.  No other JPEG decoder
is loaded, so every pixel comes from the CUDA path.  Mirrors test/extensions/common_ext_decoder_test.h
(TestSingleImage) with the bar raised from eps=1 to bit-exact against the oracle."""
import os
import subprocess

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "oracle", "_ref", "host_driver")
EXT_DIR = os.path.join(ROOT, "nvimagecodec_b200", "lib")
OFMT = {"I_RGB": oracle.FMT_I_RGB, "I_BGR": oracle.FMT_I_BGR, "P_RGB": oracle.FMT_P_RGB, "P_BGR": oracle.FMT_P_BGR, "I_Y": oracle.FMT_Y}


def _run(tmp_path, entries, options, extra=()):
    man = tmp_path / "manifest.txt"
    lines = []
    for i, (data, fmt, roi, where) in enumerate(entries):
        p = tmp_path / f"in_{i}.jpg"
        p.write_bytes(data)
        y0, x0, y1, x1 = roi if roi else (0, 0, 0, 0)
        lines.append(f"{p} {fmt} {y0} {x0} {y1} {x1} {where}")
    man.write_text("\n".join(lines) + "\n")
    out = tmp_path / "out"
    out.mkdir()
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.path.dirname(DRIVER) + ":" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([DRIVER, EXT_DIR, str(man), str(out), options, *extra], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    status = [tuple(int(v) for v in l.split()) for l in (out / "status.txt").read_text().splitlines()]
    raws = [np.fromfile(out / f"{i}.raw", dtype=np.uint8) for i in range(len(entries))]
    return status, raws, r.stdout + r.stderr


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="reference core not built (oracle/build_ref.sh)")
@pytest.mark.parametrize("fancy", [1, 0])
def test_extension_is_picked_up_and_bit_exact(tmp_path, fancy):
    from nvimagecodec_b200 import corpus

    rng = np.random.default_rng(8)
    entries = []
    for i in range(20):
        h, w = int(rng.integers(24, 400)), int(rng.integers(24, 400))
        samp = ["420", "444", "422", "440", "411"][i % 5]
        data = corpus.encode(corpus.synth(h, w, 500 + i), int(rng.integers(50, 98)), samp, int(rng.integers(0, 12)), bool(i % 7 == 3))
        fmt = ["I_RGB", "P_RGB", "I_BGR", "I_Y", "P_BGR"][i % 5]
        roi = None
        if i % 3 == 1:
            y0, x0 = int(rng.integers(0, h - 8)), int(rng.integers(0, w - 8))
            roi = (y0, x0, int(rng.integers(y0 + 1, h + 1)), int(rng.integers(x0 + 1, w + 1)))
        entries.append((data, fmt, roi, "host" if i % 4 == 2 else "device"))
    entries.append((corpus.encode(corpus.synth(1080, 1920, 77), 90, "420"), "I_RGB", None, "device"))
    entries.append((corpus.encode(corpus.synth(90, 60, 78), 90, gray=True), "I_RGB", None, "device"))
    # one scan per component (non-interleaved): same coefficients as the interleaved original, which is what the oracle decodes
    from tests import remux
    orig = {}
    for k, (h, w, samp, rb) in enumerate([(120, 200, "420", 0), (75, 61, "444", 4), (96, 64, "422", 9)]):
        d = corpus.encode(corpus.synth(h, w, 560 + k), 90, samp)
        orig[len(entries)] = d
        entries.append((remux.to_three_scans(d, rb), ["I_RGB", "P_RGB", "I_Y"][k], (5, 7, h - 9, w - 3) if k == 1 else None, "device"))
    status, raws, log = _run(tmp_path, entries, f":fancy_upsampling={fancy}")
    assert "b200 jpeg decoder registered" in log
    for idx, ((data, fmt, roi, where), st, raw) in enumerate(zip(entries, status, raws)):
        assert st[0] == 1, (st, log[-2000:])  # NVIMGCODEC_PROCESSING_STATUS_SUCCESS
        ref = oracle.decode(orig.get(idx, data), OFMT[fmt], fancy=bool(fancy), roi=roi)
        assert raw.size == ref.size
        assert np.array_equal(raw, ref.reshape(-1))


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="reference core not built (oracle/build_ref.sh)")
def test_unsupported_streams_are_declined_not_crashed(tmp_path):
    """A stream the kernels do not take (here: progressive decoding switched off, and a 4-component stream) must be refused in
    canDecode (there is no fallback decoder loaded here, so the framework reports a non-success status) while the rest of
    the batch decodes."""
    cv2 = pytest.importorskip("cv2")
    import io
    from PIL import Image
    from nvimagecodec_b200 import corpus

    img = corpus.synth(64, 96, 4)
    ok, prog = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])
    buf = io.BytesIO()
    Image.fromarray(np.dstack([img, img[..., 0]])).convert("CMYK").save(buf, format="JPEG")
    good = corpus.encode(img, 90, "420")
    status, raws, log = _run(tmp_path, [(good, "I_RGB", None, "device"), (prog.tobytes(), "I_RGB", None, "device"),
                                        (good, "I_RGB", None, "device"), (buf.getvalue(), "I_RGB", None, "device")], ":progressive=0")
    assert status[0][0] == 1 and status[2][0] == 1
    assert status[1][0] != 1 and status[3][0] != 1
    ref = oracle.decode(good, oracle.FMT_I_RGB, fancy=True)
    assert np.array_equal(raws[0], ref.reshape(-1)) and np.array_equal(raws[2], ref.reshape(-1))


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="reference core not built (oracle/build_ref.sh)")
def test_progressive_streams_through_the_framework(tmp_path):
    """SOF2 streams are taken by the extension (the reference's parser reports NVIMGCODEC_JPEG_ENCODING_PROGRESSIVE_DCT_HUFFMAN)
    and decode bit-exactly next to baseline ones; test/extensions/libjpeg_turbo_ext_decoder_test.cpp:115-130 is the
    reference's own progressive case."""
    from nvimagecodec_b200 import corpus

    entries = []
    for k, (h, w, samp, rst) in enumerate([(213, 200, "420", 0), (96, 128, "444", 3), (301, 77, "422", 0)]):
        entries.append((corpus.encode(corpus.synth(h, w, 880 + k), 88, samp, rst_interval=rst, progressive=True),
                        ["I_RGB", "I_BGR", "I_Y"][k], (11, 9, h - 20, w - 13) if k == 0 else None, ["device", "host", "device"][k]))
    entries.append((corpus.encode(corpus.synth(90, 60, 78), 90, gray=True, progressive=True), "I_RGB", None, "device"))
    entries.append((corpus.encode(corpus.synth(120, 160, 79), 90, "420"), "I_RGB", None, "device"))
    status, raws, log = _run(tmp_path, entries, ":fancy_upsampling=1")
    for (data, fmt, roi, where), st, raw in zip(entries, status, raws):
        assert st[0] == 1, (st, log[-2000:])
        ref = oracle.decode(data, OFMT[fmt], fancy=True, roi=roi)
        assert np.array_equal(raw, ref.reshape(-1))


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="reference core not built (oracle/build_ref.sh)")
def test_exif_orientation_through_the_framework(tmp_path):
    """apply_exif_orientation = 1: the reference's parser reads the EXIF tag (src/parsers/jpeg.cpp + exif.cpp), the framework
    hands the orientation to the decoder, the extension applies it while storing (libjpeg_turbo_ext would decline it)."""
    import io

    from PIL import Image
    from nvimagecodec_b200 import corpus

    entries, codes = [], []
    for code in range(1, 9):
        img = corpus.synth(45 + code, 80 - code, 400 + code)
        ex = Image.Exif()
        ex[0x0112] = code
        buf = io.BytesIO()
        Image.fromarray(img).save(buf, "JPEG", quality=88, subsampling=[0, 1, 2][code % 3], exif=ex.tobytes())
        entries.append((buf.getvalue(), ["I_RGB", "P_RGB", "I_Y"][code % 3], None, "device" if code % 2 else "host"))
        codes.append(code)
    status, raws, log = _run(tmp_path, entries, ":fancy_upsampling=1", extra=("0", "any", "exif"))
    for (data, fmt, _, _), code, st, raw in zip(entries, codes, status, raws):
        assert st[0] == 1, (code, st, log[-1500:])
        ref = oracle.apply_orientation(oracle.decode(data, OFMT[fmt], fancy=True), code, planar=fmt.startswith("P_"))
        assert raw.size == ref.size and np.array_equal(raw, ref.reshape(-1)), code


NVIMTRANS = os.path.join(ROOT, "oracle", "_ref", "nvimtrans")
NVBMP = os.path.join(ROOT, "oracle", "_ref", "extensions", "libnvbmp_ext.so")


def _read_bmp(path):
    """24-bit uncompressed BMP -> HxWx3 RGB (what the reference's nvbmp encoder writes)."""
    raw = open(path, "rb").read()
    assert raw[:2] == b"BM"
    off = int.from_bytes(raw[10:14], "little")
    w = int.from_bytes(raw[18:22], "little", signed=True)
    h = int.from_bytes(raw[22:26], "little", signed=True)
    bpp = int.from_bytes(raw[28:30], "little")
    assert bpp == 24
    stride = (w * 3 + 3) // 4 * 4
    rows = np.frombuffer(raw, dtype=np.uint8, count=stride * abs(h), offset=off).reshape(abs(h), stride)[:, :w * 3].reshape(abs(h), w, 3)
    if h > 0:
        rows = rows[::-1]
    return rows[..., ::-1]


@pytest.mark.skipif(not (os.path.exists(NVIMTRANS) and os.path.exists(NVBMP)), reason="nvimtrans not built (oracle/build_ref.sh)")
def test_nvimtrans_sample_uses_the_extension(tmp_path):
    """the reference's own CLI sample, unmodified, transcodes a directory of JPEGs
    to BMP.  The only JPEG decoder in its extension directory is ours (found by codec priority through
    NVIMGCODEC_EXTENSIONS_PATH); nvimtrans asks for planar RGB with ":fancy_upsampling=0"
    (example/nvimtrans/main.cpp:275, 516)."""
    import shutil

    from nvimagecodec_b200 import corpus

    ext = tmp_path / "extensions"
    ext.mkdir()
    shutil.copy(os.path.join(EXT_DIR, "libb200jpeg_ext.so"), ext / "libb200jpeg_ext.so")
    shutil.copy(NVBMP, ext / "libnvbmp_ext.so")
    src, dst = tmp_path / "in", tmp_path / "out"
    src.mkdir()
    dst.mkdir()
    datas = {}
    for i, (h, w, samp, rst) in enumerate([(240, 320, "420", 0), (199, 301, "444", 5), (128, 96, "422", 2), (375, 500, "420", 16)]):
        datas[f"img{i}"] = corpus.encode(corpus.synth(h, w, 640 + i), 88, samp, rst)
        (src / f"img{i}.jpg").write_bytes(datas[f"img{i}"])
    env = dict(os.environ)
    env["NVIMGCODEC_EXTENSIONS_PATH"] = str(ext)
    env["LD_LIBRARY_PATH"] = os.path.dirname(NVIMTRANS) + ":" + env.get("LD_LIBRARY_PATH", "")
    # One file per invocation: the sample's directory mode cannot work against this revision of the core
    # (example/nvimtrans/main.cpp:689 passes a NULL status array, src/nvimgcodec_capi.cpp:950 rejects NULL),
    # whichever decoder is installed.
    for name, data in datas.items():
        r = subprocess.run([NVIMTRANS, "-i", str(src / (name + ".jpg")), "-o", str(dst / (name + ".bmp"))],
                           capture_output=True, text=True, env=env, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        got = _read_bmp(dst / (name + ".bmp"))
        ref = oracle.decode(data, oracle.FMT_I_RGB, fancy=False)
        assert np.array_equal(got, ref), name


PYMOD = os.path.join(ROOT, "oracle", "_ref", "python")

_PY_SCRIPT = r'''
import sys, glob, numpy as np
from nvidia import nvimgcodec
files = sorted(glob.glob(sys.argv[1] + "/*.jpg"))
dec = nvimgcodec.Decoder()                      # binding default options: ":fancy_upsampling=0"
imgs = dec.decode([open(f, "rb").read() for f in files])
for f, im in zip(files, imgs):
    assert im is not None, f
    assert im.buffer_kind == nvimgcodec.ImageBufferKind.STRIDED_DEVICE
    np.save(f[:-4] + "_rgb.npy", np.asarray(im.cpu()))
gray = dec.decode([open(f, "rb").read() for f in files], params=nvimgcodec.DecodeParams(color_spec=nvimgcodec.ColorSpec.GRAY))
for f, im in zip(files, gray):
    np.save(f[:-4] + "_gray.npy", np.asarray(im.cpu()))
print("decoded", len(files), "with nvidia.nvimgcodec", nvimgcodec.__version__)
'''


@pytest.mark.skipif(not os.path.isdir(os.path.join(PYMOD, "nvidia", "nvimgcodec")), reason="reference python binding not built")
def test_reference_python_binding_picks_the_extension(tmp_path):
    """SURVEY.md 8(f) n1, second half: `from nvidia import nvimgcodec; nvimgcodec.Decoder().decode(...)` -- the reference's own
    pybind11 module (python/*.cpp, unmodified) over the unmodified core.  The extension directory holds only our decoder,
    so these pixels come from the CUDA path; the binding asks for I_RGB / I_Y with ":fancy_upsampling=0"."""
    import shutil
    import sys

    from nvimagecodec_b200 import corpus

    ext = tmp_path / "extensions"
    ext.mkdir()
    shutil.copy(os.path.join(EXT_DIR, "libb200jpeg_ext.so"), ext / "libb200jpeg_ext.so")
    datas = {}
    for i, (h, w, samp, rst) in enumerate([(240, 320, "420", 0), (199, 301, "444", 5), (128, 96, "422", 2), (1080, 1920, "420", 0)]):
        datas[f"img{i}"] = corpus.encode(corpus.synth(h, w, 740 + i), 88, samp, rst)
        (tmp_path / f"img{i}.jpg").write_bytes(datas[f"img{i}"])
    (tmp_path / "run.py").write_text(_PY_SCRIPT)
    env = dict(os.environ)
    env["NVIMGCODEC_EXTENSIONS_PATH"] = str(ext)
    env["PYTHONPATH"] = PYMOD + ":" + env.get("PYTHONPATH", "")
    env["LD_LIBRARY_PATH"] = os.path.join(ROOT, "oracle", "_ref") + ":" + env.get("LD_LIBRARY_PATH", "")
    r = subprocess.run([sys.executable, str(tmp_path / "run.py"), str(tmp_path)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    for name, data in datas.items():
        rgb = np.load(tmp_path / (name + "_rgb.npy"))
        assert np.array_equal(rgb, oracle.decode(data, oracle.FMT_I_RGB, fancy=False)), name
        gray = np.load(tmp_path / (name + "_gray.npy"))
        assert np.array_equal(gray.reshape(gray.shape[0], gray.shape[1]), oracle.decode(data, oracle.FMT_Y, fancy=False)), name


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="reference core not built (oracle/build_ref.sh)")
def test_out_of_bounds_region_is_filled(tmp_path):
    """Regions that reach outside the image (nvimgcodecRegion_t with the constant out-of-bounds policy): what nvjpeg_ext accepts
    and fills (extensions/nvjpeg/cuda_decoder.cpp:763-766, src/imgproc/out_of_bound_roi_fill.cpp) and libjpeg_turbo_ext declines.
    The part inside the image must equal the crop of the full decode, everything else the fill sample."""
    from nvimagecodec_b200 import corpus

    cases = [(120, 160, "420", "I_RGB", (-7, -11, 100, 130)), (97, 131, "444", "P_RGB", (10, 20, 140, 150)),
             (64, 200, "422", "I_Y", (-5, 30, 70, 260)), (80, 80, "420", "I_BGR", (-16, -16, 96, 96)),
             (75, 61, "444", "P_BGR", (5, -9, 60, 40))]
    entries = []
    for k, (h, w, samp, fmt, roi) in enumerate(cases):
        entries.append((corpus.encode(corpus.synth(h, w, 880 + k), 88, samp, rst_interval=k), fmt, roi, "device"))
    old = os.environ.get("HOST_DRIVER_FILL")
    os.environ["HOST_DRIVER_FILL"] = "77"
    try:
        status, raws, log = _run(tmp_path, entries, ":fancy_upsampling=1")
    finally:
        if old is None:
            os.environ.pop("HOST_DRIVER_FILL")
        else:
            os.environ["HOST_DRIVER_FILL"] = old
    for (data, fmt, roi, _), (h, w, _, _, _), st, raw in zip(entries, cases, status, raws):
        assert st[0] == 1, (st, log[-1500:])
        full = oracle.decode(data, OFMT[fmt], fancy=True)
        y0, x0, y1, x1 = roi
        planar = fmt.startswith("P_")
        shape = (3, y1 - y0, x1 - x0) if planar else ((y1 - y0, x1 - x0) if fmt == "I_Y" else (y1 - y0, x1 - x0, 3))
        want = np.full(shape, 77, np.uint8)
        cy0, cx0, cy1, cx1 = max(y0, 0), max(x0, 0), min(y1, h), min(x1, w)
        if planar:
            want[:, cy0 - y0:cy1 - y0, cx0 - x0:cx1 - x0] = full[:, cy0:cy1, cx0:cx1]
        else:
            want[cy0 - y0:cy1 - y0, cx0 - x0:cx1 - x0] = full[cy0:cy1, cx0:cx1]
        assert raw.size == want.size and np.array_equal(raw, want.reshape(-1)), (fmt, roi)
