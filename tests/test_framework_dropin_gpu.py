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


def _run(tmp_path, entries, options):
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
    r = subprocess.run([DRIVER, EXT_DIR, str(man), str(out), options], capture_output=True, text=True, env=env, timeout=300)
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
    status, raws, log = _run(tmp_path, entries, f":fancy_upsampling={fancy}")
    assert "b200 jpeg decoder registered" in log
    for (data, fmt, roi, where), st, raw in zip(entries, status, raws):
        assert st[0] == 1, (st, log[-2000:])  # NVIMGCODEC_PROCESSING_STATUS_SUCCESS
        ref = oracle.decode(data, OFMT[fmt], fancy=bool(fancy), roi=roi)
        assert raw.size == ref.size
        assert np.array_equal(raw, ref.reshape(-1))


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="reference core not built (oracle/build_ref.sh)")
def test_unsupported_streams_are_declined_not_crashed(tmp_path):
    """A progressive stream must be refused in canDecode (there is no fallback decoder loaded here,
    so the framework reports a non-success status) while the rest of the batch decodes."""
    cv2 = pytest.importorskip("cv2")
    from nvimagecodec_b200 import corpus

    img = corpus.synth(64, 96, 4)
    ok, prog = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])
    good = corpus.encode(img, 90, "420")
    status, raws, log = _run(tmp_path, [(good, "I_RGB", None, "device"), (prog.tobytes(), "I_RGB", None, "device"),
                                        (good, "I_RGB", None, "device")], "")
    assert status[0][0] == 1 and status[2][0] == 1
    assert status[1][0] != 1
    ref = oracle.decode(good, oracle.FMT_I_RGB, fancy=True)
    assert np.array_equal(raws[0], ref.reshape(-1)) and np.array_equal(raws[2], ref.reshape(-1))
