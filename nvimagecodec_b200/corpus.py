"""Seeded synthetic JPEG corpora for the BASELINE.json configs.

Procedural RGB images are encoded with ``cv2.imencode`` -- libjpeg-turbo's encoder,
the same engine as the reference's CPU encoder (extensions/opencv/opencv_encoder.cpp:380-437)
-- so the bitstreams are ordinary baseline SOF0 JPEGs with standard or optimised
Huffman tables, optional restart markers and any of the usual chroma samplings.
Used by tests/ and bench.py only; the decoder itself never imports this module.
"""
import hashlib

import numpy as np

SAMPLINGS = {"444": 0x111111, "422": 0x211111, "420": 0x221111, "440": 0x121111, "411": 0x411111}


def synth(height, width, seed):
    """Procedural image: low-frequency sinusoid products + 32x32 blocks + noise."""
    rng = np.random.default_rng(seed)
    yy = np.arange(height, dtype=np.float32)[:, None]
    xx = np.arange(width, dtype=np.float32)[None, :]
    img = np.empty((height, width, 3), dtype=np.float32)
    bh, bw = (height + 31) // 32, (width + 31) // 32
    for c in range(3):
        acc = np.full((height, width), 127.0, dtype=np.float32)
        for _ in range(3):
            fy, fx = rng.uniform(0.005, 0.05, size=2)
            py, px = rng.uniform(0, 2 * np.pi, size=2)
            acc += (80.0 / 3.0) * np.sin(fy * yy + py) * np.sin(fx * xx + px) * 1.7
        blocks = rng.integers(0, 60, size=(bh, bw)).astype(np.float32)
        acc += np.kron(blocks, np.ones((32, 32), dtype=np.float32))[:height, :width] - 30.0
        acc += rng.normal(0.0, 6.0, size=(height, width)).astype(np.float32)
        img[..., c] = acc
    return np.clip(img, 0, 255).astype(np.uint8)


def encode(rgb, quality=90, sampling="420", rst_interval=0, optimize=False, gray=False, progressive=False):
    import cv2

    params = [cv2.IMWRITE_JPEG_QUALITY, int(quality)]
    if not gray:
        params += [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, SAMPLINGS[sampling]]
    if rst_interval:
        params += [cv2.IMWRITE_JPEG_RST_INTERVAL, int(rst_interval)]
    if optimize:
        params += [cv2.IMWRITE_JPEG_OPTIMIZE, 1]
    if progressive:  # libjpeg's jpeg_simple_progression script (ten scans for a colour image), optimised tables per scan
        params += [cv2.IMWRITE_JPEG_PROGRESSIVE, 1]
    src = cv2.cvtColor(rgb, cv2.COLOR_RGB2GRAY) if gray else rgb[..., ::-1]
    ok, buf = cv2.imencode(".jpg", np.ascontiguousarray(src), params)
    if not ok:
        raise RuntimeError("cv2.imencode failed")
    return buf.tobytes()


def config_c2(n=256, height=1080, width=1920):
    """C2: batch 256 x 1920x1080 baseline 4:2:0 q90, no RST."""
    return [encode(synth(height, width, s), 90, "420") for s in range(n)]


def config_c3(n=256, seed=1234):
    """C3: ImageNet-shaped variable-size batch, mixed samplings, with RST markers."""
    rng = np.random.default_rng(seed)
    out = []
    samplings = ["444", "422", "420"]
    for i in range(n):
        h = int(np.clip(rng.normal(375, 0.15 * 375), 64, 1024))
        w = int(np.clip(rng.normal(500, 0.15 * 500), 64, 1024))
        samp = samplings[i % 3]
        q = [75, 85, 95][(i // 3) % 3]
        hmax = 2 if samp in ("422", "420") else 1
        row = (w + 8 * hmax - 1) // (8 * hmax)
        rst = [row, 8, 32][(i // 9) % 3]
        out.append(encode(synth(h, w, seed * 100003 + i), q, samp, rst_interval=rst))
    return out


def config_c4(n=64, height=2160, width=3840):
    """C4: 3840x2160 4:4:4 q95, no restart markers."""
    return [encode(synth(height, width, 5000 + s), 95, "444") for s in range(n)]


def config_c5(n=8192, seed=77):
    """C5: mixed batch with ROI + grayscale outputs; returns (jpegs, requests).

    request = dict(fmt=..., roi=(y0,x0,y1,x1) or None)
    """
    rng = np.random.default_rng(seed)
    base = config_c3(min(n, 512), seed=seed)
    big = [encode(synth(1080, 1920, 9000 + s), 90, "420") for s in range(max(1, min(n, 512) // 10))]
    jpegs, reqs = [], []
    from . import jpeg_host

    for i in range(n):
        data = big[i % len(big)] if i % 10 == 9 else base[i % len(base)]
        inf = jpeg_host.probe(data)
        u = rng.uniform()
        fmt = "I_Y" if u < 0.3 else "I_RGB"
        roi = None
        if rng.uniform() < 0.5:
            fh = rng.uniform(0.25, 0.75)
            fw = rng.uniform(0.25, 0.75)
            rh, rw = max(1, int(inf.height * fh)), max(1, int(inf.width * fw))
            y0 = int(rng.integers(0, inf.height - rh + 1))
            x0 = int(rng.integers(0, inf.width - rw + 1))
            roi = (y0, x0, y0 + rh, x0 + rw)
        jpegs.append(data)
        reqs.append(dict(fmt=fmt, roi=roi))
    return jpegs, reqs


def manifest(jpegs):
    return [dict(bytes=len(j), sha256=hashlib.sha256(j).hexdigest()) for j in jpegs]
