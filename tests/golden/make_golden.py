"""Generates tests/golden/: small JPEG fixtures plus the pixels libjpeg-turbo produces for them.

Run HERE (the container with /root/reference, PIL and cv2), never on the GPU box:
    python tests/golden/make_golden.py

Expected pixels come from the libjpeg-turbo binaries in the image -- Pillow (3.1.4.1) and
OpenCV (3.1.2), both ISLOW IDCT + fancy upsampling, i.e. the libjpeg_turbo_ext defaults
(extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:223-249).  The two must agree before a
fixture is written.  Set A re-encodes the reference's own decoded goldens
(resources/ref/jpeg/*.ppm; their .jpg inputs are git-LFS stubs) as the BASELINE config #1
substitute; set B is small synthetic edge cases.
"""
import hashlib
import io
import json
import os
import sys

import cv2
import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from nvimagecodec_b200 import corpus  # noqa: E402
from oracle import turbo  # noqa: E402  (the reference's own jpeg_mem.cpp over Pillow's libjpeg-turbo; oracle/build_ref.sh)

REF = "/root/reference/resources/ref/jpeg"


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def turbo_rgb(data):
    a = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
    b = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)[..., ::-1]
    assert np.array_equal(a, b), "Pillow and OpenCV libjpeg-turbo builds disagree"
    return a


def turbo_gray(data):
    # out_color_space = JCS_GRAYSCALE: the Y plane, chroma never decoded
    im = Image.open(io.BytesIO(data))
    im.draft("L", im.size)
    a = np.asarray(im.convert("L")) if im.mode != "L" else np.asarray(im)
    b = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_GRAYSCALE)
    assert np.array_equal(a, b), "Pillow and OpenCV grayscale decodes disagree"
    return a


def main():
    entries = []
    os.makedirs(os.path.join(HERE, "jpeg"), exist_ok=True)

    def add(name, data, note):
        rgb = turbo_rgb(data)
        gray = turbo_gray(data)
        with open(os.path.join(HERE, "jpeg", name + ".jpg"), "wb") as f:
            f.write(data)
        # the reference's libjpeg_turbo::Uncompress, both upsampling modes (fancy must equal Pillow/OpenCV)
        ref_fancy = turbo.decode(data, "I_RGB", fancy=True)
        assert np.array_equal(ref_fancy, rgb), "reference glue disagrees with Pillow"
        ref_box = turbo.decode(data, "I_RGB", fancy=False)
        h_, w_ = rgb.shape[:2]
        roi = (h_ // 3, w_ // 4, max(h_ // 3 + 1, h_ - h_ // 5), max(w_ // 4 + 1, w_ - w_ // 7))
        ref_roi = turbo.decode(data, "I_RGB", fancy=False, roi=roi)
        e = dict(name=name, note=note, bytes=len(data), height=int(rgb.shape[0]), width=int(rgb.shape[1]),
                 sha256_rgb_fancy=sha(rgb), sha256_y=sha(gray), sha256_rgb_box=sha(ref_box), roi=list(roi), sha256_roi_box=sha(ref_roi))
        if rgb.size <= 6000:
            e["rgb_fancy"] = rgb.reshape(-1).tolist()
        entries.append(e)

    # ---- set A: config #1 substitute (reference goldens re-encoded, 4:2:0 q90 + the other samplings)
    for stem in ["padlock-406986_640_410", "padlock-406986_640_411", "padlock-406986_640_420", "padlock-406986_640_422",
                 "padlock-406986_640_440", "padlock-406986_640_444"]:
        rgb = np.asarray(Image.open(os.path.join(REF, stem + ".ppm")).convert("RGB"))
        add("c1_" + stem + "_as420", corpus.encode(rgb, 90, "420"), "resources/ref/jpeg/%s.ppm re-encoded 4:2:0 q90" % stem)
    rgb = np.asarray(Image.open(os.path.join(REF, "padlock-406986_640_444.ppm")).convert("RGB"))
    for samp in ["444", "422", "440", "411"]:
        add("c1_padlock_%s" % samp, corpus.encode(rgb, 90, samp), "padlock 444 golden re-encoded %s q90" % samp)
    add("c1_padlock_420_rst", corpus.encode(rgb, 90, "420", rst_interval=40), "padlock re-encoded 4:2:0 q90, RST every 40 MCUs")
    add("c1_padlock_420_opt", corpus.encode(rgb, 75, "420", optimize=True), "padlock re-encoded 4:2:0 q75, optimised Huffman tables")
    add("c1_padlock_gray", corpus.encode(rgb, 90, gray=True), "padlock re-encoded as grayscale q90")

    # ---- set P: progressive (SOF2) streams.  The reference pins progressive decode with
    # test/extensions/libjpeg_turbo_ext_decoder_test.cpp:115-130 on progressive-subsampled-imagenet-n02089973_1957.jpg (a git-LFS
    # stub here); its decoded golden is re-encoded as a progressive stream, like set A.
    prog = np.asarray(Image.open(os.path.join(REF, "progressive-subsampled-imagenet-n02089973_1957.ppm")).convert("RGB"))
    add("p_imagenet_1957_420", corpus.encode(prog, 90, "420", progressive=True), "resources/ref/jpeg/progressive-...1957.ppm re-encoded progressive 4:2:0 q90")
    add("p_imagenet_1957_444_rst", corpus.encode(prog, 80, "444", rst_interval=5, progressive=True), "the same, progressive 4:4:4 q80, RST every 5 MCUs")
    add("p_padlock_422", corpus.encode(rgb, 85, "422", progressive=True), "padlock re-encoded progressive 4:2:2 q85")
    add("p_padlock_gray", corpus.encode(rgb, 90, gray=True, progressive=True), "padlock re-encoded progressive grayscale q90")
    for (h, w) in [(8, 8), (17, 9), (37, 53)]:
        img = corpus.synth(h, w, 31 * h + w)
        for samp in ["444", "420", "440"]:
            add("p_syn_%dx%d_%s" % (h, w, samp), corpus.encode(img, 75, samp, rst_interval=(2 if samp == "440" else 0), progressive=True),
                "synthetic progressive")

    # ---- set B: synthetic edge cases
    for (h, w) in [(1, 1), (8, 8), (17, 9), (37, 53), (48, 64), (2, 300), (300, 3)]:
        img = corpus.synth(h, w, 31 * h + w)
        for samp in ["444", "422", "420", "440", "411"]:
            for rst in [0, 1, 3]:
                add("syn_%dx%d_%s_r%d" % (h, w, samp, rst), corpus.encode(img, 90 if rst != 3 else 50, samp, rst, optimize=(rst == 3)),
                    "synthetic")
        add("syn_%dx%d_gray" % (h, w), corpus.encode(img, 90, gray=True), "synthetic grayscale")
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(dict(generator="tests/golden/make_golden.py",
                       engines="Pillow libjpeg-turbo 3.1.4.1 == OpenCV libjpeg-turbo 3.1.2 (ISLOW, fancy upsampling); "
                               "fancy_upsampling=0 and ROI vectors from the reference's libjpeg_turbo::Uncompress "
                               "(extensions/libjpeg_turbo/jpeg_mem.cpp) over the same Pillow binary",
                       entries=entries), f, indent=0)
    print(len(entries), "fixtures,", sum(e["bytes"] for e in entries), "jpeg bytes")


if __name__ == "__main__":
    main()
