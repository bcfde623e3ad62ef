"""Small mixed batch for compute-sanitizer: every kernel, both upsampling modes, ROI, formats, RST, fix-up path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import oracle
from nvimagecodec_b200 import Decoder, corpus, _lib
srcs, fmts, rois = [], [], []
rng = np.random.default_rng(0)
for i, (h, w, samp, rst) in enumerate([(97, 131, "420", 0), (64, 200, "444", 3), (33, 57, "422", 1), (120, 90, "440", 0), (75, 75, "411", 2), (300, 500, "420", 0), (1, 1, "420", 0), (17, 9, "444", 0)]):
    srcs.append(corpus.encode(corpus.synth(h, w, i), 85, samp, rst, optimize=bool(i & 1)))
    fmts.append(["I_RGB", "P_BGR", "I_Y", "I_YUV", "P_RGB", "I_BGR", "I_RGB", "P_YUV"][i])
    rois.append((h // 4, w // 5, h - h // 6, w - w // 7) if i in (0, 5) else None)
srcs.append(corpus.encode(corpus.synth(50, 70, 9), 90, gray=True)); fmts.append("I_RGB"); rois.append(None)
bad = 0
for opts, fancy in ((":fancy_upsampling=1", True), (":fancy_upsampling=0 b200_jpeg_decoder:debug_no_warm=1 b200_jpeg_decoder:subseq_bits=1024", False)):
    dec = Decoder(0, options=opts)
    outs = dec.decode(srcs, fmts, rois)
    torch.cuda.synchronize()
    omap = {"I_RGB": 0, "I_BGR": 1, "P_RGB": 2, "P_BGR": 3, "I_Y": 4, "I_YUV": 5, "P_YUV": 6}
    for d, f, r, o, st in zip(srcs, fmts, rois, outs, dec.last_status):
        if st != _lib.OK or not np.array_equal(o.cpu().numpy(), oracle.decode(d, omap[f], fancy, r)):
            bad += 1
    dec.close()
print("sanitize case: mismatches", bad)
sys.exit(1 if bad else 0)
