import sys, time, io
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from nvimagecodec_b200 import Decoder, corpus, _lib
import oracle
dec = Decoder(0, options=":fancy_upsampling=1")
bad = 0; tot = 0
cases = []
for (h, w) in [(37, 53), (48, 64), (375, 500), (17, 9), (8, 8), (1, 1), (300, 3), (1080, 1920)]:
    img = corpus.synth(h, w, h * 1000 + w)
    for samp in ["444", "422", "420", "440", "411"]:
        for rst in [0, 1, 3]:
            cases.append((corpus.encode(img, 90, samp, rst, samp == "422"), (h, w, samp, rst)))
    cases.append((corpus.encode(img, 90, gray=True), (h, w, "gray", 0)))
srcs = [c[0] for c in cases]
t0 = time.time()
outs = dec.decode(srcs, "I_RGB")
torch.cuda.synchronize()
print("decode time", time.time() - t0, "statuses", set(dec.last_status), "resync", dec.stat(_lib.STAT_RESYNC_IMAGES))
for (data, tag), o, st in zip(cases, outs, dec.last_status):
    tot += 1
    ref = oracle.decode(data, oracle.FMT_I_RGB, True)
    if o is None:
        bad += 1; print("FAILED status", tag, st); continue
    got = o.cpu().numpy()
    d = int(np.abs(ref.astype(int) - got.astype(int)).max())
    if d:
        bad += 1
        # coefficient check
        inf = oracle.info(data)
        print("MISMATCH", tag, "maxdiff", d, "nbad", int((ref != got).sum()))
print("cases", tot, "bad", bad)
