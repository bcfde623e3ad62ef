import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from nvimagecodec_b200 import Decoder, corpus
from oracle import turbo
dec = Decoder(0, options=":fancy_upsampling=1")
for name, (h, w, q, samp, rst) in {"1080p 420 q90": (1080, 1920, 90, "420", 0), "500x375 420 q85 rst8": (375, 500, 85, "420", 8), "4K 444 q95": (2160, 3840, 95, "444", 0)}.items():
    d = corpus.encode(corpus.synth(h, w, 3), q, samp, rst)
    pin = torch.frombuffer(bytearray(d), dtype=torch.uint8).pin_memory().numpy()
    b = dec.make_batch([pin], "I_RGB")
    for _ in range(5): dec.run(b)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(50): dec.run(b)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 50
    t1 = time.perf_counter(); 
    for _ in range(5): turbo.decode(d, "I_RGB", True)
    cpu = (time.perf_counter() - t1) / 5
    print("%-22s batch=1 GPU %.3f ms  (CPU reference path, 1 thread: %.2f ms)" % (name, dt * 1e3, cpu * 1e3))
