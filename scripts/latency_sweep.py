"""Single-image latency per subsequence length (0 = the decoder's own choice)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nvimagecodec_b200 import Decoder, corpus
cases = {"1080p 420 q90": (1080, 1920, 90, "420", 0), "500x375 420 q85 rst8": (375, 500, 85, "420", 8), "4K 444 q95": (2160, 3840, 95, "444", 0),
         "8K x 4K 420 q90": (4320, 7680, 90, "420", 0)}
data = {k: corpus.encode(corpus.synth(h, w, 3), q, s, r) for k, (h, w, q, s, r) in cases.items()}
for sb in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0,512,1024,2048").split(",")]:
    dec = Decoder(0, options=":fancy_upsampling=1" + (" :subseq_bits=%d" % sb if sb else ""))
    for name, d in data.items():
        pin = torch.frombuffer(bytearray(d), dtype=torch.uint8).pin_memory().numpy()
        b = dec.make_batch([pin], "I_RGB")
        for _ in range(5): dec.run(b)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(30): dec.run(b)
        torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 30
        print("subseq_bits=%-4d %-22s %.3f ms  resync=%d" % (sb, name, dt * 1e3, dec.stat(5)))
    dec.close()
