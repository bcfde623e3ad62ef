import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nvimagecodec_b200 import Decoder, corpus
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
jp = bench.make_corpus(n, os.cpu_count())
dec = Decoder(0, options=":fancy_upsampling=1")
b = dec.make_batch(jp, "I_RGB")
dec.prepare(b); dec.execute(); torch.cuda.synchronize()
ct, rr, rep, mx = [dec.stat(i) for i in (8, 9, 10, 11)]
print("C2: CTAs %d repair rounds/CTA %.2f repaired subseq/CTA %.2f (of 248) max rounds %d" % (ct, rr / ct, rep / ct, mx))
dec.finish(b)
j4 = [corpus.encode(corpus.synth(2160, 3840, 5000 + s), 95, "444") for s in range(4)]
b = dec.make_batch(j4, "I_RGB")
dec.prepare(b); dec.execute(); torch.cuda.synchronize()
ct, rr, rep, mx = [dec.stat(i) for i in (8, 9, 10, 11)]
print("C4: CTAs %d repair rounds/CTA %.2f repaired subseq/CTA %.2f max rounds %d" % (ct, rr / ct, rep / ct, mx))
print(dec.finish(b), "resync", dec.stat(5))
