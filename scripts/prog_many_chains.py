"""GPU box: many more progressive chains than the GPU holds at once (restart intervals cut every scan into chains): the
launch must finish (a waiting chain's predecessor always has the lower CTA index) and the pixels must match the oracle."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import oracle  # noqa: E402
from nvimagecodec_b200 import Decoder, corpus  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 600
base = [corpus.encode(corpus.synth(96 + 8 * (k % 5), 128 + 16 * (k % 3), 7000 + k), 85, ("420", "444")[k % 2], rst_interval=1 + k % 3, progressive=True)
        for k in range(24)]
srcs = [base[i % len(base)] for i in range(n)]
d = Decoder(0, options=":fancy_upsampling=1")
t0 = time.perf_counter()
outs = d.decode(srcs, "I_RGB")
import torch  # noqa: E402

torch.cuda.synchronize()
dt = time.perf_counter() - t0
assert all(s == 0 for s in d.last_status)
refs = [oracle.decode(b) for b in base]
for i in range(0, n, 7):
    assert np.array_equal(outs[i].cpu().numpy(), refs[i % len(base)]), i
print("ok:", n, "images,", d.stat(0), "launches,", round(dt * 1e3, 1), "ms")
