"""GPU box: damaged progressive streams (bytes flipped anywhere incl. scan headers and tables, truncation, dropped slices) in
batches next to pristine ones: no crash, no hang, the pristine neighbours stay bit-exact, the decoder stays usable."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import oracle  # noqa: E402
from nvimagecodec_b200 import Decoder, corpus  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
rng = np.random.default_rng(99)
seeds = [corpus.encode(corpus.synth(72 + 16 * k, 88 + 24 * k, 60 + k), 82, ("420", "444", "422")[k % 3], rst_interval=(0, 2, 5)[k % 3], progressive=True)
         for k in range(6)]
good = seeds[0]
ref = oracle.decode(good)
d = Decoder(0, options=":fancy_upsampling=1")
counts = {}
for b0 in range(0, n, 100):
    srcs = [good]
    for it in range(b0, b0 + 100):
        b = bytearray(seeds[it % len(seeds)])
        kind = it % 4
        if kind == 0:
            for _ in range(1 + it % 5):
                b[int(rng.integers(2, len(b)))] = int(rng.integers(0, 256))
        elif kind == 1:
            marks = [i for i in range(len(b) - 1) if b[i] == 0xFF and b[i + 1] in (0xC2, 0xC4, 0xDA, 0xDB, 0xDD)]
            m = marks[int(rng.integers(0, len(marks)))]
            b[min(len(b) - 1, m + int(rng.integers(2, 24)))] = int(rng.integers(0, 256))
        elif kind == 2:
            b = b[:int(rng.integers(4, len(b)))]
        else:
            a = int(rng.integers(2, len(b) - 2))
            del b[a:a + int(rng.integers(1, 64))]
        srcs.append(bytes(b))
    srcs.append(good)
    outs = d.decode(srcs, "I_RGB")
    for s in d.last_status:
        counts[s] = counts.get(s, 0) + 1
    assert d.last_status[0] == 0 and d.last_status[-1] == 0
    assert np.array_equal(outs[0].cpu().numpy(), ref) and np.array_equal(outs[-1].cpu().numpy(), ref)
print("ok", n, "damaged streams; statuses:", counts)
