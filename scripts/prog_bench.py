"""GPU box: progressive (SOF2) decode timing -- kernels only (bitstreams resident), batches of 1080p 4:2:0 q90 and of
ImageNet-shaped images.  Usage: python scripts/prog_bench.py [n_1080p]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch

from nvimagecodec_b200 import Decoder, corpus


def timed(dec, srcs, reps=5):
    pinned = [torch.frombuffer(bytearray(j), dtype=torch.uint8).pin_memory().numpy() for j in srcs]
    batch = dec.make_batch(pinned, "I_RGB", None, to_host=False)
    dec.prepare(batch)
    for _ in range(2):
        dec.execute()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        dec.execute()
    e1.record()
    torch.cuda.synchronize()
    st = dec.finish(batch)
    assert all(s == 0 for s in st), st
    mpix = sum(i.width * i.height for i in batch.infos) / 1e6
    ms = e0.elapsed_time(e1) / reps
    t0 = time.perf_counter()
    st = dec.run(batch)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    return {"images": len(srcs), "ms_kernels": round(ms, 3), "gpix_per_s": round(mpix / ms, 2), "ms_call_wall": round(wall, 3), "compressed_mb": round(sum(len(s) for s in srcs) / 1e6, 2)}


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    dec = Decoder(0, options=":fancy_upsampling=1")
    base = [corpus.synth(1080, 1920, s) for s in range(min(n, 8))]
    prog = [corpus.encode(base[i % len(base)], 90, "420", progressive=True) for i in range(n)]
    seq = [corpus.encode(base[i % len(base)], 90, "420") for i in range(n)]
    out = {}
    for k in (1, 8, n):
        out["1080p_progressive_x%d" % k] = timed(dec, prog[:k])
    out["1080p_baseline_x%d" % n] = timed(dec, seq)
    rng = np.random.default_rng(3)
    small = []
    for i in range(256):
        h, w = int(375 * rng.uniform(0.85, 1.15)), int(500 * rng.uniform(0.85, 1.15))
        small.append(corpus.encode(corpus.synth(h, w, 1000 + i), int(rng.integers(75, 96)), ("420", "444", "422")[i % 3], progressive=True))
    out["imagenet_like_progressive_x256"] = timed(dec, small)
    # the reference's CPU path on the same progressive files (one thread), when the glue is built
    try:
        from oracle import turbo

        t0 = time.perf_counter()
        for s in prog[:4]:
            turbo.decode(s, "I_RGB", fancy=True)
        out["cpu_reference_1080p_progressive_ms_per_image_1_thread"] = round((time.perf_counter() - t0) * 1e3 / 4, 2)
    except Exception as e:  # noqa: BLE001
        out["cpu_reference"] = "unavailable: %s" % e
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
