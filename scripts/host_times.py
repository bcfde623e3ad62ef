"""Host-side cost of the three phases of a batch (256 images of the named config, pinned bitstreams, pixels left in HBM)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nvimagecodec_b200 import Decoder
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else 'c2'
jp = bench.make_workload(wl, 256, os.cpu_count())[0]
pinned = [torch.frombuffer(bytearray(j), dtype=torch.uint8).pin_memory() for j in jp]
srcs = [t.numpy() for t in pinned]
dec = Decoder(0, options=":fancy_upsampling=1")
for n in (256, 32, 1):
    b = dec.make_batch(srcs[:n], "I_RGB")
    acc = [0.0, 0.0, 0.0, 0.0]
    for it in range(8):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); dec.prepare(b); t1 = time.perf_counter()
        torch.cuda.synchronize(); t2 = time.perf_counter()
        dec.execute(); t3 = time.perf_counter()
        torch.cuda.synchronize(); t4 = time.perf_counter()
        dec.finish(b); t5 = time.perf_counter()
        if it >= 3:
            for i, v in enumerate((t1 - t0, t2 - t1, t3 - t2, t4 - t3 + 0 * t5)):
                acc[i] += v / 5
            acc3 = t5 - t4
    print("n=%d prepare host %.3f ms, H2D drain %.3f ms, execute launch %.3f ms, kernels drain %.3f ms, finish %.3f ms" % (n, acc[0] * 1e3, acc[1] * 1e3, acc[2] * 1e3, acc[3] * 1e3, acc3 * 1e3))
