"""One process per GPU (torchrun): b200jpegDecodeBatch, pinned bitstreams in, pinned pixels out, per-rank time per batch."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from nvimagecodec_b200 import Decoder

rank, local = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
jp = bench.make_corpus(256, 8)
torch.cuda.set_device(local)
dec = Decoder(local, options=":fancy_upsampling=1")
pinned = [torch.frombuffer(bytearray(j), dtype=torch.uint8).pin_memory() for j in jp]
srcs = [t.numpy() for t in pinned]
for to_host in (True, False):
    b = dec.make_batch(srcs, "I_RGB", to_host=to_host)
    for _ in range(3):
        dec.run(b)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        dec.run(b)
    torch.cuda.synchronize()
    print("rank", rank, "to_host", to_host, "ms/batch %.2f" % ((time.perf_counter() - t0) / 5 * 1e3), flush=True)
