import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nvimagecodec_b200 import Decoder
import bench
jp = bench.make_corpus(256, os.cpu_count())
pinned = [torch.frombuffer(bytearray(j), dtype=torch.uint8).pin_memory() for j in jp]
srcs = [t.numpy() for t in pinned]
mpix = 256 * 1920 * 1080 / 1e6
for mb, im, ln in [(24, 64, 3), (24, 64, 4), (24, 64, 6), (16, 32, 4), (16, 32, 6), (16, 32, 8), (32, 64, 6), (8, 16, 8)]:
    dec = Decoder(0, options=f":fancy_upsampling=1 b200_jpeg_decoder:chunk_mb={mb} b200_jpeg_decoder:chunk_images={im} b200_jpeg_decoder:lanes={ln}")
    for to_host in (False, True):
        b = dec.make_batch(srcs, "I_RGB", to_host=to_host)
        for _ in range(3): dec.run(b)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(6): dec.run(b)
        torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 6
        print("lanes", ln, "chunk_mb", mb, "chunk_images", im, "host_out" if to_host else "dev_out", "%.2f ms" % (dt * 1e3), "%.0f MPix/s" % (mpix / dt))
    dec.close()
