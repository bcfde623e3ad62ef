#!/bin/bash
# GPU box: per-kernel times on C2 for each subsequence length (and any extra decoder options in OPTS)
cd "$(dirname "$0")/.."
for sb in 512 1024 2048 4096; do
python - <<PY
import sys,torch
sys.path.insert(0,'.')
import bench
from nvimagecodec_b200 import Decoder
jp=bench.make_corpus(256,16)
dec=Decoder(0,options=":fancy_upsampling=1 b200_jpeg_decoder:subseq_bits=$sb${OPTS}")
b=dec.make_batch(jp,"I_RGB")
dec.set_profiling(True)
dec.prepare(b)
for i in range(3): dec.execute()
torch.cuda.synchronize()
acc={}
for i in range(10):
    dec.execute(); torch.cuda.synchronize()
    for k,v in dec.stage_times().items(): acc[k]=acc.get(k,0)+v/10
print("subseq=$sb total=%.3f"%sum(acc.values()), {k:round(v,3) for k,v in acc.items() if v>0.01})
PY
done
