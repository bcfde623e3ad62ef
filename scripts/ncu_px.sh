#!/bin/bash
# GPU box: one ncu capture of the fused pixel kernel per B200JPEG_PX_DEBUG value given (0 = whole kernel, 2 = phase A only, 4 = phase B only)
cd "$(dirname "$0")/.."
TAG=${TAG:-px}
for d in "$@"; do
B200JPEG_PX_DEBUG=$d ncu --set full --clock-control none --import-source on -k regex:pixels -c 1 -f -o gpurun_out/${TAG}_d$d python - > gpurun_out/${TAG}_d$d.log 2>&1 <<PY
import sys,torch
sys.path.insert(0,'.')
import bench
from nvimagecodec_b200 import Decoder
jp=bench.make_corpus(256,16)
dec=Decoder(0,options=":fancy_upsampling=1")
b=dec.make_batch(jp,"I_RGB")
dec.prepare(b)
for i in range(2): dec.execute()
torch.cuda.synchronize()
PY
ls -la gpurun_out/${TAG}_d$d.ncu-rep
done
