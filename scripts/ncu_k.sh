#!/bin/bash
# GPU box: one `ncu --set full` capture of the kernels matching $1 (regex) on the C2 batch.  TAG names the report; OPTS = decoder options.
cd "$(dirname "$0")/.."
TAG=${TAG:-k}
ncu --set full --clock-control none --import-source on -k regex:$1 -c ${COUNT:-1} -f -o gpurun_out/${TAG} python - > gpurun_out/${TAG}.log 2>&1 <<PY
import sys,torch
sys.path.insert(0,'.')
import bench
from nvimagecodec_b200 import Decoder
jp=bench.make_corpus(256,16)
dec=Decoder(0,options=":fancy_upsampling=1${OPTS}")
b=dec.make_batch(jp,"I_RGB")
dec.prepare(b)
for i in range(2): dec.execute()
torch.cuda.synchronize()
PY
ls -la gpurun_out/${TAG}.ncu-rep
