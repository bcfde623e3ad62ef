#!/bin/bash
# GPU box: `ncu --set full` capture of the progressive scan kernel (every wave) on N 1080p progressive images.  TAG names the report.
cd "$(dirname "$0")/.."
TAG=${TAG:-prog}
N=${N:-4}
ncu --set full --clock-control none --import-source on -k regex:prog_ -c ${COUNT:-8} -f -o gpurun_out/${TAG} python - > gpurun_out/${TAG}.log 2>&1 <<PY
import sys,torch
sys.path.insert(0,'.')
from nvimagecodec_b200 import Decoder, corpus
jp=[corpus.encode(corpus.synth(1080,1920,s),90,"420",progressive=True) for s in range($N)]
dec=Decoder(0,options=":fancy_upsampling=1${OPTS}")
b=dec.make_batch(jp,"I_RGB")
dec.prepare(b)
dec.execute()
torch.cuda.synchronize()
PY
ls -la gpurun_out/${TAG}.ncu-rep
