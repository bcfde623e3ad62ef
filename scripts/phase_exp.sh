cd /root/repo
for d in 0 2 4; do
  B200JPEG_PX_DEBUG=$d python - <<PY
import os,sys,torch,time
sys.path.insert(0,'.')
import bench
from nvimagecodec_b200 import Decoder,_lib
jp=bench.make_corpus(256,16)
dec=Decoder(0,options=":fancy_upsampling=1")
b=dec.make_batch(jp,"I_RGB")
dec.set_profiling(True)
dec.prepare(b)
for i in range(3): dec.execute()
torch.cuda.synchronize()
acc={}
for i in range(10):
    dec.execute(); torch.cuda.synchronize()
    for k,v in dec.stage_times().items(): acc[k]=acc.get(k,0)+v/10
print("PX_DEBUG=$d", {k:round(v,3) for k,v in acc.items() if v>0.01})
PY
done
