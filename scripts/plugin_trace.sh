#!/bin/bash
# GPU box: the plugin path (host_driver -> nvimgcodecDecoderDecode -> extension) on the C2 batch with host-phase tracing
cd "$(dirname "$0")/.."
python - <<PY
import sys,os
sys.path.insert(0,'.')
import bench
jp=bench.make_corpus(256,16)
os.makedirs('/tmp/c2',exist_ok=True)
with open('/tmp/c2/manifest_device.txt','w') as m, open('/tmp/c2/manifest_host.txt','w') as mh:
    for i,d in enumerate(jp):
        p='/tmp/c2/%d.jpg'%i
        open(p,'wb').write(d)
        m.write(p+' I_RGB 0 0 0 0 device\n'); mh.write(p+' I_RGB 0 0 0 0 host\n')
PY
mkdir -p /tmp/c2/out
export LD_LIBRARY_PATH=$PWD/oracle/_ref:/usr/local/cuda/lib64:$LD_LIBRARY_PATH
export HOST_DRIVER_CRC=1000000 HOST_DRIVER_PINNED=1
for where in device host; do
echo "== $where outputs ${OPTS}"
B200JPEG_TRACE=1 oracle/_ref/host_driver nvimagecodec_b200/lib /tmp/c2/manifest_$where.txt /tmp/c2/out ":fancy_upsampling=1${OPTS}" 6 2>&1 | grep -E "b200jpeg\]|BENCH" | tail -8
done
