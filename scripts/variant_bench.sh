#!/bin/bash
# stage times for each prebuilt variant of the library under scripts/variants/ (experiment helper)
cd "$(dirname "$0")/.."
cp nvimagecodec_b200/lib/libb200jpeg_ext.so /tmp/orig.so
for v in scripts/variants/*.so; do
  echo "=== $v"
  cp "$v" nvimagecodec_b200/lib/libb200jpeg_ext.so
  python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],3), [(s['kernel'].split('(')[1][:-1], s['ms']) for s in d['stages']])"
done
cp /tmp/orig.so nvimagecodec_b200/lib/libb200jpeg_ext.so
