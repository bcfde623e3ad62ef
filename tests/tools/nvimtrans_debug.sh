#!/bin/bash
cd "$(dirname "$0")/../.."
T=$(mktemp -d); mkdir -p $T/ext $T/in $T/out
cp nvimagecodec_b200/lib/libb200jpeg_ext.so oracle/_ref/extensions/libnvbmp_ext.so $T/ext/
python - <<PY
import sys; sys.path.insert(0, '.')
from nvimagecodec_b200 import corpus
open('$T/in/a.jpg','wb').write(corpus.encode(corpus.synth(240,320,1),88,'420'))
open('$T/in/b.jpg','wb').write(corpus.encode(corpus.synth(199,301,2),88,'444',5))
PY
export NVIMGCODEC_EXTENSIONS_PATH=$T/ext LD_LIBRARY_PATH=$PWD/oracle/_ref:$LD_LIBRARY_PATH
oracle/_ref/nvimtrans -i $T/in -o $T/out -c bmp -b 2 -v 4 2>&1 | tail -40
ls -la $T/out
