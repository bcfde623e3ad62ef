#!/bin/bash
# Stepwise run of the reference's Python binding (oracle/_ref/python) against this extension, to localise a crash.
cd "$(dirname "$0")/../.."
R=$PWD
mkdir -p gpurun_out /tmp/extdir
cp nvimagecodec_b200/lib/libb200jpeg_ext.so /tmp/extdir/
export PYTHONPATH=$R/oracle/_ref/python LD_LIBRARY_PATH=$R/oracle/_ref PYNVIMGCODEC_VERBOSITY=${V:-3}
cat > /tmp/segv_bt.c <<'C'
#define _GNU_SOURCE
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>
static void h(int s){ void* b[64]; int n = backtrace(b, 64); backtrace_symbols_fd(b, n, 2); _exit(139); }
__attribute__((constructor)) static void init(void){ void* b[2]; backtrace(b, 2); signal(SIGSEGV, h); }
C
gcc -shared -fPIC -o /tmp/segv_bt.so /tmp/segv_bt.c
step() { echo "=== $1"; shift; LD_PRELOAD=/tmp/segv_bt.so timeout 120 python -c "$@" 2>&1 | tail -40; echo "rc=${PIPESTATUS[0]}"; }
step "import, no extension path" "
from nvidia import nvimgcodec
print('imported', nvimgcodec.__version__)"
export NVIMGCODEC_EXTENSIONS_PATH=/tmp/extdir
step "import with extension path" "
from nvidia import nvimgcodec
print('imported', nvimgcodec.__version__)"
step "Decoder()" "
from nvidia import nvimgcodec
d = nvimgcodec.Decoder()
print('decoder ok')"
step "decode one" "
import numpy as np, glob
from nvidia import nvimgcodec
d = nvimgcodec.Decoder()
f = sorted(glob.glob('$R/tests/golden/jpeg/*.jpg'))[0]
im = d.decode(open(f,'rb').read())
print(im.shape, im.dtype)
a = np.asarray(im.cpu())
print(a.shape, a.mean())"
