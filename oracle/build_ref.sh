#!/usr/bin/env bash
# Builds, from the reference's OWN sources where they lie under /root/reference, the pieces the
# tests and the CPU baseline need.  Outputs go only to oracle/_ref/ (git-ignored, shipped to the
# GPU box).  Nothing is copied into the repository; the reference's CMake build is not run
# (it needs zstd/OpenCV/network, see DESIGN.md "Reference builds").
#
#   oracle/_ref/libnvimgcodec.so   the unmodified nvImageCodec core (src/*.cpp, src/parsers, src/imgproc)
#                                  -- the host framework that loads our extension in tests/
#   oracle/_ref/host_driver        tests/host/host_driver.cpp linked against it (our own file)
#   oracle/_ref/nvimtrans          example/nvimtrans/main.cpp, unmodified, + oracle/_ref/extensions/libnvbmp_ext.so (extensions/nvbmp)
#   oracle/_ref/python/nvidia/nvimgcodec/           the reference's pybind11 module (python/*.cpp), unmodified: `from nvidia import nvimgcodec`
#   oracle/_ref/extensions_nvjpeg/libnvjpeg_ext.so   extensions/nvjpeg (unmodified) over the toolkit's libnvjpeg: GPU comparison point
#   oracle/_ref/libref_turbo.so    the reference's CPU JPEG glue (extensions/libjpeg_turbo/jpeg_mem.cpp, jpeg_handle.cpp)
#                                  + oracle/ref_turbo_shim.cpp, compiled against the reconstructed libjpeg62 ABI header
#                                  (oracle/turbo_abi/) and linked to the libjpeg-turbo binary bundled with Pillow
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
OBJ="$OUT/obj"
GEN="$OUT/gen"
CUDA="${CUDA_HOME:-/usr/local/cuda}"
JOBS="${JOBS:-$(nproc)}"
[ -d "$REF/src" ] || { echo "reference tree not found at $REF; keeping prebuilt oracle/_ref" >&2; exit 0; }
mkdir -p "$OBJ" "$GEN"

sed -e 's/@PROJECT_VERSION_MAJOR@/0/;s/@PROJECT_VERSION_MINOR@/8/;s/@PROJECT_VERSION_PATCH@/0/;s/@PROJECT_VERSION_TWEAK@/0/' \
    "$REF/include/nvimgcodec_version.h.in" > "$GEN/nvimgcodec_version.h.tmp"
cmp -s "$GEN/nvimgcodec_version.h.tmp" "$GEN/nvimgcodec_version.h" 2>/dev/null || mv "$GEN/nvimgcodec_version.h.tmp" "$GEN/nvimgcodec_version.h"
rm -f "$GEN/nvimgcodec_version.h.tmp"

INC="-I$GEN -I$REF/include -I$REF/src -I$REF -I$REF/external/NVTX/c/include -I$REF/external/boost/preprocessor/include -I$CUDA/include"
# -fvisibility=hidden as in the reference's CMakeLists.txt:314 (the core and the Python binding both define nvimgcodec::Image)
CXXFLAGS="-std=c++20 -O2 -fPIC -fvisibility=hidden -DNDEBUG -w"

SRCS="nvimgcodec_capi plugin_framework codec_registry codec image_decoder image_encoder image_parser code_stream file_io_stream
 std_file_io_stream mmaped_file_io_stream image image_parser_factory image_encoder_factory image_decoder_factory image_generic_decoder
 image_generic_encoder nvimgcodec_director processing_results decode_state_batch encode_state_batch thread_pool default_executor
 builtin_modules parsers/bmp parsers/exif parsers/jpeg parsers/jpeg2k parsers/tiff parsers/png parsers/pnm parsers/webp
 parsers/parsers_ext_module imgproc/pinned_buffer imgproc/device_buffer imgproc/exception imgproc/out_of_bound_roi_fill
 imgproc/stream_device dynlink/dynlink_cuda"

{
  echo "all: objs"
  OBJS=""
  for s in $SRCS; do
    [ -f "$REF/src/$s.cpp" ] || continue
    o="$OBJ/$(echo "$s" | tr '/' '_').o"
    OBJS="$OBJS $o"
    printf '%s: %s\n\tg++ %s %s -c $< -o $@\n' "$o" "$REF/src/$s.cpp" "$CXXFLAGS" "$INC"
  done
  o="$OBJ/imgproc_convert_kernel_gpu.o"
  OBJS="$OBJS $o"
  printf '%s: %s\n\tnvcc -gencode arch=compute_100a,code=sm_100a -std=c++20 -O2 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -DNDEBUG -w %s -c $< -o $@\n' \
      "$o" "$REF/src/imgproc/convert_kernel_gpu.cu" "$INC"
  echo "objs:$OBJS"
} > "$OUT/ref.mk"
make -s -j"$JOBS" -f "$OUT/ref.mk"

OBJS=$(ls "$OBJ"/*.o)   # (the turbo glue objects use the .to suffix so they stay out of libnvimgcodec.so)
g++ -shared -o "$OUT/libnvimgcodec.so" $OBJS -L"$CUDA/lib64" -L"$CUDA/lib64/stubs" -lcudart_static -lcuda -ldl -lpthread -lrt
if [ -f "$HERE/../tests/host/host_driver.cpp" ]; then
  g++ -std=c++17 -O2 -I"$GEN" -I"$REF/include" -I"$CUDA/include" "$HERE/../tests/host/host_driver.cpp" -o "$OUT/host_driver" \
      -L"$OUT" -lnvimgcodec -Wl,-rpath,'$ORIGIN' -Wl,--allow-shlib-undefined -L"$CUDA/lib64" -lcudart_static -ldl -lpthread -lrt
fi
# ---- the nvimtrans sample and the dependency-free nvbmp extension (its output encoder), both unmodified:
#      the named drop-in consumer of SURVEY.md 8(f) n1
mkdir -p "$OUT/extensions"
for f in nvbmp_ext encoder decoder ext_module; do
  g++ $CXXFLAGS $INC -I"$REF/extensions/nvbmp" -c "$REF/extensions/nvbmp/$f.cpp" -o "$OBJ/nvbmp_$f.bo"
done
g++ -shared -o "$OUT/extensions/libnvbmp_ext.so" "$OBJ"/nvbmp_*.bo -lpthread
g++ -std=c++20 -O2 -w $INC -I"$REF/example/nvimtrans" "$REF/example/nvimtrans/main.cpp" -o "$OUT/nvimtrans" \
    -L"$OUT" -lnvimgcodec -Wl,-rpath,'$ORIGIN' -Wl,--allow-shlib-undefined -L"$CUDA/lib64" -lcudart_static -ldl -lpthread -lrt

# ---- the reference's Python binding (python/*.cpp, pybind11 from external/pybind11, unmodified).  Its dlpack submodule is
#      absent from the tree; the same public header ships with tvm_ffi in this image.
PYINC="$(python - <<'PY'
import sysconfig
print(sysconfig.get_paths()["include"])
PY
)"
PYEXT="$(python - <<'PY'
import sysconfig
print(sysconfig.get_config_var("EXT_SUFFIX"))
PY
)"
DLPACK="$(python - <<'PY'
import glob, os, sys
c = glob.glob(os.path.join(sys.prefix, "lib", "python*", "site-packages", "tvm_ffi", "3rdparty", "dlpack", "include"))
print(c[0] if c else "")
PY
)"
if [ -n "$DLPACK" ] && [ -f "$PYINC/Python.h" ] && [ -d "$REF/external/pybind11/include/pybind11" ]; then
  PYOUT="$OUT/python/nvidia/nvimgcodec"
  mkdir -p "$PYOUT" "$OBJ/py"
  for f in "$REF"/python/*.cpp; do
    b="$(basename "$f" .cpp)"
    [ "$OBJ/py/$b.po" -nt "$f" ] && continue
    g++ -std=c++20 -O1 -fPIC -fvisibility=hidden -w -DNDEBUG $INC -I"$REF/python" -I"$REF/external/pybind11/include" -I"$DLPACK" -I"$PYINC" -c "$f" -o "$OBJ/py/$b.po" &
  done
  wait
  g++ -shared -o "$PYOUT/nvimgcodec_impl$PYEXT" "$OBJ"/py/*.po "$OBJ"/imgproc_*.o "$OBJ/dynlink_dynlink_cuda.o" -L"$OUT" -lnvimgcodec \
      -Wl,-rpath,'$ORIGIN/../../..' -Wl,--allow-shlib-undefined -L"$CUDA/lib64" -L"$CUDA/lib64/stubs" -lcudart_static -lcuda -ldl -lpthread -lrt
  sed -e 's/@GIT_SHA@/36b4ee8/' "$REF/python/__init__.py.in" > "$PYOUT/__init__.py"
  cp "$REF/python/nvidia/__init__.py" "$OUT/python/nvidia/__init__.py"
fi

# ---- the reference's nvjpeg_ext (wrapper over the closed nvJPEG library of the CUDA toolkit), unmodified: the GPU
#      comparison point of SURVEY.md 8(d).  Linked directly against libnvjpeg (no dlopen stubs: those need libclang).
if [ -f "$CUDA/include/nvjpeg.h" ] && ls "$CUDA"/lib64/libnvjpeg.so* >/dev/null 2>&1; then
  mkdir -p "$OUT/extensions_nvjpeg"
  NVJ_DEFS="-DNVJPEG_LOSSLESS_SUPPORTED=0 -DNVJPEG_HW_ENCODER_SUPPORTED=0 -DCUDA_VERSION_MAJOR=12 -DCUDA_VERSION_MINOR=9"
  for f in cuda_decoder hw_decoder cuda_encoder type_convert nvjpeg_ext nvjpeg_utils exception ext_module; do
    g++ $CXXFLAGS $NVJ_DEFS $INC -I"$REF/extensions" -I"$REF/extensions/nvjpeg" -c "$REF/extensions/nvjpeg/$f.cpp" -o "$OBJ/nvjpeg_$f.jo" &
  done
  wait
  g++ -shared -o "$OUT/extensions_nvjpeg/libnvjpeg_ext.so" "$OBJ"/nvjpeg_*.jo "$OBJ"/imgproc_*.o "$OBJ/dynlink_dynlink_cuda.o" \
      -L"$CUDA/lib64" -L"$CUDA/lib64/stubs" -lnvjpeg -lcudart_static -lcuda -ldl -lpthread -lrt
fi

# ---- the reference's own CPU JPEG path over the libjpeg-turbo binary that ships with Pillow
TURBO_SO="$(python - <<'PY'
import glob, os, sys
try:
    import PIL
    c = sorted(glob.glob(os.path.join(os.path.dirname(PIL.__file__), "..", "pillow.libs", "libjpeg-*.so.62*")))
    print(os.path.realpath(c[0]) if c else "")
except Exception:
    print("")
PY
)"
if [ -n "$TURBO_SO" ] && [ -f "$HERE/ref_turbo_shim.cpp" ]; then
  TINC="-I$HERE/turbo_abi $INC -I$REF/extensions/libjpeg_turbo"
  g++ $CXXFLAGS -fvisibility=hidden $TINC -c "$REF/extensions/libjpeg_turbo/jpeg_mem.cpp" -o "$OBJ/turbo_jpeg_mem.to"
  g++ $CXXFLAGS -fvisibility=hidden $TINC -c "$REF/extensions/libjpeg_turbo/jpeg_handle.cpp" -o "$OBJ/turbo_jpeg_handle.to"
  g++ $CXXFLAGS -fvisibility=hidden $TINC -c "$HERE/ref_turbo_shim.cpp" -o "$OBJ/turbo_shim.to"
  g++ -shared -o "$OUT/libref_turbo.so" "$OBJ/turbo_jpeg_mem.to" "$OBJ/turbo_jpeg_handle.to" "$OBJ/turbo_shim.to" "$TURBO_SO" \
      -Wl,-rpath,"$(dirname "$TURBO_SO")" -lpthread
  echo "$TURBO_SO" > "$OUT/libref_turbo.engine"
else
  echo "Pillow's libjpeg-turbo not found: libref_turbo.so not built" >&2
fi
echo "built $(ls "$OUT" | grep -v -E '^(obj|gen|ref.mk)$' | tr '\n' ' ')"
