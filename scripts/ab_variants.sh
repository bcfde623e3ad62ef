#!/bin/bash
# GPU box: parity (decode tests) and per-kernel times of the default build and of every variants/*.so, one after the other.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${TAG:-ab}
for lib in default variants/*.so; do
  [ "$lib" = "default" ] && unset B200JPEG_LIB || export B200JPEG_LIB=$PWD/$lib
  name=$(basename "$lib" .so)
  python -m pytest tests/test_decode_gpu.py tests/test_fullsize_gpu.py -x -q -m gpu 2>&1 | tail -3 | sed "s/^/[$name] /"
  python bench.py --steps 10 --warmup 3 --no-extras ${BENCH_ARGS} > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err || tail -c 600 gpurun_out/${TAG}_${name}.err
  python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_${name}.json"))
print("[$name]", round(d["value"]), "MPix/s", round(d["ms_per_step"], 3), "ms; capi dev-out", round(d["e2e_capi_device_out"]["value"]), "plugin dev-out", round(d["e2e_device_out"]["value"]), "plugin host-out", round(d["e2e"]["value"]))
print("[$name]", " ".join("%s=%.3f" % (s["kernel"], s["ms"]) for s in d["stages"]))
PY
done
