#!/bin/bash
# GPU box: parity tests, then a short bench; prints the per-kernel table.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -5
python bench.py --steps ${STEPS:-10} --warmup 3 ${BENCH_ARGS:-} > gpurun_out/bench_new.json 2> gpurun_out/bench_new.err || tail -c 600 gpurun_out/bench_new.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_new.json"))
print("value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), "dev_out", round(d["e2e_device_out"]["value"]))
for s in d["stages"]:
    print(s)
PY
