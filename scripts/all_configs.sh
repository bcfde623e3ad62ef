#!/bin/bash
# one bench line per BASELINE config (c2 is the headline); writes gpurun_out/bench_<cfg>.json
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for w in ${CONFIGS:-c2 c3 c4 c5}; do
  python bench.py --steps ${STEPS:-10} --warmup 3 --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || tail -c 400 gpurun_out/bench_$w.err
  python - $w <<'PY'
import json, sys
d = json.load(open("gpurun_out/bench_%s.json" % sys.argv[1]))
print(sys.argv[1], "value", round(d["value"]), "img/s", round(d["images_per_s"]), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]),
      "dev_out", round(d["e2e_device_out"]["value"]), "cpu", round(d["cpu_baseline"]["value"]))
print("   ", [(s["kernel"].split("(")[1][:-1], s["ms"]) for s in d["stages"]])
PY
done
