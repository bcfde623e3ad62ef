#!/bin/bash
# GPU box, one GPU: the round's measurement set.  TAG names the files (gpurun_out/<TAG>_*).
cd "$(dirname "$0")/.."
TAG=${TAG:-r1m}
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
for w in c2 c3 c4 c5; do
  python bench.py --steps 20 --warmup 3 --workload $w > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err || tail -c 400 gpurun_out/${TAG}_bench_$w.err
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>/dev/null
python tests/tools/latency.py > gpurun_out/${TAG}_latency.txt 2>&1
python tests/tools/compare_nvjpeg.py > gpurun_out/${TAG}_compare_nvjpeg.txt 2>&1
# launch list (cold-cache, serialised: shares only) and one full-set capture of every kernel -- after the same command exited 0 above
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/${TAG}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -c 10 -o gpurun_out/${TAG}_prof python bench.py --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_full.log 2>&1
ls -la gpurun_out/${TAG}_*
python - <<PY
import json
for w in ("c2", "c3", "c4", "c5"):
    d = json.load(open("gpurun_out/${TAG}_bench_%s.json" % w))
    print(w, round(d["value"]), round(d["images_per_s"]), round(d["ms_per_step"], 3), round(d["e2e"]["value"]), round(d["e2e_device_out"]["value"]), round(d["cpu_baseline"]["value"]), d["clocks"])
PY
