#!/bin/bash
# GPU box, one GPU: the round's measurement set.  TAG names the files (gpurun_out/<TAG>_*).
cd "$(dirname "$0")/.."
TAG=${TAG:-r2m}
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 > gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench_c2.err || tail -c 400 gpurun_out/${TAG}_bench_c2.err
for w in c3 c4 c5; do
  python bench.py --steps 10 --warmup 3 --no-extras --workload $w > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err || tail -c 400 gpurun_out/${TAG}_bench_$w.err
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>/dev/null
python tests/tools/latency.py > gpurun_out/${TAG}_latency.txt 2>&1
python scripts/prog_bench.py 256 > gpurun_out/${TAG}_progressive.json 2> gpurun_out/${TAG}_progressive.err || tail -c 400 gpurun_out/${TAG}_progressive.err
python scripts/prog_one.py 90 2>&1 | grep "chain" | tail -10 > gpurun_out/${TAG}_progressive_chains.txt
# (compute-sanitizer is closed on this GPU pool: no memcheck / racecheck log can be produced here; see profiles/README.md)
# launch list (cold-cache, serialised: shares only) and one full-set capture of every kernel -- after the same command exited 0 above
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/${TAG}_ncu_launches.log 2>&1
TAG=${TAG}_prof COUNT=10 scripts/ncu_k.sh "."
TAG=${TAG}_prog N=4 COUNT=2 scripts/ncu_prog.sh
ls -la gpurun_out/${TAG}_*
python - <<PY
import json
for w in ("c2", "c3", "c4", "c5"):
    d = json.load(open("gpurun_out/${TAG}_bench_%s.json" % w))
    print(w, round(d["value"]), round(d["images_per_s"]), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]), "dev-out", round(d["e2e_device_out"]["value"]), "capi", round(d["e2e_capi"]["value"]), round(d["e2e_capi_device_out"]["value"]), d["clocks"])
PY
