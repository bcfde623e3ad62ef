#!/bin/bash
# kernel-only stage times for a list of decoder option strings (experiment helper): opt_sweep.sh "opt1" "opt2" ...
cd "$(dirname "$0")/.."
for o in "$@"; do
  echo "=== $o"
  python bench.py --steps 5 --warmup 3 --workload ${WORKLOAD:-c2} --options " $o" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],3), [(s['kernel'].split('(')[1][:-1], s['ms']) for s in d['stages']])"
done
