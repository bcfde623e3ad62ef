#!/bin/bash
# e2e numbers per decoder option string (experiment helper): WORKLOAD=c3 e2e_sweep.sh "opt1" "opt2" ...
cd "$(dirname "$0")/.."
for o in "$@"; do
  python bench.py --steps 5 --warmup 3 --workload ${WORKLOAD:-c2} --options " $o" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('${WORKLOAD:-c2}', '$o', 'kernels', round(d['value']), 'e2e', round(d['e2e']['value']), 'dev_out', round(d['e2e_device_out']['value']))"
done
