#!/bin/bash
# kernel-only stage times for each subsequence length (experiment helper)
cd "$(dirname "$0")/.."
for sb in 1024 2048 4096; do
  echo "=== subseq_bits=$sb"
  python bench.py --steps 5 --warmup 3 --options " :subseq_bits=$sb" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],3), [(s['kernel'], s['ms']) for s in d['stages']])"
done
python scripts/sync_stats.py 2>&1 | tail -3
