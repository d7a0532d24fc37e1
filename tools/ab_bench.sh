#!/bin/bash
# A/B of alternative builds of libsrb200.so (build_ab/*.so) on one GPU box: prints fps and stage times.
for so in "$@"; do
  SRB200_LIB=$PWD/$so python bench.py --steps 100 --no-cpu-baseline --workload ${WORKLOAD:-city4k} 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()})"
done
