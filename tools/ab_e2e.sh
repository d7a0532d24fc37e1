#!/bin/bash
# A/B of alternative builds on the end-to-end loop: prints value, e2e and the float4 read-back variant.
for so in "$@"; do
  SRB200_LIB=$PWD/$so python bench.py --steps 100 --e2e-steps 60 --no-cpu-baseline --workload ${WORKLOAD:-city4k} 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$so', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'float4 sync', round(d['e2e']['colorbuffer_float4_sync']['value'],1), 'raster', round(d['stage_ms']['tile_raster']*1000,1))"
done
