#!/bin/bash
mkdir -p gpurun_out/r02p
O=gpurun_out/r02p
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 --e2e-steps 60 --no-other-gather > $O/bench_n2.json 2> $O/bench_n2.err || tail -8 $O/bench_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02p/bench_n2.json').read().strip().splitlines()[-1])
print(d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e'].get('host_us_per_call_rank0'), d['e2e'].get('cuts_tile_rows'))
print(d['stage_ms_ranks'])
PY
