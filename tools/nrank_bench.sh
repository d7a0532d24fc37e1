#!/bin/bash
# N-rank bench helper for gpurun calls: source it, then `run <N> <port> [bench.py flags]`.
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $1 --steps 300 --e2e-steps 20 "${@:3}" 2>>gpurun_out/nn.err | tail -1 | tee gpurun_out/scale_n$1_$2.json | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$*', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['config']['bands']['cuts_tile_rows'], 'via gpu0', round(d.get('e2e_via_gpu0_gather',{}).get('value') or 0,1), 'verified', d['e2e'].get('rows_verified'), 'gather', d['config']['bands'].get('gather_format'), d['config']['bands'].get('gather_verified'), 'rgba8 extra', (d.get('value_rgba8_gather') or {}).get('value'), (d.get('value_rgba8_gather') or {}).get('verified'))
for r in d['stage_ms_ranks']: print('   ', {k: round(v*1000,1) for k,v in r.items()})"; }
