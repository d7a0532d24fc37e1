#!/bin/bash
mkdir -p gpurun_out/r02n8b
O=gpurun_out/r02n8b
runN() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $2 --warmup 5 $3 > $O/bench_$1.json 2> $O/bench_$1.err || tail -8 $O/bench_$1.err; }
runN city4k_n8 8 "--steps 300 --e2e-steps 40"
runN stress8k_n8 8 "--workload stress8k --steps 60 --e2e-steps 16 --no-other-gather"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02n8b/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        b=d['config_detail']['bands']; o=d.get('value_other_gather') or {}
        print(f.split('/')[-1], 'value', round(d['value'],1), 'other', round(o.get('value',0),1), 'e2e', round(d['e2e']['value'],1), 'verified', b['gather_verified'], d['e2e'].get('rows_verified'), 'cuts', b['cuts_tile_rows'])
        for t in (b.get('cuts_refinement') or []): print('    ', t)
    except Exception as ex:
        print(f, 'FAILED', ex)
PY
