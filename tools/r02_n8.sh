#!/bin/bash
# 8-GPU lines of the head build
mkdir -p gpurun_out/r02n8
O=gpurun_out/r02n8
runN() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $2 --warmup 5 $3 > $O/bench_$1.json 2> $O/bench_$1.err || tail -8 $O/bench_$1.err; }
runN city4k_n8 8 "--steps 300 --e2e-steps 60"
runN city4k_n4 4 "--steps 300 --e2e-steps 60"
runN stress8k_n8 8 "--workload stress8k --steps 60 --e2e-steps 20"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02n8/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        o=d.get('value_other_gather') or {}
        e=d['e2e']; r8=e.get('rgba8_present_parallel') or {}
        print(f.split('/')[-1], 'value', round(d['value'],1), 'other', o.get('format'), round(o.get('value',0),1), o.get('verified'), 'e2e', round(e['value'],1), e.get('rows_verified'), 'e2e rgba8 parallel', round(r8.get('value',0),1), r8.get('rows_verified'), 'verified', d['config_detail']['bands']['gather_verified'], 'cuts', d['config_detail']['bands'].get('cuts_tile_rows'))
        for r,s in enumerate(d['stage_ms_ranks']): print('    r',r,{k: round(v*1000,1) for k,v in s.items()})
    except Exception as ex:
        print(f, 'FAILED', ex)
PY
