#!/bin/bash
mkdir -p gpurun_out/r02rf
O=gpurun_out/r02rf
N=$1
runN() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $2 --warmup 5 $3 > $O/bench_$1.json 2> $O/bench_$1.err || tail -8 $O/bench_$1.err; }
runN city4k_n${N} $N "--steps 300 --e2e-steps 40"
runN city4k_n${N}_norefine $N "--steps 300 --e2e-steps 40 --refine-cuts 0 --no-other-gather"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02rf/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        b=d['config_detail']['bands']
        print(f.split('/')[-1], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'verified', b['gather_verified'], d['e2e'].get('rows_verified'), 'cuts', b['cuts_tile_rows'])
        for t in (b.get('cuts_refinement') or []): print('    ', t)
    except Exception as ex:
        print(f, 'FAILED', ex)
PY
