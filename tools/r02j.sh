#!/bin/bash
# 2 GPUs: no-alloc test alone (fresh process), then the whole GPU suite (it runs last there), then N=2 bench
mkdir -p gpurun_out/r02j
O=gpurun_out/r02j
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> $O/pytest_noalloc_alone.log
tail -3 $O/pytest_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
run2() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 300 --warmup 5 --e2e-steps 50 $2 > $O/bench_$1.json 2> $O/bench_$1.err || tail -8 $O/bench_$1.err; }
run2 n2_peer ""
run2 n2_side "--frame-barrier side --no-other-gather"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02j/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        o=d.get('value_other_gather') or {}
        print(f.split('/')[-1], 'value', round(d['value'],1), 'other', o.get('format'), round(o.get('value',0),1), o.get('verified'), 'e2e', round(d['e2e']['value'],1), d['e2e'].get('rows_verified'), 'verified', d['config_detail']['bands']['gather_verified'])
    except Exception as e:
        print(f, 'FAILED', e)
PY
