#!/bin/bash
# 2-GPU validation of the head build: the multi-GPU oracle test on different devices, bench at N=2
mkdir -p gpurun_out/r02n2
O=gpurun_out/r02n2
timeout 600 python -m pytest tests/test_gpu_bands_fused.py -m gpu -x -q > $O/pytest_multi.log 2>&1; echo "rc=$?" >> $O/pytest_multi.log
tail -3 $O/pytest_multi.log
runN() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $2 --warmup 5 $3 > $O/bench_$1.json 2> $O/bench_$1.err || tail -8 $O/bench_$1.err; }
runN city4k_n2 2 "--steps 300 --e2e-steps 60"
runN stress8k_n2 2 "--workload stress8k --steps 60 --e2e-steps 20 --no-other-gather"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02n2/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        o=d.get('value_other_gather') or {}
        e=d['e2e']; r8=e.get('rgba8_present_parallel') or {}
        print(f.split('/')[-1], 'value', round(d['value'],1), 'other', o.get('format'), round(o.get('value',0),1), o.get('verified'), 'e2e', round(e['value'],1), e.get('rows_verified'), 'e2e rgba8 parallel', round(r8.get('value',0),1), r8.get('rows_verified'), 'verified', d['config_detail']['bands']['gather_verified'])
        for r,s in enumerate(d['stage_ms_ranks']): print('    r',r,{k: round(v*1000,1) for k,v in s.items()})
    except Exception as ex:
        print(f, 'FAILED', ex)
PY
