#!/bin/bash
# 2 GPUs: multi-GPU parity test, N=1 e2e check, N=2 bench with the three frame-barrier placements
mkdir -p gpurun_out/r02e
O=gpurun_out/r02e
nvidia-smi --query-gpu=index,name --format=csv > $O/gpus.txt
timeout 600 python -m pytest tests/test_gpu_bands_fused.py -m gpu -x -q > $O/pytest_multi.log 2>&1; echo "rc=$?" >> $O/pytest_multi.log
tail -5 $O/pytest_multi.log
for wl in city4k gun1080; do
  timeout 600 python bench.py --workload $wl --steps 300 --warmup 5 --no-cpu-baseline > $O/bench_${wl}_n1.json 2> $O/bench_${wl}_n1.err || tail -5 $O/bench_${wl}_n1.err
done
run2() { # name, extra args
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 300 --warmup 5 $2 > $O/bench_$1.json 2> $O/bench_$1.err || tail -8 $O/bench_$1.err
}
run2 n2_side ""
run2 n2_inline "--frame-barrier inline --no-other-gather"
run2 n2_none "--frame-barrier none --no-other-gather"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02e/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        o=d.get('value_other_gather') or {}
        print(f.split('/')[-1], 'value', round(d['value'],1), 'other', o.get('format'), round(o.get('value',0),1), o.get('verified'), 'e2e', round(d['e2e']['value'],1), d['e2e'].get('rows_verified'), 'rgba8', round((d.get('e2e_via_gpu0_gather') or d['e2e'])['rgba8_present']['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()})
        if d.get('stage_ms_ranks'):
            for r,s in enumerate(d['stage_ms_ranks']): print('    r',r,{k: round(v*1000,1) for k,v in s.items()})
            print('   cuts', d['config_detail']['bands']['cuts_tile_rows'], 'e2e cuts', d['e2e'].get('cuts_tile_rows'), 'verified', d['config_detail']['bands']['gather_verified'])
    except Exception as e:
        print(f, 'FAILED', e)
PY
