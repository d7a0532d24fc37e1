#!/bin/bash
# Final 1-GPU evidence of the round: smoke, GPU suite, the default bench line (as the driver runs it), the other workloads,
# the reference arm.
mkdir -p gpurun_out/r02f1
O=gpurun_out/r02f1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log; tail -5 $O/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python bench.py > $O/bench_city4k.json 2> $O/bench_city4k.err || tail -5 $O/bench_city4k.err
python bench.py --impl reference --steps 20 --warmup 3 > $O/bench_city4k_reference.json 2> $O/ref.err
for wl in gun1080 stress8k stress8k_ext; do
  steps=1000; [ $wl != gun1080 ] && steps=100
  python bench.py --workload $wl --steps $steps > $O/bench_$wl.json 2> $O/bench_$wl.err || tail -5 $O/bench_$wl.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02f1/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if d.get('impl')=='reference':
            print(f.split('/')[-1], 'REFERENCE', round(d['value'],2), d['cpu_baseline']['cores'], d['steps'], d['warmup']); continue
        print(f.split('/')[-1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'rgba8', round(d['e2e']['rgba8_present']['value'],1), 'cb', round(d['e2e']['colorbuffer_float4_sync']['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1), 'frac', round(d['roofline']['frac'],3), 'cpu', d.get('cpu_baseline',{}).get('value'), d['clocks'])
    except Exception as e:
        print(f, 'FAILED', e)
PY
