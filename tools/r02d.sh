#!/bin/bash
mkdir -p gpurun_out/r02d
O=gpurun_out/r02d
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
for wl in city4k gun1080 stress8k; do
  steps=300; [ $wl = stress8k ] && steps=60
  timeout 600 python bench.py --workload $wl --steps $steps --warmup 5 --no-cpu-baseline > $O/bench_$wl.json 2> $O/bench_$wl.err || tail -5 $O/bench_$wl.err
done
timeout 200 python tools/band_probe.py 68 70 city4k 2>&1 | tee $O/probe_68_70.txt
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02d/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'rgba8', round(d['e2e']['rgba8_present']['value'],1), 'cb', round(d['e2e']['colorbuffer_float4_sync']['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))
    except Exception as e:
        print(f, 'FAILED', e)
PY
