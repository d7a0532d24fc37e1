#!/bin/bash
mkdir -p gpurun_out/r02w
O=gpurun_out/r02w
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for args in "0 0 gun1080" "0 0 city4k" "68 69 city4k" "88 111 city4k" "0 0 stress8k"; do
  timeout 300 python tools/band_probe.py $args 2>&1 | tee -a $O/probe.txt
done
for wl in gun1080 city4k; do
  timeout 600 python bench.py --workload $wl --steps 300 --warmup 5 --no-cpu-baseline > $O/bench_$wl.json 2> $O/bench_$wl.err || tail -5 $O/bench_$wl.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02w/bench_*.json')):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f.split('/')[-1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'rgba8', round(d['e2e']['rgba8_present']['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))
PY
