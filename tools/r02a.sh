#!/bin/bash
# round 2, run a: GPU tests, then the three workloads with the head build and two A/B builds of the direct threshold
mkdir -p gpurun_out/r02a
O=gpurun_out/r02a
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -15 $O/pytest_gpu.log
for wl in city4k stress8k gun1080; do
  steps=300; [ $wl = stress8k ] && steps=60
  timeout 600 python bench.py --workload $wl --steps $steps --warmup 5 --no-cpu-baseline > $O/bench_$wl.json 2> $O/bench_$wl.err
  for v in direct0 direct8; do
    SRB200_LIB=$PWD/build_ab/$v.so timeout 600 python bench.py --workload $wl --steps $steps --warmup 5 --no-cpu-baseline > $O/bench_${wl}_$v.json 2> $O/bench_${wl}_$v.err
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02a/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1), d['config_detail']['triangles_direct'], d['config_detail']['tile_pairs'])
    except Exception as e:
        print(f, 'FAILED', e)
PY
