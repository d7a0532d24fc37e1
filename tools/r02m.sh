#!/bin/bash
mkdir -p gpurun_out/r02m
O=gpurun_out/r02m
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
bash tools/checked_suite.sh $O
for wl in city4k stress8k_ext stress8k; do
  steps=200; [ $wl != city4k ] && steps=40
  timeout 600 python bench.py --steps $steps --e2e-steps 10 --no-cpu-baseline --workload $wl 2>$O/err_$wl.txt | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/bench.txt
done
