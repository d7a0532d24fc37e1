#!/bin/bash
mkdir -p gpurun_out/r02t
O=gpurun_out/r02t
SRB200_LIB=$PWD/build_ab/bulk.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tiny_and_split.py tests/test_gpu_edges.py -m gpu -x -q 2>&1 | tail -2 | tee $O/pytest_bulk.log
for rep in 1 2; do
for wl in city4k gun1080 stress8k; do
for so in softwarerenderer_b200/libsrb200.so build_ab/bulk.so; do
  steps=300; [ $wl = stress8k ] && steps=40
  SRB200_LIB=$PWD/$so timeout 300 python bench.py --steps $steps --e2e-steps 5 --no-cpu-baseline --workload $wl 2>$O/err.txt | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/ab.txt
done; done; done
