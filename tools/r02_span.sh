#!/bin/bash
# Span resolve: parity subset on the default build, then A/B of the variants on three workloads.
mkdir -p gpurun_out/r02sp
O=gpurun_out/r02sp
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tiny_and_split.py tests/test_gpu_edges.py tests/test_gpu_states.py -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for wl in city4k gun1080 stress8k; do
  steps=200; [ $wl = gun1080 ] && steps=1000; [ $wl = stress8k ] && steps=60
  for so in build_ab/span0.so build_ab/span32.so softwarerenderer_b200/libsrb200.so build_ab/span256.so build_ab/span0.so softwarerenderer_b200/libsrb200.so; do
    SRB200_LIB=$PWD/$so python bench.py --steps $steps --no-cpu-baseline --workload $wl 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/ab.txt
  done
done
