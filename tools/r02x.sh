#!/bin/bash
mkdir -p gpurun_out/r02x
O=gpurun_out/r02x
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for wl in city4k gun1080 stress8k; do
for so in build_ab/hiz0.so build_ab/hiz16.so softwarerenderer_b200/libsrb200.so build_ab/hiz64.so; do
  steps=300; [ $wl = stress8k ] && steps=40
  SRB200_LIB=$PWD/$so timeout 300 python bench.py --steps $steps --e2e-steps 5 --no-cpu-baseline --workload $wl 2>$O/err.txt | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/ab.txt
done; done
for so in build_ab/hiz0.so softwarerenderer_b200/libsrb200.so; do
  echo "== $so" | tee -a $O/probe.txt
  SRB200_LIB=$PWD/$so timeout 300 python tools/band_probe.py 68 69 city4k 2>&1 | tail -1 | tee -a $O/probe.txt
  SRB200_LIB=$PWD/$so timeout 300 python tools/band_probe.py 68 70 city4k 2>&1 | tail -1 | tee -a $O/probe.txt
done
