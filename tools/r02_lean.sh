#!/bin/bash
mkdir -p gpurun_out/r02ln
O=gpurun_out/r02ln
for so in softwarerenderer_b200/libsrb200.so build_ab/big1.so build_ab/big2.so build_ab/fillrefs.so build_ab/both.so softwarerenderer_b200/libsrb200.so; do
  echo "== $so" | tee -a $O/probe.txt
  for b in "0 67" "68 69" "70 74" "88 111"; do
    SRB200_LIB=$PWD/$so python tools/band_probe.py $b 2>/dev/null | tee -a $O/probe.txt
  done
done
for wl in city4k gun1080; do
  steps=200; [ $wl = gun1080 ] && steps=1000
  for so in softwarerenderer_b200/libsrb200.so build_ab/fillrefs.so softwarerenderer_b200/libsrb200.so build_ab/fillrefs.so; do
    SRB200_LIB=$PWD/$so python bench.py --steps $steps --no-cpu-baseline --workload $wl 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/ab.txt
  done
done
