#!/bin/bash
# Extra CTAs of the tile kernel sized per frame: parity on the split / band tests, A/B against the previous head.
mkdir -p gpurun_out/r02ex
O=gpurun_out/r02ex
timeout 900 python -m pytest tests/test_gpu_tiny_and_split.py tests/test_gpu_bands_fused.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
SRB200_LIB=$PWD/build_ab/checked.so timeout 600 python -m pytest tests/test_gpu_tiny_and_split.py tests/test_gpu_bands_fused.py -m gpu -x -q > $O/pytest_checked.log 2>&1; echo "rc=$?" >> $O/pytest_checked.log
tail -2 $O/pytest_checked.log
for wl in city4k gun1080; do
  steps=200; [ $wl = gun1080 ] && steps=1000
  for so in build_ab/prev.so softwarerenderer_b200/libsrb200.so build_ab/prev.so softwarerenderer_b200/libsrb200.so; do
    SRB200_LIB=$PWD/$so python bench.py --steps $steps --no-cpu-baseline --workload $wl 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/ab.txt
  done
done
for so in build_ab/prev.so softwarerenderer_b200/libsrb200.so; do
  echo "== $so" | tee -a $O/probe.txt
  for b in "68 69" "0 67" "70 74"; do
    SRB200_LIB=$PWD/$so PROBE=depth python tools/band_probe.py $b 2>/dev/null | tee -a $O/probe.txt
  done
done
