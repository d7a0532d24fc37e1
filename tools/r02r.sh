#!/bin/bash
mkdir -p gpurun_out/r02r
O=gpurun_out/r02r
for rep in 1 2; do
for so in build_ab/head.so softwarerenderer_b200/libsrb200.so build_ab/precheck.so; do
  for wl in city4k; do
    SRB200_LIB=$PWD/$so timeout 300 python bench.py --steps 300 --e2e-steps 5 --no-cpu-baseline --workload $wl 2>$O/err.txt | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/ab.txt
  done
done
done
SRB200_LIB=$PWD/build_ab/precheck.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tiny_and_split.py -m gpu -x -q 2>&1 | tail -2
for so in softwarerenderer_b200/libsrb200.so build_ab/precheck.so; do
SRB200_LIB=$PWD/$so timeout 300 python bench.py --steps 40 --e2e-steps 5 --no-cpu-baseline --workload stress8k 2>$O/err.txt | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('stress8k $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()})" | tee -a $O/ab.txt
done
