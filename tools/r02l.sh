#!/bin/bash
mkdir -p gpurun_out/r02l
O=gpurun_out/r02l
for wl in city4k gun1080; do
  for so in softwarerenderer_b200/libsrb200.so build_ab/unroll.so build_ab/unroll4.so; do
    SRB200_LIB=$PWD/$so timeout 300 python bench.py --steps 200 --e2e-steps 5 --no-cpu-baseline --workload $wl 2>$O/err.txt | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl $so', round(d['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1))" | tee -a $O/ab.txt
  done
done
timeout 600 python tools/sanitize_case.py > $O/sanitize_plain.log 2>&1 && \
timeout 1500 compute-sanitizer --tool memcheck --leak-check no --error-exitcode 7 python tools/sanitize_case.py > $O/memcheck.log 2>&1; echo "memcheck rc=$?" | tee -a $O/memcheck.log
tail -3 $O/sanitize_plain.log; tail -6 $O/memcheck.log
