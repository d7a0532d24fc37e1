#!/bin/bash
# Evidence for the head build: smoke, full GPU suite, bench lines, reference arm, then ncu (launch list + full set of the
# tile and setup kernels) of the same frame loop after it ran plainly.
mkdir -p gpurun_out/r02hd
O=gpurun_out/r02hd
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log; tail -2 $O/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python bench.py > $O/bench_city4k.json 2> $O/bench_city4k.err || tail -5 $O/bench_city4k.err
python bench.py --impl reference --steps 20 --warmup 3 > $O/bench_city4k_reference.json 2> $O/ref.err
for wl in gun1080 stress8k stress8k_ext; do
  steps=1000; [ $wl != gun1080 ] && steps=100
  python bench.py --workload $wl --steps $steps > $O/bench_$wl.json 2> $O/bench_$wl.err || tail -5 $O/bench_$wl.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02hd/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if d.get('impl')=='reference':
            print(f.split('/')[-1], 'REFERENCE', round(d['value'],2), d['cpu_baseline']['cores'], d['steps'], d['warmup']); continue
        print(f.split('/')[-1], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'rgba8', round(d['e2e']['rgba8_present']['value'],1), {k: round(v*1000,1) for k,v in d['stage_ms'].items()}, 'lat', round(d['frame_latency']['median_ms']*1000,1), 'frac', round(d['roofline']['frac'],3), 'cpu', d.get('cpu_baseline',{}).get('value'), d['clocks'])
    except Exception as e:
        print(f, 'FAILED', e)
PY
python tools/ncu_frame.py city4k 3 > $O/plain_city4k.log 2>&1 || { echo plain run failed; tail -5 $O/plain_city4k.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_city4k.csv python tools/ncu_frame.py city4k 3 > $O/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"vis_tile_kernel|setup_kernel" -s 4 -c 4 -o $O/full_city4k python tools/ncu_frame.py city4k 3 > $O/ncu3.log 2>&1
ls -la $O | head -30; tail -2 $O/ncu3.log
