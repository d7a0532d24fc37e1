#!/bin/bash
# 8 GPUs: multi-GPU parity test, city4k N=8 / N=4 / N=2, stress8k + stress8k_ext N=8, NUMA binding off for comparison
mkdir -p gpurun_out/r02o
O=gpurun_out/r02o
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_bands_fused.py -m gpu -x -q > $O/pytest_multi.log 2>&1; echo "rc=$?" >> $O/pytest_multi.log
tail -3 $O/pytest_multi.log
runN() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $2 --warmup 5 $3 > $O/bench_$1.json 2> $O/bench_$1.err || tail -8 $O/bench_$1.err; }
runN city4k_n8 8 "--steps 300 --e2e-steps 60"
runN city4k_n8_nonuma 8 "--steps 100 --e2e-steps 60 --no-numa-bind --no-other-gather"
runN city4k_n4 4 "--steps 300 --e2e-steps 60"
runN city4k_n2 2 "--steps 300 --e2e-steps 60"
runN stress8k_n8 8 "--workload stress8k --steps 60 --e2e-steps 20"
runN stress8k_ext_n8 8 "--workload stress8k_ext --steps 60 --e2e-steps 20"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02o/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        o=d.get('value_other_gather') or {}
        print(f.split('/')[-1], 'value', round(d['value'],1), 'other', o.get('format'), round(o.get('value',0),1), o.get('verified'), 'e2e', round(d['e2e']['value'],1), d['e2e'].get('rows_verified'), d['e2e'].get('numa_bound_cpus'), 'rgba8', round((d.get('e2e_via_gpu0_gather') or d['e2e'])['rgba8_present']['value'],1), 'verified', d['config_detail']['bands']['gather_verified'])
        print('   cuts', d['config_detail']['bands']['cuts_tile_rows'], 'e2e cuts', d['e2e'].get('cuts_tile_rows'))
    except Exception as e:
        print(f, 'FAILED', e)
PY
