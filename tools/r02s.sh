#!/bin/bash
mkdir -p gpurun_out/r02s
O=gpurun_out/r02s
timeout 900 python -m pytest tests/test_ingest.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3 | tee $O/pytest_part.log
for t in 1 8 16; do tools/micro/per_draw_submit $t | tee -a $O/per_draw_submit.txt; done
ORC_TRACE=1 python bench.py --impl reference --steps 6 --warmup 2 > $O/ref_city4k.json 2> $O/ref_city4k.err; tail -3 $O/ref_city4k.err; cut -c1-260 $O/ref_city4k.json
python bench.py --impl reference --workload gun1080 --steps 20 --warmup 3 | cut -c1-200
python bench.py --impl reference --workload stress8k --steps 2 --warmup 1 | cut -c1-200
nproc
