#!/bin/bash
# ncu: launch list + full set of the dominant kernels, city4k and stress8k (all ncu runs of a call count as one)
mkdir -p gpurun_out/r02q
O=gpurun_out/r02q
python tools/ncu_frame.py city4k 3 > $O/plain_city4k.log 2>&1 && python tools/ncu_frame.py stress8k 2 > $O/plain_stress8k.log 2>&1 || { echo plain run failed; tail -5 $O/plain_*.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_city4k.csv python tools/ncu_frame.py city4k 3 > $O/ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_stress8k.csv python tools/ncu_frame.py stress8k 2 > $O/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"vis_tile_kernel|setup_kernel" -s 4 -c 4 -o $O/full_city4k python tools/ncu_frame.py city4k 3 > $O/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"vis_tile_kernel|setup_kernel" -s 4 -c 2 -o $O/full_stress8k python tools/ncu_frame.py stress8k 2 > $O/ncu4.log 2>&1
ls -la $O; tail -3 $O/ncu3.log
