#!/bin/bash
mkdir -p gpurun_out/r02c
O=gpurun_out/r02c
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
export SRB200_LIB=$PWD/build_ab/timing.so
for args in "67 68 0" "67 68 192" "68 70 0" "68 70 192" "0 0 0"; do
  echo "== tile_times $args"; timeout 200 python tools/tile_times.py $args 2>&1 | tee -a $O/tile_times.txt
done
