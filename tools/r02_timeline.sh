#!/bin/bash
mkdir -p gpurun_out/r02tl
O=gpurun_out/r02tl
export SRB200_LIB=$PWD/build_ab/timeline.so
for args in "0 67 city4k 0 1" "68 69 city4k 0 1" "0 0 gun1080 0 1" "0 0 city4k 0 1" "0 67 city4k 0 0"; do
  n=$(echo $args | tr ' ' '_')
  python tools/timeline.py $args > $O/tl_$n.txt 2>&1; tail -1 $O/tl_$n.txt
done
sed -n 20,40p "$O/tl_0_67_city4k_0_1.txt"
