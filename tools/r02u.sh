#!/bin/bash
mkdir -p gpurun_out/r02u
O=gpurun_out/r02u
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
for so in build_ab/head.so softwarerenderer_b200/libsrb200.so; do
  echo "== $so" | tee -a $O/probe.txt
  SRB200_LIB=$PWD/$so timeout 300 python tools/band_probe.py 69 97 stress8k 2>&1 | tee -a $O/probe.txt
  SRB200_LIB=$PWD/$so timeout 300 python tools/band_probe.py 68 69 city4k 2>&1 | tee -a $O/probe.txt
  SRB200_LIB=$PWD/$so timeout 300 python tools/band_probe.py 88 111 city4k 2>&1 | tee -a $O/probe.txt
done
