#!/bin/bash
mkdir -p gpurun_out/r02z
O=gpurun_out/r02z
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tiny_and_split.py -m gpu -x -q 2>&1 | tail -2
for band in "68 69" "68 70" "67 68" "70 74"; do timeout 300 python tools/band_probe.py $band city4k 2>&1 | tee -a $O/probe.txt; done
timeout 300 python tools/band_probe.py 0 0 city4k 2>&1 | tee -a $O/probe.txt
