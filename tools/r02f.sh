#!/bin/bash
mkdir -p gpurun_out/r02f
O=gpurun_out/r02f
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
timeout 300 python tools/band_probe.py 0 0 city4k 2>&1 | tee $O/probe_full.txt
timeout 300 python tools/band_probe.py 67 68 city4k 2>&1 | tee $O/probe_67.txt
timeout 300 python tools/band_probe.py 68 69 city4k 2>&1 | tee $O/probe_68.txt
timeout 300 python tools/band_probe.py 0 67 city4k 2>&1 | tee $O/probe_0_67.txt
timeout 300 python tools/band_probe.py 0 0 gun1080 2>&1 | tee $O/probe_gun.txt
