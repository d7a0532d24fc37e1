#!/bin/bash
mkdir -p gpurun_out/r02n
O=gpurun_out/r02n
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
bash tools/checked_suite.sh $O
