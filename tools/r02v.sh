#!/bin/bash
mkdir -p gpurun_out/r02v
O=gpurun_out/r02v
for lean in 0 1; do
  echo "== lean $lean" | tee -a $O/probe.txt
  for band in "68 69" "0 67" "88 111" "70 74"; do
    SRB_EXP_LEAN=$lean timeout 300 python tools/band_probe.py $band city4k 2>&1 | tee -a $O/probe.txt
  done
  SRB_EXP_LEAN=$lean timeout 300 python tools/band_probe.py 69 97 stress8k 2>&1 | tee -a $O/probe.txt
done
for lean in 0 2; do
  echo "== full frame lean $lean" | tee -a $O/probe.txt
  SRB_EXP_LEAN=$lean timeout 300 python tools/band_probe.py 0 0 city4k 2>&1 | tee -a $O/probe.txt
  SRB_EXP_LEAN=$lean timeout 300 python tools/band_probe.py 0 0 gun1080 2>&1 | tee -a $O/probe.txt
done
