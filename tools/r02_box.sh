#!/bin/bash
# Block culling by boxes instead of spheres: parity (band tests, edges, the tests the last full run did not reach), A/B per band.
mkdir -p gpurun_out/r02bx
O=gpurun_out/r02bx
timeout 900 python -m pytest tests/test_gpu_bands_fused.py tests/test_gpu_edges.py tests/test_gpu_tiny_and_split.py tests/test_ingest.py tests/test_raycast.py tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
SRB200_LIB=$PWD/build_ab/checked.so timeout 600 python -m pytest tests/test_gpu_bands_fused.py tests/test_gpu_edges.py -m gpu -x -q > $O/pytest_checked.log 2>&1; echo "rc=$?" >> $O/pytest_checked.log
tail -2 $O/pytest_checked.log
for so in build_ab/prev.so softwarerenderer_b200/libsrb200.so; do
  echo "== $so" | tee -a $O/probe.txt
  for b in "0 67" "67 68" "68 69" "69 70" "70 74" "74 88" "88 111"; do
    SRB200_LIB=$PWD/$so python tools/band_probe.py $b 2>/dev/null | tee -a $O/probe.txt
  done
  SRB200_LIB=$PWD/$so python tools/band_probe.py 0 135 stress8k 2>/dev/null | tee -a $O/probe.txt
  SRB200_LIB=$PWD/$so python tools/band_probe.py 135 270 stress8k 2>/dev/null | tee -a $O/probe.txt
done
