#!/bin/bash
# The GPU test suite + tools/sanitize_case.py against a -DSRB_CHECKED build (device asserts on every computed index).
# compute-sanitizer is closed on this pool (profiles/r02l_compute_sanitizer_closed_on_this_pool.log); this is the
# substitute for memcheck.  build:  make -C softwarerenderer_b200/csrc OUT=../../build_ab/checked.so EXTRA=-DSRB_CHECKED
O=${1:-gpurun_out/checked}
mkdir -p $O
export SRB200_LIB=$PWD/build_ab/checked.so
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_checked.log 2>&1; echo "rc=$?" >> $O/pytest_gpu_checked.log
tail -4 $O/pytest_gpu_checked.log
timeout 600 python tools/sanitize_case.py > $O/sanitize_case_checked.log 2>&1; echo "rc=$?" >> $O/sanitize_case_checked.log
tail -2 $O/sanitize_case_checked.log
