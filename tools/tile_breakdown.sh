#!/bin/bash
# Where the time of the visibility tile kernel goes: builds diagnostic variants of libsrb200.so (wrong images,
# never shipped) into build_ab/ and prints fps + stage times of each on the GPU box.
#   here:        bash tools/tile_breakdown.sh build
#   on the box:  gpurun -- 'bash tools/tile_breakdown.sh run'
# full - noshade = shading; noshade - noshade_noatomic = shared-memory atomicMax; noshade_noatomic - noshade_nowalk =
# the coverage walk; noshade_nowalk = bin/record loads, scan, item expansion and the framebuffer stores.
set -e
cd "$(dirname "$0")/.."
if [ "$1" = build ]; then
  mkdir -p build_ab
  for v in "noshade:-DSRB_EXP_SKIP_SHADE" "noshade_noatomic:-DSRB_EXP_SKIP_SHADE -DSRB_EXP_NO_ATOMIC" \
           "noshade_nowalk:-DSRB_EXP_SKIP_SHADE -DSRB_EXP_NO_WALK" "noatomic:-DSRB_EXP_NO_ATOMIC" \
           "allempty:-DSRB_EXP_ALL_EMPTY" "skipempty:-DSRB_EXP_SKIP_EMPTY"; do
    make -s -C softwarerenderer_b200/csrc OUT=../../build_ab/${v%%:*}.so EXTRA="${v#*:}" > /dev/null
  done
  ls build_ab
else
  bash tools/ab_bench.sh softwarerenderer_b200/libsrb200.so build_ab/noshade.so build_ab/noshade_noatomic.so \
       build_ab/noshade_nowalk.so build_ab/noatomic.so build_ab/allempty.so build_ab/skipempty.so
fi
