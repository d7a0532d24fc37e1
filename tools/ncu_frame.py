#!/usr/bin/env python
"""A few frames of one workload for ncu (no timing here: numbers under a profiler are never bench values).
usage: ncu_frame.py [workload] [replays]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import softwarerenderer_b200 as srb  # noqa: E402
from softwarerenderer_b200 import scene as S  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "city4k"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
scene = bench.build_scene(wl)
win = srb.MainWindow(scene.width, scene.height)
S.submit_scene(win, scene)
win.Sync()
S.submit_scene(win, scene)          # capacities settled
win.Sync()
for _ in range(n):
    win.ReplayFrame()
    win.Sync()
st = win.Stats()
print(wl, "launches per frame", st.kernel_launches, "pairs", st.tile_pairs)
win.Close()
