#!/usr/bin/env python
"""Where the frames of a replay loop sit in time: geometry chain start/end and tile kernel start/end of every frame,
stamped on the device (diagnostic build: make -C softwarerenderer_b200/csrc OUT=../../build_ab/timeline.so
EXTRA=-DSRB_TIMELINE, run with SRB200_LIB=build_ab/timeline.so).
usage: timeline.py r0 r1 [workload] [depth] [graphs]   (r0 == r1 == 0: full frame)"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import softwarerenderer_b200 as srb  # noqa: E402
from softwarerenderer_b200 import _native as N  # noqa: E402
from softwarerenderer_b200 import scene as S  # noqa: E402

r0, r1 = int(sys.argv[1]), int(sys.argv[2])
wl = sys.argv[3] if len(sys.argv) > 3 else "city4k"
depth = int(sys.argv[4]) if len(sys.argv) > 4 else 0
graphs = int(sys.argv[5]) if len(sys.argv) > 5 else 1
scene = bench.build_scene(wl)
L = N.lib()
L.srb_debug_timeline.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
win = srb.MainWindow(scene.width, scene.height)
ctx = win._ctx
win.SetPipelineDepth(depth)
win.SetGraphs(bool(graphs))
if r1 > r0:
    win.SetBand(r0, r1)
S.submit_scene(win, scene)
S.submit_scene(win, scene)
for _ in range(200):
    N.check(ctx, L.srb_replay_frame(ctx))
win.Sync()
K = 48
out = np.zeros((K, 4), dtype=np.uint64)
n = L.srb_debug_timeline(ctx, out.ctypes.data_as(C.c_void_p), K)
t = out[:n].astype(np.int64)
t0 = t[0, 0]
print(f"# {wl} rows {r0}-{r1} depth {depth} graphs {graphs}: frame, geometry start..end, tile start..end (us since the first), "
      "geometry us, tile us, tile start - previous tile end")
for i in range(n):
    g0, g1, a, b = [(x - t0) / 1000.0 for x in t[i]]
    gap = (t[i, 2] - t[i - 1, 3]) / 1000.0 if i else 0.0
    print(f"{i:3d}  G {g0:8.1f} .. {g1:8.1f}   T {a:8.1f} .. {b:8.1f}   geom {g1 - g0:6.1f}  tile {b - a:6.1f}  gap {gap:6.1f}")
per = (t[-1, 3] - t[8, 3]) / 1000.0 / (n - 9)
print(f"# period {per:.1f} us; mean geometry {np.mean(t[8:, 1] - t[8:, 0]) / 1000:.1f} us, mean tile {np.mean(t[8:, 3] - t[8:, 2]) / 1000:.1f} us, "
      f"mean gap between tile kernels {np.mean(t[9:, 2] - t[8:-1, 3]) / 1000:.1f} us")
win.Close()
