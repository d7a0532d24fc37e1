#!/usr/bin/env python
"""Small workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck; ONE tool per gpurun call):
Gun 640x360 through both pipelines, wireframe, a pipelined burst of alternating scenes and replays, a crowd of tiny
triangles with split bins (several CTAs merging into one key slice), a band context, a near-clip scene, raycasts
(growable and static scratch) and the present paths.  Every image is compared with the oracle."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import softwarerenderer_b200 as srb  # noqa: E402
from oracle import pyoracle as O  # noqa: E402
from softwarerenderer_b200 import _native as N  # noqa: E402
from softwarerenderer_b200 import scene as S  # noqa: E402
from tests.test_gpu_tiny_and_split import _tiny_scene  # noqa: E402
from tests.util import compare  # noqa: E402

W, H = 640, 360
gun = S.scene_model("Gun", W, H)
city = S.scene_city(W, H, n_instances=8)
wire = S.scene_model("Gun", W, H, view=S.gun_camera(3))
wire.debug_mode = 1
tiny = _tiny_scene(5, 2, W=256, H=160, n=2500, straddle=True)
want = {id(s): O.render_scene(s, threads=4)[:2] for s in (gun, city, wire, tiny)}


def check(win, sc):
    compare(win.ColorBuffer, win.DepthBuffer, *want[id(sc)])


win = srb.MainWindow(W, H)
for pipeline in (1, 2):
    win.SetPipeline(pipeline)
    S.submit_scene(win, gun); check(win, gun)
    S.submit_scene(win, city); check(win, city)
win.SetPipeline(0)
S.submit_scene(win, wire); check(win, wire)
for depth in (2, 4):                                   # pipelined bursts, no host synchronisation in between
    win.SetPipelineDepth(depth)
    for k in range(7):
        S.submit_scene(win, gun if k % 2 == 0 else city)
    check(win, gun)
    for _ in range(6):
        win.ReplayFrame()
    check(win, gun)
win.SetBand(5, 14)                                     # a band context
S.submit_scene(win, city)
assert np.array_equal(win.DepthBuffer[80:224], want[id(city)][1][80:224])
out = np.zeros((H, W, 3), np.float32)
win.PresentRgb(out)
out8 = win.PresentRgba8()
win.Close()

win = srb.MainWindow(256, 160)                         # split bins: several CTAs per tile, merged through global atomics
win.SetBinSplit(64, 64, 32)
S.submit_scene(win, tiny); check(win, tiny)
st = win.Stats()
assert st.split_parts > 0, st.split_parts
for _ in range(4):
    win.ReplayFrame()
check(win, tiny)
m = S.load_model("Gun").Meshes[0]
rng = np.random.default_rng(1)
c = np.array(m.SphereBounds.Center)
o = (c + rng.normal(size=(300, 3)) * 3 * m.SphereBounds.Radius).astype(np.float32)
d = (c - o).astype(np.float32)
I = np.eye(4, dtype=np.float32)
a = srb.Physics.RaycastBatch(win, o, d, m.Vertices, m.Indices, I, 0)
b = np.zeros(300, dtype=N.RAY_HIT_DTYPE)
rays = np.ascontiguousarray(np.concatenate([o, d], axis=1))
N.check(win._ctx, N.lib().srb_raycast_static(win._ctx, win.mesh_handle(m.Vertices, m.Indices), N.ptr(np.ascontiguousarray(I.reshape(16))),
                                            N.ptr(rays), 300, 0, N.ptr(b)))
assert a.tobytes() == b.tobytes() and a["hit"].sum() > 50
win.Close()
print("sanitize_case: all images equal to the oracle; split parts", st.split_parts)
