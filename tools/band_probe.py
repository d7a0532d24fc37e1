#!/usr/bin/env python
"""One GPU, one band of the city4k frame (what a band rank of an N-GPU run owns): per-stage device times and the
replay rate with bin splitting off / on.  usage: band_probe.py r0 r1 [workload]"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import softwarerenderer_b200 as srb  # noqa: E402
from softwarerenderer_b200 import _native as N  # noqa: E402
from softwarerenderer_b200 import scene as S  # noqa: E402

r0, r1 = int(sys.argv[1]), int(sys.argv[2])
wl = sys.argv[3] if len(sys.argv) > 3 else "city4k"
scene = bench.build_scene(wl)
L = N.lib()
VARIANTS = {"split": (((0, 0, 64), 0, 1), ((128, 128, 64), 0, 1), ((256, 256, 128), 0, 1), ((512, 512, 128), 0, 1)),
            "depth": (((256, 256, 128), 0, 1), ((256, 256, 128), 2, 1), ((256, 256, 128), 3, 1), ((256, 256, 128), 4, 1)),
            "default": (((256, 256, 128), 0, 1),)}
for split, depth, graphs in VARIANTS[os.environ.get("PROBE", "default")]:
    win = srb.MainWindow(scene.width, scene.height)
    ctx = win._ctx
    win.SetBinSplit(*split)
    win.SetPipelineDepth(depth)
    win.SetGraphs(bool(graphs))
    if r1 > r0:
        win.SetBand(r0, r1)
    S.submit_scene(win, scene)
    S.submit_scene(win, scene)
    st = win.Stats()
    for _ in range(20):
        N.check(ctx, L.srb_replay_frame(ctx))
    win.Sync()
    th = time.perf_counter()
    for _ in range(24):                       # host cost of queueing a frame (nothing to wait for: the queue is empty)
        N.check(ctx, L.srb_replay_frame(ctx))
    host_us = (time.perf_counter() - th) / 24 * 1e6
    win.Sync()
    t0 = time.perf_counter()
    K = 300
    for _ in range(K):
        N.check(ctx, L.srb_replay_frame(ctx))
    win.Sync()
    fps = K / (time.perf_counter() - t0)
    N.check(ctx, L.srb_set_profiling(ctx, 1))
    acc = np.zeros(7)
    buf = (C.c_float * 7)()
    for i in range(33):
        N.check(ctx, L.srb_replay_frame(ctx))
        N.check(ctx, L.srb_get_stage_ms(ctx, buf, 7))
        if i >= 3:
            acc += np.array(list(buf))
    N.check(ctx, L.srb_set_profiling(ctx, 0))
    print(f"band {r0}-{r1} split {split} depth {depth} graphs {graphs}: {fps:8.0f} frames/s  parts {st.split_parts:5d}  pairs {st.tile_pairs}  host {host_us:.1f} us/frame  "
          + " ".join(f"{n}={a / 30 * 1000:.1f}" for n, a in zip(bench.STAGES, acc)), flush=True)
    win.Close()
