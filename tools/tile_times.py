#!/usr/bin/env python
"""Diagnostic: per-CTA timeline of the visibility tile kernel (needs a -DSRB_TILE_TIMING build,
selected with SRB200_LIB).  Prints how CTA durations distribute over bin length and when the
last CTAs of each class finish."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import softwarerenderer_b200 as srb  # noqa: E402
from softwarerenderer_b200 import _native as N  # noqa: E402
from softwarerenderer_b200 import scene as S  # noqa: E402

# usage: tile_times.py [band_row0 band_row1 [split_min [part_entries]]]
sc = S.scene_city(3840, 2160)
win = srb.MainWindow(sc.width, sc.height)
if len(sys.argv) > 2 and int(sys.argv[2]) > int(sys.argv[1]):
    win.SetBand(int(sys.argv[1]), int(sys.argv[2]))
if len(sys.argv) > 3:
    win.SetBinSplit(int(sys.argv[3]), int(sys.argv[3]), int(sys.argv[4]) if len(sys.argv) > 4 else 128)
S.submit_scene(win, sc)
for _ in range(3):
    win.ReplayFrame()
win.Sync()
n = ((sc.width + 15) // 16) * ((sc.height + 15) // 16) + 2048       # + the extra CTAs of split bins (launched first)
buf = np.zeros((n, 4), dtype=np.uint64)
L = N.lib()
L.srb_debug_tile_times.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
N.check(win._ctx, L.srb_debug_tile_times(win._ctx, N.ptr(buf), n))
buf = buf[buf[:, 0] != 0]                                             # CTAs that ran (band tiles, used extra slots)
n = buf.shape[0]
t0, t1 = buf[:, 0].astype(np.int64), buf[:, 1].astype(np.int64)
cnt = (buf[:, 2] & np.uint64(0xffffffff)).astype(np.int64)
smid = (buf[:, 3] & np.uint64(0xff)).astype(np.int64)
pa = ((buf[:, 3] >> np.uint64(8)) & np.uint64(0xfffff)).astype(np.int64) / 1000.0    # start -> first chunk loaded + scanned
pb = ((buf[:, 3] >> np.uint64(28)) & np.uint64(0xfffff)).astype(np.int64) / 1000.0   # start -> walk done
base = t0.min()
dur = (t1 - t0) / 1000.0
print(f"kernel span {(t1.max() - base) / 1000.0:.1f} us, CTAs {n}, sum of CTA durations {dur.sum() / 1000.0:.1f} ms")
edges = [0, 1, 8, 32, 96, 192, 512, 1024, 1 << 30]
for a, b in zip(edges[:-1], edges[1:]):
    m = (cnt >= a) & (cnt < b)
    if m.any():
        print(f"bin [{a:5d},{b:10d}): tiles {m.sum():6d}  dur mean {dur[m].mean():7.2f} max {dur[m].max():7.2f} us  "
              f"start mean {(t0[m] - base).mean() / 1000:7.1f}  end max {(t1[m] - base).max() / 1000:7.1f} us  "
              f"share of CTA time {dur[m].sum() / dur.sum():5.1%}  "
              f"phases load {pa[m].mean():5.2f} walk {(pb[m] - pa[m]).mean():5.2f} resolve {(dur[m] - pb[m]).mean():5.2f}")
# concurrency over time
ts = np.linspace(0, (t1.max() - base), 30)
act = [int(((t0 - base) <= x).sum() - ((t1 - base) <= x).sum()) for x in ts]
print("active CTAs over time:", act)
busy = np.zeros(smid.max() + 1)
for s_ in range(len(busy)):
    m = smid == s_
    if m.any():
        busy[s_] = (t1[m].max() - t0[m].min()) / 1000.0
print("per-SM span us: min %.1f mean %.1f max %.1f" % (busy[busy > 0].min(), busy[busy > 0].mean(), busy.max()))
win.Close()
