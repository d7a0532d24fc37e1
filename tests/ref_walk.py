"""Independent pure-Python/NumPy-float32 restatement of the coverage walk of
Rasterizer.RasterizeTriangle (Rasterizer.cs:427-535), for tiny cases only.

Used to pin the C++ oracle: same semantics written a second time, in a different
language, straight from the reference lines.  Input is a triangle already in
SCREEN space in the rasterizer's working order (s0, s1, s2 = outputs {v2, v1, v0}).
"""
import math

import numpy as np

F = np.float32
TILE = 16


def coverage_by_tile_walk(tri_screen_submitted, W, H):
    """tri_screen_submitted: 3x2 screen coordinates in SUBMISSION order (v0, v1, v2).
    Returns the boolean coverage mask (H, W)."""
    t = np.asarray(tri_screen_submitted, dtype=np.float32)
    s0, s1, s2 = t[2], t[1], t[0]                                  # :367 reorder {v2, v1, v0}
    cov = np.zeros((H, W), dtype=bool)
    area = (s2[0] - s0[0]) * (s1[1] - s0[1]) - (s2[1] - s0[1]) * (s1[0] - s0[0])      # :561-563 on (s0,s1,s2)
    if area == 0:
        return cov
    minX = max(int(math.floor(min(s0[0], s1[0], s2[0]))), 0)       # :432-442
    maxX = min(int(math.ceil(max(s0[0], s1[0], s2[0]))), W - 1)
    minY = max(int(math.floor(min(s0[1], s1[1], s2[1]))), 0)
    maxY = min(int(math.ceil(max(s0[1], s1[1], s2[1]))), H - 1)
    if minX > maxX or minY > maxY:
        return cov
    a01, b01 = s0[1] - s1[1], s1[0] - s0[0]                        # :445-447
    a12, b12 = s1[1] - s2[1], s2[0] - s1[0]
    a20, b20 = s2[1] - s0[1], s0[0] - s2[0]
    for ty in range(minY // TILE, maxY // TILE + 1):               # :449-462
        for tx in range(minX // TILE, maxX // TILE + 1):
            x0, x1 = max(minX, tx * TILE), min(maxX, tx * TILE + TILE - 1)
            y0, y1 = max(minY, ty * TILE), min(maxY, ty * TILE + TILE - 1)
            fx, fy = F(x0), F(y0)
            w0r = a12 * (fx - s1[0]) + b12 * (fy - s1[1])          # :481-483
            w1r = a20 * (fx - s2[0]) + b20 * (fy - s2[1])
            w2r = a01 * (fx - s0[0]) + b01 * (fy - s0[1])
            for y in range(y0, y1 + 1):
                w0, w1, w2 = w0r, w1r, w2r
                for x in range(x0, x1 + 1):
                    if (w0 >= 0 and w1 >= 0 and w2 >= 0) or (w0 <= 0 and w1 <= 0 and w2 <= 0):   # :493-494
                        cov[y, x] = True
                    w0, w1, w2 = F(w0 + a12), F(w1 + a20), F(w2 + a01)                           # :527-529
                w0r, w1r, w2r = F(w0r + b12), F(w1r + b20), F(w2r + b01)                         # :532-534
    return cov
