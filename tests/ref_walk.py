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


def wireframe_depth(tri_screen_submitted, depths_submitted, W, H, depth_test="lequal"):
    """Independent float32 restatement of DebugMode.Wireframe for ONE triangle on a cleared target
    (Rasterizer.cs:419-425 -> three DrawLine calls, Rasterizer.cs:232-340).  `depths_submitted` are
    the per-vertex depths[] values (ndc.z+1)/2 in SUBMISSION order.  Returns (depth[H,W], hits[H,W]):
    the stored depth (clear = float.MinValue) and how many of the three lines touched each pixel.
    Alpha is assumed non-zero (every touched pixel that passes the depth test is written)."""
    t = np.asarray(tri_screen_submitted, dtype=np.float32)
    s = [t[2], t[1], t[0]]                                          # :367 reorder {v2, v1, v0}
    dd = np.asarray(depths_submitted, dtype=np.float32)
    d0, d1 = dd[2], dd[1]                                           # every call reads depths[0], depths[1] (:315)
    depth = np.full((H, W), F(-3.4028235e38), dtype=np.float32)
    hits = np.zeros((H, W), dtype=np.int32)
    for a, b in ((0, 1), (1, 2), (2, 0)):                           # :421-423
        p0, p1 = s[a], s[b]
        minX = int(max(min(p0[0], p1[0]), F(0)))                    # :242-245 (truncation)
        maxX = int(min(max(p0[0], p1[0]), F(W - 1)))
        minY = int(max(min(p0[1], p1[1]), F(0)))
        maxY = int(min(max(p0[1], p1[1]), F(H - 1)))
        if minX > maxX or minY > maxY:
            continue
        dx, dy = F(p1[0] - p0[0]), F(p1[1] - p0[1])
        len_sq = F(F(dx * dx) + F(dy * dy))
        for y in range(minY, maxY + 1):
            for x in range(minX, maxX + 1):
                cx, cy = F(F(x) + F(0.5)), F(F(y) + F(0.5))
                px, py = F(cx - p0[0]), F(cy - p0[1])
                tt = F(0)
                if len_sq > 0:
                    tt = F(F(F(px * dx) + F(py * dy)) / len_sq)
                tt = max(F(0), min(F(1), tt))
                ddx = F(cx - F(p0[0] + F(tt * dx)))
                ddy = F(cy - F(p0[1] + F(tt * dy)))
                if F(F(ddx * ddx) + F(ddy * ddy)) <= F(0.25):
                    hits[y, x] += 1
                    nd = F(F(1) / F(F(d0 * F(F(1) - tt)) + F(d1 * tt)))
                    old = depth[y, x]
                    ok = nd >= old if depth_test == "lequal" else True       # :542-559 (LessEqual is new >= old)
                    if ok:
                        depth[y, x] = nd
    return depth, hits
