"""Shared parity checks: GPU framebuffer vs oracle framebuffer."""
import numpy as np

CLEAR_DEPTH = np.float32(-3.4028235e38)


def to_u8(c):
    return (np.clip(np.nan_to_num(c, nan=0.0), 0.0, 1.0) * 255.0 + 0.5).astype(np.int32)


def compare(gc, gd, oc, od, min_frac=0.999):
    """North-star gate: coverage mask and depth winner bit-exact; RGBA8 within +-1 LSB
    on >= 99.9 % of pixels.  Returns a dict of statistics (asserts on failure)."""
    cov_g, cov_o = gd != CLEAR_DEPTH, od != CLEAR_DEPTH
    n_cov_diff = int((cov_g != cov_o).sum())
    n_depth_diff = int((gd != od).sum())
    d8 = np.abs(to_u8(gc) - to_u8(oc)).max(axis=-1)
    frac_ok = float((d8 <= 1).mean())
    stats = dict(coverage=float(cov_o.mean()), coverage_mismatch=n_cov_diff, depth_mismatch=n_depth_diff,
                 rgba8_within1=frac_ok, rgba8_max_err=int(d8.max()),
                 float_max_err=float(np.nanmax(np.abs(gc - oc))), float_exact=bool(np.array_equal(gc, oc)))
    assert n_cov_diff == 0, stats
    assert n_depth_diff == 0, stats
    assert frac_ok >= min_frac, stats
    return stats


# ---------------------------------------------------------------------------
# Tiny hand-built scenes: triangles given directly in SCREEN pixels.  With
# model = view = proj = identity the stock vertex shader leaves clip = (x,y,z,1),
# so sx = (x*0.5+0.5)*W and sy = (1-(y*0.5+0.5))*H (Rasterizer.cs:383-386) invert
# exactly for power-of-two W,H and coordinates that are multiples of 1/4.
# ---------------------------------------------------------------------------
def screen_tri_scene(W, H, tris, colors=None, uvs=None, ndc_z=0.5, cull=0, depth=2, blend=1, texture=None,
                     separate_draws=False):
    """tris: list of 3x2 screen-pixel coordinates (or 3x3 with a per-vertex NDC z).
    colors: per-triangle RGBA (flat shaded) or per-vertex 3x4.  Returns a Scene."""
    from softwarerenderer_b200 import numerics as nm
    from softwarerenderer_b200.scene import DrawItem, Mesh, Scene
    from softwarerenderer_b200.shaders import Material, Shaders
    I = nm.identity()
    pos, col, uv = [], [], []
    for k, t in enumerate(tris):
        t = np.asarray(t, dtype=np.float64)
        for j in range(3):
            sx, sy = t[j, 0], t[j, 1]
            z = t[j, 2] if t.shape[1] > 2 else (ndc_z[k] if np.ndim(ndc_z) else ndc_z)
            pos.append((2.0 * sx / W - 1.0, 1.0 - 2.0 * sy / H, z))
            c = (1, 1, 1, 1) if colors is None else colors[k]
            c = np.asarray(c, dtype=np.float32)
            col.append(c[j] if c.ndim == 2 else c)
            uv.append((0.0, 0.0) if uvs is None else uvs[k][j])
    n = len(tris)
    draws = []
    groups = [[k] for k in range(n)] if separate_draws else [list(range(n))]
    for g in groups:
        sel = [3 * k + j for k in g for j in range(3)]
        v = Shaders.make_vertices(np.asarray(pos)[sel], uv=np.asarray(uv)[sel], color=np.asarray(col)[sel])
        idx = np.arange(len(sel), dtype=np.uint16)
        draws.append(DrawItem(Mesh(v, idx, Material(), None, texture), I, I, I, cull=cull, depth=depth, blend=blend))
    # fog_start > every clip z used here -> fog factor clamps to 1 -> no fog unless a test sets it
    return Scene(W, H, draws, clear_color=(0.0, 0.0, 0.0, 0.0), fog_start=10.0, fog_end=20.0)
