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
