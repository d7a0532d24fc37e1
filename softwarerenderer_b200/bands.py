"""Sort-first screen-space partition for N GPUs.

Bands are cut on 16-row tile boundaries -- mandatory for bit-exactness, because the
reference seeds its edge accumulators at max(minY, tileStartY) (Rasterizer.cs:466-483);
a cut inside a tile row would change the rounding of every pixel below it.  Each rank
renders its band with the whole pipeline (geometry replicated) and the bands are
gathered into rank 0's framebuffer: the only exchange step of the path.
"""
from __future__ import annotations

TILE = 16


def band_rows(tiles_y: int, n: int, r: int):
    """Tile-row range [r0, r1) of rank r of n, balanced by tile rows."""
    if n <= 0 or not 0 <= r < n:
        raise ValueError("band_rows: need 0 <= r < n")
    base, rem = divmod(tiles_y, n)
    r0 = r * base + min(r, rem)
    return r0, r0 + base + (1 if r < rem else 0)


def band_pixels(height: int, n: int, r: int):
    """Pixel-row range [y0, y1) of rank r of n for a target `height` rows tall."""
    tiles_y = (height + TILE - 1) // TILE
    r0, r1 = band_rows(tiles_y, n, r)
    return r0 * TILE, min(r1 * TILE, height)


def gather_bands(fb, height: int, rank: int, world: int, dist=None):
    """Gather every rank's band of `fb` ([H, W, C] tensor, full-size on every rank) into
    rank 0's `fb` with grouped point-to-point sends (bands differ in size, so this is a
    gather-v, not an all-gather).  Works on NCCL (device tensors, over NVLink) and gloo."""
    if world == 1:
        return
    if dist is None:
        import torch.distributed as dist
    ops = []
    if rank == 0:
        for r in range(1, world):
            a, b = band_pixels(height, world, r)
            if b > a:
                ops.append(dist.P2POp(dist.irecv, fb[a:b], r))
    else:
        a, b = band_pixels(height, world, rank)
        if b > a:
            ops.append(dist.P2POp(dist.isend, fb[a:b], 0))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
