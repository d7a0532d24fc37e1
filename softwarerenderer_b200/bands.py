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


def _cuts_for_limit(w, n, limit, remote_extra):
    """Greedy fill: band k takes rows while its cost stays <= limit (rows of bands k >= 1 cost
    remote_extra more).  Returns the cuts, or None when the rows do not fit into n bands."""
    rows, cuts, r = len(w), [0], 0
    for k in range(n):
        extra = remote_extra if k > 0 else 0.0
        acc, first = 0.0, r
        last_allowed = rows - (n - 1 - k)                 # leave a row for each remaining band
        while r < last_allowed and (r == first or acc + w[r] + extra <= limit):
            acc += w[r] + extra
            r += 1
        cuts.append(r)
    return cuts if r == rows else None


def balanced_cuts(row_weights, n: int, remote_extra: float = 0.0):
    """Cut points [c_0=0, c_1, ..., c_n=rows] over tile rows such that every band carries about
    the same weight (the cost model of the caller: e.g. bin pairs + a constant per tile).  Every
    band gets at least one row when rows >= n.  Deterministic: all ranks compute the same cuts.
    remote_extra > 0: every row of bands 1..n-1 costs that much more (their pixels travel to
    rank 0 over NVLink); the cuts then minimise the largest band cost, so band 0 grows."""
    w = [float(x) for x in row_weights]
    rows = len(w)
    if n <= 0:
        raise ValueError("balanced_cuts: n must be positive")
    total = sum(w)
    if total <= 0.0 or rows < n:
        return [band_rows(rows, n, r)[0] for r in range(n)] + [rows]
    if remote_extra > 0.0 and n > 1:
        lo, hi = 0.0, total + remote_extra * rows
        best = _cuts_for_limit(w, n, hi, remote_extra)
        for _ in range(48):
            mid = 0.5 * (lo + hi)
            c = _cuts_for_limit(w, n, mid, remote_extra)
            if c is None:
                lo = mid
            else:
                hi, best = mid, c
        return best
    cuts, acc, r = [0], 0.0, 0
    for k in range(1, n):
        target = total * k / n
        while r < rows and (acc + w[r] * 0.5 <= target or r < cuts[-1] + 1):
            acc += w[r]
            r += 1
        r = min(r, rows - (n - k))            # leave a row for each remaining band
        acc = sum(w[:r])
        cuts.append(r)
    cuts.append(rows)
    return cuts


def rebalance_from_periods(cuts, periods, row_weights):
    """One round of measured re-balancing.  `periods[r]` is the time rank r needed for its band `cuts[r]..cuts[r+1]`
    (any unit); it is spread over the band's rows in proportion to the model's `row_weights`, and the cuts are balanced
    again on those measured row costs.  Deterministic: every rank computes the same cuts from the same gathered
    periods.  Returns the new cuts (possibly the old ones: a fixed point)."""
    n = len(periods)
    if len(cuts) != n + 1:
        raise ValueError("rebalance_from_periods: need one period per band")
    w = [float(x) for x in row_weights]
    if cuts[0] != 0 or cuts[-1] != len(w) or any(cuts[k + 1] <= cuts[k] for k in range(n)):
        raise ValueError("rebalance_from_periods: cuts must be increasing from 0 to the number of rows")
    cost = [0.0] * len(w)
    for r in range(n):
        a, b = cuts[r], cuts[r + 1]
        tot = sum(w[a:b])
        for i in range(a, b):
            cost[i] = float(periods[r]) * (w[i] / tot if tot > 0.0 else 1.0 / (b - a))
    return balanced_cuts(cost, n)


def relieve_worst_band(cuts, periods):
    """Local search step after the model-based rounds: candidate cuts that take ONE tile row off the band with the
    longest measured period, handing it to a neighbour -- the faster neighbour first.  Returns a list of 0..2 candidate
    cut lists (empty when the worst band is a single row: a tile row cannot be divided).  Deterministic."""
    n = len(periods)
    if len(cuts) != n + 1:
        raise ValueError("relieve_worst_band: need one period per band")
    r = max(range(n), key=lambda k: (periods[k], -k))
    if cuts[r + 1] - cuts[r] <= 1:
        return []
    cands = []
    if r > 0:
        c = list(cuts); c[r] += 1                      # first row to the band above
        cands.append((periods[r - 1], c))
    if r < n - 1:
        c = list(cuts); c[r + 1] -= 1                  # last row to the band below
        cands.append((periods[r + 1], c))
    cands.sort(key=lambda t: t[0])
    return [c for _, c in cands]


def band_pixels(height: int, n: int, r: int):
    """Pixel-row range [y0, y1) of rank r of n for a target `height` rows tall."""
    tiles_y = (height + TILE - 1) // TILE
    r0, r1 = band_rows(tiles_y, n, r)
    return r0 * TILE, min(r1 * TILE, height)


def cut_pixels(height: int, cuts, r: int):
    """Pixel-row range [y0, y1) of band r for explicit tile-row cuts (balanced_cuts)."""
    return cuts[r] * TILE, min(cuts[r + 1] * TILE, height)


def gather_bands(fb, height: int, rank: int, world: int, dist=None, cuts=None):
    """Gather every rank's band of `fb` ([H, W, C] tensor, full-size on every rank) into
    rank 0's `fb` with grouped point-to-point sends (bands differ in size, so this is a
    gather-v, not an all-gather).  Works on NCCL (device tensors, over NVLink) and gloo.
    `cuts`: explicit tile-row cuts (default: the equal split of band_rows)."""
    if world == 1:
        return
    if dist is None:
        import torch.distributed as dist
    span = (lambda r: cut_pixels(height, cuts, r)) if cuts is not None else (lambda r: band_pixels(height, world, r))
    ops = []
    if rank == 0:
        for r in range(1, world):
            a, b = span(r)
            if b > a:
                ops.append(dist.P2POp(dist.irecv, fb[a:b], r))
    else:
        a, b = span(rank)
        if b > a:
            ops.append(dist.P2POp(dist.isend, fb[a:b], 0))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
