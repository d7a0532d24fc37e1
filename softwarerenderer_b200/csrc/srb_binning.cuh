// srb_binning.cuh -- triangle-setup kernel and the count / scan / fill binning
// passes.  Included by srb200.cu only.
#pragma once
#include "srb_geometry.cuh"

namespace srb {

// Tile lists are built in three passes: setup counts (RED atomics), one CTA
// scans, fill scatters.  Distant instances put thousands of consecutive
// triangles into the same tile, so per-tile atomics are warp-aggregated with
// match.any: one atomic per distinct tile per warp instead of one per lane.
constexpr int kSmallTiles = 4;     // triangles touching <= this many tiles are walked by their own lane

// k-th tile of a packed range (row-major), or -1.
__device__ __forceinline__ int small_tile(uint32_t tx, uint32_t ty, int n, int k, int tilesX) {
    if (k >= n) return -1;
    int x0 = tx & 0xffff, nx = (int)(tx >> 16) - x0 + 1, y0 = ty & 0xffff;
    int yy = k / nx, xx = k - yy * nx;
    return (y0 + yy) * tilesX + x0 + xx;
}

// Visit every tile of the accepted triangles held by the lanes of one warp.
// MODE 0 = count (tileCount += 1), MODE 1 = fill (bins[cursor++] = seq).
// All 32 lanes must call this.
template <int MODE>
__device__ __forceinline__ void warp_bin_tiles(bool valid, uint32_t seq, uint32_t tx, uint32_t ty, int tilesX,
                                               uint32_t* __restrict__ tileCounter, uint32_t* __restrict__ bins,
                                               uint32_t cap) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int x0 = tx & 0xffff, x1 = tx >> 16, y0 = ty & 0xffff, y1 = ty >> 16;
    const int nx = x1 - x0 + 1, n = valid ? nx * (y1 - y0 + 1) : 0;
    const int nSmall = n <= kSmallTiles ? n : 0;
    int maxSmall = nSmall;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) maxSmall = max(maxSmall, __shfl_xor_sync(full, maxSmall, o));
    for (int k = 0; k < maxSmall; ++k) {
        const int tile = small_tile(tx, ty, nSmall, k, tilesX);
        const unsigned grp = __match_any_sync(full, tile);
        if (tile >= 0) {
            const int leader = __ffs(grp) - 1;
            const int cnt = __popc(grp);
            if (MODE == 0) {
                if (lane == leader) atomicAdd(tileCounter + tile, (uint32_t)cnt);
            } else {
                uint32_t base = 0;
                if (lane == leader) base = atomicAdd(tileCounter + tile, (uint32_t)cnt);
                base = __shfl_sync(grp, base, leader);
                const uint32_t slot = base + (uint32_t)__popc(grp & ((1u << lane) - 1u));
                if (slot < cap) bins[slot] = seq;
            }
        }
    }
    // Larger triangles: the whole warp walks one triangle's tiles (distinct
    // addresses per lane), four atomics in flight per lane.
    unsigned big = __ballot_sync(full, n > kSmallTiles);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const int bx0 = __shfl_sync(full, x0, src), by0 = __shfl_sync(full, y0, src);
        const int bnx = __shfl_sync(full, nx, src), bn = __shfl_sync(full, n, src);
        const uint32_t bseq = __shfl_sync(full, seq, src);
        for (int i0 = lane; i0 < bn; i0 += 128) {
            int tile[4];
            uint32_t slot[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + 32 * u;
                const int yy = i / bnx, xx = i - yy * bnx;
                tile[u] = i < bn ? (by0 + yy) * tilesX + bx0 + xx : -1;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (tile[u] >= 0) {
                    if (MODE == 0) atomicAdd(tileCounter + tile[u], 1u);
                    else slot[u] = atomicAdd(tileCounter + tile[u], 1u);
                }
            if (MODE == 1) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (tile[u] >= 0 && slot[u] < cap) bins[slot[u]] = bseq;
            }
        }
    }
}

// ---------------------------------------------------------------------------
// Setup kernel: one thread per submitted triangle, one CTA per block-table
// entry (a run of <= kGeomBlock triangles of ONE draw, so the draw lookup is a
// single load and a culled draw retires whole CTAs at once).  Body of
// RenderMesh's Parallel.For lambda (Rasterizer.cs:200-229) with the vertex
// shader hoisted into vertex_kernel: behind tests, optional near clip,
// DrawTriangle, cull, bbox; then per-tile counting for the bins.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kGeomBlock)
setup_kernel(const DrawRec* __restrict__ draws, const int2* __restrict__ tblocks, FrameParams fp,
             SrbUniforms U, const float4* __restrict__ clipIn, const float4* __restrict__ wnIn,
             RecStore recs, ClipAttr* __restrict__ clipAttrs, TriRef* __restrict__ refs,
             uint32_t* __restrict__ tileCount, Counters* cnt) {
    const unsigned full = 0xffffffffu;
    const int2 blk = __ldg(tblocks + blockIdx.x);
    const int di = blk.x;
    const DrawRec& d = draws[di];
    if (!d.visible) return;                                   // uniform over the CTA
    const int t = blk.y + threadIdx.x;
    SetupOut out;
    out.n = 0;
    out.seq[0] = out.seq[1] = 0; out.tx[0] = out.tx[1] = 0; out.ty[0] = out.ty[1] = 0;
    if (t < d.n_triangles) {
        const int g = d.tbase + t;
        int i3[3] = {(int)__ldg(d.idx + 3 * t), (int)__ldg(d.idx + 3 * t + 1), (int)__ldg(d.idx + 3 * t + 2)};
        float4 c0 = clipIn[d.vbase + i3[0]], c1 = clipIn[d.vbase + i3[1]], c2 = clipIn[d.vbase + i3[2]];
        bool b0 = c0.w <= 0.0f, b1 = c1.w <= 0.0f, b2 = c2.w <= 0.0f;     // :208-210
        if (b0 && b1 && b2) {
            // all behind: dropped (:212-215)
        } else if (b0 || b1 || b2) {
            setup_clipped(d, di, d.vbase, i3, clipIn, wnIn, fp, U, (uint32_t)g, recs, clipAttrs, cnt, out);
        } else {
            float4 c[3] = {c2, c1, c0};                                    // :367
            TriRec r;
            int minX, maxX, minY, maxY;
            if (setup_tri(c, fp.W, fp.H, d.cull, r, minX, maxX, minY, maxY)) {
                int ty0 = max(minY / kTile, fp.bandRow0), ty1 = min(maxY / kTile, fp.bandRow1 - 1);
                if (ty0 <= ty1) {
                    r.meta = (uint32_t)di;
                    uint32_t seq = 2u * (uint32_t)g;
                    store_rec(recs, seq, r);
                    out.n = 1;
                    out.seq[0] = seq;
                    out.tx[0] = (uint32_t)(minX / kTile) | ((uint32_t)(maxX / kTile) << 16);
                    out.ty[0] = (uint32_t)ty0 | ((uint32_t)ty1 << 16);
                }
            }
        }
    }
    // Warp-aggregated append to the compact accepted-triangle list.
    const int lane = threadIdx.x & 31;
    int n = out.n;
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(full, incl, o);
        if (lane >= o) incl += v;
    }
    int total = __shfl_sync(full, incl, 31);
    if (total == 0) return;                                   // uniform over the warp
    uint32_t base = 0;
    if (lane == 31) base = atomicAdd(&cnt->nRefs, (uint32_t)total);
    base = __shfl_sync(full, base, 31);
    for (int k = 0; k < n; ++k) {
        TriRef rf;
        rf.seq = out.seq[k]; rf.tx = out.tx[k]; rf.ty = out.ty[k]; rf._pad = 0;
        refs[base + (uint32_t)(incl - n + k)] = rf;
    }
    // Count pass.
    warp_bin_tiles<0>(out.n > 0, out.seq[0], out.tx[0], out.ty[0], fp.tilesX, tileCount, nullptr, 0u);
    if (__ballot_sync(full, out.n > 1))
        warp_bin_tiles<0>(out.n > 1, out.seq[1], out.tx[1], out.ty[1], fp.tilesX, tileCount, nullptr, 0u);
}

// ---------------------------------------------------------------------------
// Exclusive scan of the per-tile counts over this context's band: one CTA,
// kScanPer consecutive tiles per thread held in registers (independent loads,
// block scans only -- no atomics: heavy tiles cluster on the horizon and would
// serialise on a counter).  Also emits the launch order of the tile kernel:
// heavy bins first, then medium, then the rest in natural order, so the few
// 1000-triangle tiles start in the first wave and never form the tail.
// tileStart[t] = offset of bin t, tileCursor[t] = same (the fill pass bumps it
// so it ends as the bin's end offset).
// ---------------------------------------------------------------------------
constexpr uint32_t kHeavyBin = 192, kMediumBin = 32;
constexpr int kScanIters = 16;                         // per chunk each warp scans 16 x 32 consecutive tiles

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// All accesses are lane-contiguous: a single SM retires about one memory request per cycle,
// so a per-thread strided layout (32 sectors per instruction) would bound this kernel.
__global__ void __launch_bounds__(1024)
scan_tiles_kernel(const uint32_t* __restrict__ tileCount, uint32_t* __restrict__ tileStart,
                  uint32_t* __restrict__ tileCursor, uint32_t* __restrict__ order, int tile0, int nTiles,
                  Counters* cnt, uint32_t pairCap) {
    __shared__ uint32_t wsum[3][32];
    __shared__ uint32_t tot[2];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int tileEnd = tile0 + nTiles;
    // pass 1: how many heavy / medium tiles the band has (fixes the three regions of order[])
    uint32_t nh = 0, nm = 0;
    for (int i = tile0 + (int)threadIdx.x; i < tileEnd; i += 1024) {
        const uint32_t c = __ldg(tileCount + i);
        nh += c >= kHeavyBin;
        nm += (c >= kMediumBin && c < kHeavyBin);
    }
    nh = warp_sum(nh); nm = warp_sum(nm);
    if (lane == 0) { wsum[0][wid] = nh; wsum[1][wid] = nm; }
    __syncthreads();
    if (wid == 0) {
        uint32_t a = warp_sum(wsum[0][lane]), b = warp_sum(wsum[1][lane]);
        if (lane == 0) { tot[0] = a; tot[1] = b; }
    }
    __syncthreads();
    const uint32_t th = tot[0], tm = tot[1];
    // pass 2: offsets and order, chunk by chunk; warp w owns kScanIters*32 consecutive tiles of the chunk
    uint32_t carry = 0, carryH = 0, carryM = 0;
    for (int chunk = tile0; chunk < tileEnd; chunk += 1024 * kScanIters) {
        const int wb = chunk + wid * (32 * kScanIters);
        uint32_t c[kScanIters];
        uint32_t sum = 0, sh = 0, smd = 0;
#pragma unroll
        for (int k = 0; k < kScanIters; ++k) {
            const int i = wb + k * 32 + lane;
            c[k] = i < tileEnd ? __ldg(tileCount + i) : 0u;
            sum += c[k];
            sh += c[k] >= kHeavyBin;
            smd += (c[k] >= kMediumBin && c[k] < kHeavyBin);
        }
        sum = warp_sum(sum); sh = warp_sum(sh); smd = warp_sum(smd);
        __syncthreads();                                   // wsum of the previous round fully consumed
        if (lane == 0) { wsum[0][wid] = sum; wsum[1][wid] = sh; wsum[2][wid] = smd; }
        __syncthreads();
        uint32_t run = carry, oh = carryH, om = carryM;    // exclusive offsets of this warp's segment
        {
            const uint32_t a = wsum[0][lane], b = wsum[1][lane], d = wsum[2][lane];
            const unsigned below = lane < wid ? full : 0u;
            run += warp_sum(a & below); oh += warp_sum(b & below); om += warp_sum(d & below);
            carry += warp_sum(a); carryH += warp_sum(b); carryM += warp_sum(d);
        }
        uint32_t orest = (uint32_t)(min(wb, tileEnd) - tile0) - oh - om;      // earlier tiles that are neither
#pragma unroll
        for (int k = 0; k < kScanIters; ++k) {
            const int i = wb + k * 32 + lane;
            uint32_t incl = c[k];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(full, incl, o);
                if (lane >= o) incl += t;
            }
            const bool valid = i < tileEnd;
            const bool hv = valid && c[k] >= kHeavyBin, md = valid && c[k] >= kMediumBin && !hv;
            const unsigned mh = __ballot_sync(full, hv), mm = __ballot_sync(full, md), mv = __ballot_sync(full, valid);
            const unsigned lt = (1u << lane) - 1u;
            if (valid) {
                const uint32_t excl = run + incl - c[k];
                tileStart[i] = excl;
                tileCursor[i] = excl;
                uint32_t pos;
                if (hv) pos = oh + __popc(mh & lt);
                else if (md) pos = th + om + __popc(mm & lt);
                else pos = th + tm + orest + __popc(mv & ~mh & ~mm & lt);
                order[pos] = (uint32_t)i;
            }
            run += __shfl_sync(full, incl, 31);
            oh += __popc(mh); om += __popc(mm); orest += __popc(mv & ~mh & ~mm);
        }
    }
    if (threadIdx.x == 0) {
        cnt->nPairs = carry;
        if (carry > pairCap) atomicOr(&cnt->overflow, 1u);
    }
}

// ---------------------------------------------------------------------------
// Fill pass: grid-stride over the compact accepted-triangle list; bins receive
// seq ids in arbitrary order (the visibility path is order-independent; the
// ordered path sorts each bin by seq before use).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
fill_bins_kernel(const TriRef* __restrict__ refs, const Counters* __restrict__ cnt, FrameParams fp,
                 uint32_t* __restrict__ tileCursor, uint32_t* __restrict__ bins) {
    const uint32_t nRefs = cnt->nRefs;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; (i & ~31u) < nRefs; i += stride) {
        const bool valid = i < nRefs;
        TriRef rf;
        rf.seq = 0; rf.tx = 0; rf.ty = 0; rf._pad = 0;
        if (valid) rf = refs[i];
        warp_bin_tiles<1>(valid, rf.seq, rf.tx, rf.ty, fp.tilesX, tileCursor, bins, fp.pairCap);
    }
}

}  // namespace srb
