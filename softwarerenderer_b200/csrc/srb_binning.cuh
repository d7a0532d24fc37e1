// srb_binning.cuh -- triangle-setup kernel and the count / scan / fill binning
// passes.  Included by srb_kernels.cu only.
#pragma once
#include "srb_geometry.cuh"

namespace srb {

// Visit every tile of every accepted triangle held by the lanes of one warp.
// Triangles touching <= kSmallTiles tiles are walked by their own lane; larger
// ones are walked by the whole warp (ballot over the "big" lanes), so that a
// screen-filling triangle does not serialise on one thread.  All 32 lanes must
// call this (no early returns before it).
constexpr int kSmallTiles = 4;
template <typename F>
__device__ __forceinline__ void warp_for_each_tile(bool valid, uint32_t seq, uint32_t tx, uint32_t ty,
                                                   int tilesX, F f) {
    const unsigned full = 0xffffffffu;
    int x0 = tx & 0xffff, x1 = tx >> 16, y0 = ty & 0xffff, y1 = ty >> 16;
    int nx = x1 - x0 + 1, n = valid ? nx * (y1 - y0 + 1) : 0;
    if (n > 0 && n <= kSmallTiles) {
        for (int y = y0; y <= y1; ++y)
            for (int x = x0; x <= x1; ++x) f(y * tilesX + x, seq);
    }
    unsigned big = __ballot_sync(full, n > kSmallTiles);
    const int lane = threadIdx.x & 31;
    while (big) {
        int src = __ffs(big) - 1;
        big &= big - 1;
        int bx0 = __shfl_sync(full, x0, src), by0 = __shfl_sync(full, y0, src);
        int bnx = __shfl_sync(full, nx, src), bn = __shfl_sync(full, n, src);
        uint32_t bseq = __shfl_sync(full, seq, src);
        for (int i = lane; i < bn; i += 32) {
            int yy = i / bnx, xx = i - yy * bnx;
            f((by0 + yy) * tilesX + bx0 + xx, bseq);
        }
    }
}

// ---------------------------------------------------------------------------
// Setup kernel: one thread per submitted triangle.  Body of RenderMesh's
// Parallel.For lambda (Rasterizer.cs:200-229) with the vertex shader hoisted
// into vertex_kernel: behind tests, optional near clip, DrawTriangle, cull,
// bbox; then per-tile counting for the bins.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
setup_kernel(const DrawRec* __restrict__ draws, const int32_t* __restrict__ tbase, FrameParams fp,
             SrbUniforms U, const float4* __restrict__ clipIn, const float4* __restrict__ wnIn,
             TriRec* __restrict__ recs, ClipAttr* __restrict__ clipAttrs, TriRef* __restrict__ refs,
             uint32_t* __restrict__ tileCount, Counters* cnt) {
    const unsigned full = 0xffffffffu;
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    SetupOut out;
    out.n = 0;
    if (g < fp.totalT) {
        int di = find_owner(tbase, fp.nDraws, g);
        const DrawRec& d = draws[di];
        if (d.visible) {
            int t = g - d.tbase;
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
                        recs[seq] = r;
                        out.n = 1;
                        out.seq[0] = seq;
                        out.tx[0] = (uint32_t)(minX / kTile) | ((uint32_t)(maxX / kTile) << 16);
                        out.ty[0] = (uint32_t)ty0 | ((uint32_t)ty1 << 16);
                    }
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
    uint32_t base = 0;
    if (lane == 31 && total > 0) base = atomicAdd(&cnt->nRefs, (uint32_t)total);
    base = __shfl_sync(full, base, 31);
    for (int k = 0; k < n; ++k) {
        TriRef rf;
        rf.seq = out.seq[k]; rf.tx = out.tx[k]; rf.ty = out.ty[k]; rf._pad = 0;
        refs[base + (uint32_t)(incl - n + k)] = rf;
    }
    // Count pass.
    auto count = [&](int tile, uint32_t) { atomicAdd(tileCount + tile, 1u); };
    warp_for_each_tile(out.n > 0, out.seq[0], out.tx[0], out.ty[0], fp.tilesX, count);
    unsigned second = __ballot_sync(full, out.n > 1);
    if (second) warp_for_each_tile(out.n > 1, out.seq[1], out.tx[1], out.ty[1], fp.tilesX, count);
}

// ---------------------------------------------------------------------------
// Exclusive scan of the per-tile counts (single CTA; <= 130k tiles at 8K).
// tileStart[t] = offset of bin t, tileCursor[t] = same (fill pass bumps it so
// it ends as the bin's end offset).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
scan_tiles_kernel(const uint32_t* __restrict__ tileCount, uint32_t* __restrict__ tileStart,
                  uint32_t* __restrict__ tileCursor, int nTiles, Counters* cnt, uint32_t pairCap) {
    __shared__ uint32_t warpSums[32];
    __shared__ uint32_t carry, chunkTotal;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nTiles; base += 1024) {
        int i = base + threadIdx.x;
        uint32_t v = i < nTiles ? tileCount[i] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(full, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warpSums[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            uint32_t ws = warpSums[lane];
            uint32_t wi = ws;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(full, wi, o);
                if (lane >= o) wi += t;
            }
            warpSums[lane] = wi - ws;     // exclusive per-warp offset
            if (lane == 31) chunkTotal = wi;
        }
        __syncthreads();
        uint32_t excl = carry + warpSums[wid] + incl - v;
        if (i < nTiles) { tileStart[i] = excl; tileCursor[i] = excl; }
        __syncthreads();
        if (threadIdx.x == 0) carry += chunkTotal;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        cnt->nPairs = carry;
        if (carry > pairCap) atomicOr(&cnt->overflow, 1u);
    }
}

// ---------------------------------------------------------------------------
// Fill pass: one thread per accepted triangle (compact list); bins receive
// seq ids in arbitrary order (the visibility path is order-independent; the
// ordered path sorts each bin by seq before use).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
fill_bins_kernel(const TriRef* __restrict__ refs, const Counters* __restrict__ cnt, FrameParams fp,
                 uint32_t* __restrict__ tileCursor, uint32_t* __restrict__ bins) {
    uint32_t nRefs = cnt->nRefs;
    uint32_t warpBase = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u;
    if (warpBase >= nRefs) return;                      // whole warp exits together
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = i < nRefs;
    TriRef rf;
    rf.seq = 0; rf.tx = 0; rf.ty = 0; rf._pad = 0;
    if (valid) rf = refs[i];
    const uint32_t cap = fp.pairCap;
    auto put = [&](int tile, uint32_t seq) {
        uint32_t slot = atomicAdd(tileCursor + tile, 1u);
        if (slot < cap) bins[slot] = seq;
    };
    warp_for_each_tile(valid, rf.seq, rf.tx, rf.ty, fp.tilesX, put);
}

}  // namespace srb
