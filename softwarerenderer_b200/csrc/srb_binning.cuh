// srb_binning.cuh -- triangle-setup kernel and the count / scan / fill binning
// passes.  Included by srb200.cu only.
#pragma once
#include "srb_geometry.cuh"

namespace srb {

// Tile lists are built in three passes: setup counts (RED atomics), one CTA
// scans, fill scatters.  Distant instances put thousands of consecutive
// triangles into the same tile, so per-tile atomics are warp-aggregated with
// match.any: one atomic per distinct tile per warp instead of one per lane.

__device__ __forceinline__ int tile_range_count(uint32_t tx, uint32_t ty) {
    return ((int)(tx >> 16) - (int)(tx & 0xffff) + 1) * ((int)(ty >> 16) - (int)(ty & 0xffff) + 1);
}

// k-th tile of a packed range (row-major), or -1.
__device__ __forceinline__ int small_tile(uint32_t tx, uint32_t ty, int n, int k, int tilesX) {
    if (k >= n) return -1;
    int x0 = tx & 0xffff, nx = (int)(tx >> 16) - x0 + 1, y0 = ty & 0xffff;
    int yy = k / nx, xx = k - yy * nx;
    return (y0 + yy) * tilesX + x0 + xx;
}

// Small triangles (<= kSmallTiles tiles) go through a CTA-wide aggregation table in shared
// memory: (tile -> number of pairs from this CTA), open addressing keyed by the tile id.  Distant
// instances put thousands of consecutive triangles into the same few tiles, and atomics on one
// L2 address retire serially (tens of ns each): aggregated per CTA, a hot tile receives one
// global atomic per CTA instead of one per triangle.
constexpr int kAggSlots = 1024;            // >= the small pairs one CTA pass can produce (256 threads x 4)
struct TileAgg {
    uint32_t key[kAggSlots];               // tile + 1, 0 = empty
    uint32_t val[kAggSlots];               // pairs counted; in fill mode replaced by the bin offset of the run
};
__device__ __forceinline__ void agg_clear(TileAgg& a) {
    for (int i = threadIdx.x; i < kAggSlots; i += blockDim.x) { a.key[i] = 0u; a.val[i] = 0u; }
}
// Returns slot | rank << 16 (rank = position of this pair within the CTA's run for that tile).
__device__ __forceinline__ uint32_t agg_insert(TileAgg& a, int tile) {
    uint32_t h = (uint32_t)tile & (kAggSlots - 1);
    const uint32_t want = (uint32_t)tile + 1u;
    while (true) {
        const uint32_t k = atomicCAS(&a.key[h], 0u, want);
        if (k == 0u || k == want) break;
        h = (h + 1u) & (kAggSlots - 1);
    }
    return h | (atomicAdd(&a.val[h], 1u) << 16);
}
// One global atomic per distinct tile.  Count mode adds to tileCount; fill mode reserves the run
// in the bin and leaves its offset in val[].
template <int MODE>
__device__ __forceinline__ void agg_flush(TileAgg& a, uint32_t* __restrict__ tileCounter) {
    for (int i = threadIdx.x; i < kAggSlots; i += blockDim.x) {
        const uint32_t k = a.key[i];
        if (k) {
            const uint32_t base = atomicAdd(tileCounter + (k - 1u), a.val[i]);
            SRB_ASSERT(MODE == 0 || base + a.val[i] >= base);
            if (MODE == 1) a.val[i] = base;
        }
    }
}

// Large triangles (more than kSmallTiles tiles) are appended to a global queue by the setup
// kernel and binned by their own kernel, one warp per triangle and one lane per tile of its
// bounding box, four atomics in flight per lane.  Nearby meshes are runs of hundreds of large
// triangles: handled inside the CTA that owns them they made a 40 us tail; as a queue they
// spread over every SM.  Each (triangle, tile) pair passes the conservative tile test first
// (about a quarter of the bbox tiles of large triangles drop out).
// MODE 0 = count (tileCount += 1), MODE 1 = fill (bins[cursor++] = seq).
constexpr int kBigSplit = 128;             // tiles per queue entry = one pass of the warp's 4 x 32 lanes
__device__ __forceinline__ int big_rows(uint32_t tx) {                 // tile rows per entry
    return max(1, kBigSplit / ((int)(tx >> 16) - (int)(tx & 0xffff) + 1));
}
__device__ __forceinline__ int big_parts(uint32_t tx, uint32_t ty) {   // queue entries of one large triangle
    const int rows = big_rows(tx), ny = (int)(ty >> 16) - (int)(ty & 0xffff) + 1;
    return (ny + rows - 1) / rows;
}
// (Measured, no gain: keeping the count pass's verdicts in the entry for the fill pass -- the passes wait
// on their atomics, not on the test.)
template <int MODE>
__device__ __forceinline__ void bin_big_body(const BigEntry* __restrict__ q, const Counters* __restrict__ cnt,
                                             const FrameParams& fp, uint32_t* __restrict__ tileCounter,
                                             uint32_t* __restrict__ bins, const uint32_t block, const uint32_t nBlocks) {
    if (cnt->overflow) return;
    const uint32_t nq = min(cnt->nBig, fp.bigCap);
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (block * blockDim.x + threadIdx.x) >> 5, nWarps = (nBlocks * blockDim.x) >> 5;
    for (uint32_t e = warp; e < nq; e += nWarps) {
        const float4* p = reinterpret_cast<const float4*>(q + e);
        const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
        const uint32_t bx = __float_as_uint(b.z), by = __float_as_uint(b.w);
        const uint32_t seq = __float_as_uint(c.x), tx = __float_as_uint(c.y), ty = __float_as_uint(c.z);
        const int x0 = tx & 0xffff, nx = (int)(tx >> 16) - x0 + 1, y0 = ty & 0xffff;
        const int n = nx * ((int)(ty >> 16) - y0 + 1);
        const int minX = bx & 0xffffu, maxX = bx >> 16, minY = by & 0xffffu, maxY = by >> 16;
        for (int i0 = lane; i0 < n; i0 += 128) {
            int tile[4];
            uint32_t slot[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + 32 * u;
                const int yy = i / nx, xx = i - yy * nx;
                const int tX = x0 + xx, tY = y0 + yy;
                SRB_ASSERT(i >= n || (tX < fp.tilesX && tY >= fp.bandRow0 && tY < fp.bandRow1));
                const bool keep = i < n && (fp.wireframe ||     // lines are not bound by the fill test; the bbox tiles are kept
                                           tile_may_cover(a.x, a.y, a.z, a.w, b.x, b.y,
                                                          max(minX, tX * kTile), min(maxX, tX * kTile + kTile - 1),
                                                          max(minY, tY * kTile), min(maxY, tY * kTile + kTile - 1)));
                tile[u] = keep ? tY * fp.tilesX + tX : -1;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (tile[u] >= 0) {
                    if (MODE == 0) {
                        atomicAdd(tileCounter + tile[u], 1u);
                        bins[tile[u]] = 1u;        // count pass: `bins` is the per-tile "holds a large triangle" flag array
                    } else {
                        slot[u] = atomicAdd(tileCounter + tile[u], 1u);
                    }
                }
            if (MODE == 1) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (tile[u] >= 0 && slot[u] < fp.pairCap) bins[slot[u]] = seq;
            }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(256)
bin_big_kernel(const BigEntry* __restrict__ q, const Counters* __restrict__ cnt, FrameParams fp,
               uint32_t* __restrict__ tileCounter, uint32_t* __restrict__ bins) {
    bin_big_body<MODE>(q, cnt, fp, tileCounter, bins, blockIdx.x, gridDim.x);
}

// ---------------------------------------------------------------------------
// Setup kernel: one thread per submitted triangle, one CTA per block-table
// entry (a run of <= kGeomBlock triangles of ONE draw, so the draw lookup is a
// single load and a culled draw retires whole CTAs at once).  Body of
// RenderMesh's Parallel.For lambda (Rasterizer.cs:200-229) with the vertex
// shader hoisted into vertex_kernel: behind tests, optional near clip,
// DrawTriangle, cull, bbox; then per-tile counting for the bins.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kGeomBlock, 8)
setup_kernel(const DrawRec* __restrict__ draws, const int2* __restrict__ tblocks, FrameParams fp,
             SrbUniforms U, const float4* __restrict__ clipIn, const float4* __restrict__ wnIn,
             RecStore recs, ClipAttr* __restrict__ clipAttrs, TriRef* __restrict__ refs, BigEntry* __restrict__ bigQ,
             uint32_t* __restrict__ tileCount, Counters* cnt) {
    const unsigned full = 0xffffffffu;
    __shared__ TileAgg agg;
    __shared__ uint32_t sNoCover;
    // tblocks is the live table the frustum kernel of this frame compacted (visible draws only); stride loop and
    // early first-entry load as in vertex_kernel.
    int2 blk = tblocks[blockIdx.x];
    const uint32_t nLive = cnt->nTLive;
  for (uint32_t bi = blockIdx.x; bi < nLive; bi += gridDim.x, blk = tblocks[min(bi, nLive - 1)]) {
    const int di = blk.x;
    const DrawRec& d = draws[di];
    // (A band context's table holds only blocks that may touch the band: band_blocks_kernel.)
    agg_clear(agg);
    if (threadIdx.x == 0) sNoCover = 0u;
    __syncthreads();
    const int t = blk.y + threadIdx.x;
    SetupOut out;
    out.n = 0;
    out.nNoCover = 0;
    out.bigMask = 0u;
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
                r.meta = (uint32_t)di;
                accept_tri(r, 2u * (uint32_t)g, minX, maxX, minY, maxY, fp, recs, out);
            }
        }
    }
    // Append to the compact accepted-triangle list (the fill pass walks it) and to the large-triangle queue.
    // Both lengths live in one 64-bit word, so a warp reserves its share of both with a single atomic
    // (every reservation is a round trip on the same L2 address).  (Measured: one reservation per CTA
    // instead -- two more barriers -- city4k setup 44.3 -> 45.8 us, stress8k 703 -> 656 us.)
    const int lane = threadIdx.x & 31;
    // A large triangle takes one queue entry per <= kBigSplit tiles (whole tile rows): bin_big_kernel gives an
    // entry to one warp, and the few triangles that span thousands of tiles would otherwise be its tail.
    int nb = 0;
    for (int k = 0; k < out.n; ++k)
        if (out.bigMask >> k & 1u) nb += big_parts(out.tx[k], out.ty[k]);
    const int n = out.n;
    // list entries in the low word, queue entries in the high word -- the layout of the counter pair itself, so that
    // no count can carry into the other (a triangle of a 65535-px target can take thousands of queue entries)
    const unsigned long long mine = (unsigned long long)(uint32_t)n | ((unsigned long long)(uint32_t)nb << 32);
    unsigned long long incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long v = __shfl_up_sync(full, incl, o);
        if (lane >= o) incl += v;
    }
    if (out.nNoCover) atomicAdd(&sNoCover, (uint32_t)out.nNoCover);      // statistics; flushed once per CTA below
    const unsigned long long tot = __shfl_sync(full, incl, 31);
    if (tot) {                                                // uniform over the warp
        unsigned long long base = 0ull;
        if (lane == 31) base = atomicAdd(reinterpret_cast<unsigned long long*>(&cnt->nRefs), tot);
        base = __shfl_sync(full, base, 31);
        const unsigned long long excl = incl - mine;
        uint32_t ri = (uint32_t)base + (uint32_t)excl, bi = (uint32_t)(base >> 32) + (uint32_t)(excl >> 32);
        for (int k = 0; k < n; ++k, ++ri) {
            SRB_ASSERT(ri < 2u * (uint32_t)fp.totalT + 64u);
            TriRef rf;
            rf.seq = out.seq[k]; rf.tx = out.tx[k]; rf.ty = out.ty[k]; rf._pad = 0;
            refs[ri] = rf;
            if (out.bigMask >> k & 1u) {
                const int parts = big_parts(out.tx[k], out.ty[k]);
                if (bi + parts > fp.bigCap) { atomicOr(&cnt->overflow, 4u); bi += parts; continue; }
                // the record this thread stored a moment ago: plain (coherent) loads, not the read-only path
                const volatile float4* q = recs.R + 3 * rec_slot(recs, out.seq[k]);
                const float ax = q[0].x, ay = q[0].y, az = q[0].z, aw = q[0].w, bx = q[1].x, by = q[1].y;
                const float cz = q[2].z, cw = q[2].w;
                const int y0 = out.ty[k] & 0xffff, y1 = out.ty[k] >> 16, rows = big_rows(out.tx[k]);
                for (int y = y0; y <= y1; y += rows) {
                    BigEntry& e = bigQ[bi++];
                    e.g = BinGeom{ax, ay, az, aw, bx, by, __float_as_uint(cz), __float_as_uint(cw)};
                    e.seq = out.seq[k]; e.tx = out.tx[k];
                    e.ty = (uint32_t)y | ((uint32_t)min(y + rows - 1, y1) << 16);
                    e._pad = 0u;
                }
            }
        }
    }
    // Count pass: small triangles through the aggregation table (large ones: bin_big_kernel<0>).
    for (int k = 0; k < out.n; ++k) {
        const int nt = tile_range_count(out.tx[k], out.ty[k]);
        if (nt <= kSmallTiles)
            for (int j = 0; j < nt; ++j) {
                const int tile = small_tile(out.tx[k], out.ty[k], nt, j, fp.tilesX);
                SRB_ASSERT(tile >= fp.bandRow0 * fp.tilesX && tile < fp.bandRow1 * fp.tilesX);
                agg_insert(agg, tile);
            }
    }
    __syncthreads();
    agg_flush<0>(agg, tileCount);
    if (threadIdx.x == 0 && sNoCover) atomicAdd(&cnt->nNoCover, sNoCover);
    __syncthreads();                                          // the table and sNoCover are reused by the next block
  }
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
// Bin splitting.  A bin of >= splitMin entries is rasterized by several CTAs, each walking partSize of its entries
// into its own shared-memory keys and merging them into one global 256-key slice with 64-bit atomic max; the CTA
// that finishes last (a counter per slice) resolves the tile.  The extra launch-order entries (one per part beyond
// the first) live behind the one-per-tile entries and are launched FIRST.  At most kMaxExtra of them: partSize grows
// with the frame's total of split entries so that they always fit.  A distant crowd puts 500-1000 triangles into
// single tiles; one CTA walks such a bin in 128-entry chunks for 40-80 us, which is the floor of any band that owns
// one (multi-GPU) however few tiles the band has.
constexpr uint32_t kMaxExtra = 4096;
constexpr int kScanIters = 16;                         // per chunk each warp scans 16 x 32 consecutive tiles

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One thread-block cluster of kScanCtas CTAs (one launch, one cluster barrier): CTA r owns a
// contiguous range of tiles, totals are exchanged through distributed shared memory.  All
// accesses are lane-contiguous: an SM retires about one memory request per cycle, so a
// per-thread strided layout (32 sectors per instruction) would bound this kernel.
constexpr int kScanCtas = 8;

// The frame's counters are final when this kernel ends (nothing after it writes them).  It publishes them to the
// host itself, as plain stores into mapped pinned memory: a copy-engine transfer on this stream would queue behind
// whatever the D2H engine is doing -- the 100 MB present copy of the previous frame, for one -- and hold up the
// rest of the frame with it.
__global__ void publish_counters_kernel(const Counters* __restrict__ cnt, uint32_t* __restrict__ hostCnt) {
    hostCnt[threadIdx.x] = reinterpret_cast<const uint32_t*>(cnt)[threadIdx.x];
    __threadfence_system();
}

// SPLIT: bins may be split (visibility path); the plain instance carries none of that code.
template <bool SPLIT>
__global__ void __cluster_dims__(kScanCtas, 1, 1) __launch_bounds__(1024)
scan_tiles_kernel(const uint32_t* __restrict__ tileCount, uint32_t* __restrict__ tileStart,
                  uint32_t* __restrict__ tileCursor, uint4* __restrict__ order, int tile0, int nTiles,
                  Counters* cnt, uint32_t pairCap, const uint32_t splitMin_, const uint32_t splitPart,
                  uint32_t* __restrict__ hostCnt, const uint32_t* __restrict__ tileHasBig, const uint32_t maxExtra) {
    const uint32_t splitMin = SPLIT ? splitMin_ : 0u;
    SRB_ASSERT(!SPLIT || (maxExtra >= 1u && maxExtra <= kMaxExtra));
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ uint32_t wsum[4][32];
    __shared__ uint32_t ctaTot[4];                      // this CTA's range totals: pairs, heavy, medium, entries of split bins (read remotely)
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int rank = (int)cluster.block_rank();
    const int per = (((nTiles + kScanCtas - 1) / kScanCtas) + 31) & ~31;
    const int r0 = min(tile0 + rank * per, tile0 + nTiles), r1 = min(r0 + per, tile0 + nTiles);
    // phase 1: totals of this CTA's range
    uint32_t ns = 0, nh = 0, nm = 0, nsp = 0;
    for (int i0 = r0 + (int)threadIdx.x; i0 < r1; i0 += 4 * 1024) {
        uint32_t c[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) c[u] = i0 + u * 1024 < r1 ? __ldg(tileCount + i0 + u * 1024) : 0u;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            ns += c[u];
            nh += c[u] >= kHeavyBin;
            nm += (c[u] >= kMediumBin && c[u] < kHeavyBin);
            if (SPLIT) nsp += (splitMin != 0u && c[u] >= splitMin) ? c[u] : 0u;
        }
    }
    ns = warp_sum(ns); nh = warp_sum(nh); nm = warp_sum(nm);
    if (SPLIT) nsp = warp_sum(nsp);
    if (lane == 0) { wsum[0][wid] = ns; wsum[1][wid] = nh; wsum[2][wid] = nm; wsum[3][wid] = nsp; }
    __syncthreads();
    if (wid == 0) {
        uint32_t a = warp_sum(wsum[0][lane]), b = warp_sum(wsum[1][lane]), d = warp_sum(wsum[2][lane]);
        uint32_t e = SPLIT ? warp_sum(wsum[3][lane]) : 0u;
        if (lane == 0) { ctaTot[0] = a; ctaTot[1] = b; ctaTot[2] = d; ctaTot[3] = e; }
    }
    cluster.sync();
    uint32_t carry = 0, carryH = 0, carryM = 0, tp = 0, th = 0, tm = 0, tsp = 0;
#pragma unroll
    for (int r = 0; r < kScanCtas; ++r) {
        const uint32_t* t = cluster.map_shared_rank(ctaTot, r);
        const uint32_t a = t[0], b = t[1], d = t[2];
        if (r < rank) { carry += a; carryH += b; carryM += d; }
        tp += a; th += b; tm += d;
        if (SPLIT) tsp += t[3];
    }
    // Entries per part: a multiple of 32, at least splitPart, and large enough that the parts beyond the first of
    // all split bins fit maxExtra -- the extra CTAs this frame's tile kernel is launched with, sized by the host from
    // the frames before (<= kMaxExtra) -- since sum over bins of ceil(c/ps) - 1 <= sum of c/ps.
    const uint32_t partSize = 32u * max(max(splitPart, 32u) / 32u, (tsp + 32u * maxExtra - 1u) / (32u * maxExtra));
    cluster.sync();                                     // nobody leaves (or reuses wsum) while its totals are being read
    // phase 2: offsets and order; warp w owns kScanIters*32 consecutive tiles of each chunk
    for (int chunk = r0; chunk < r1; chunk += 1024 * kScanIters) {
        const int wb = chunk + wid * (32 * kScanIters);
        uint32_t c[kScanIters];
        uint32_t sum = 0, sh = 0, smd = 0, bigBits = 0;
#pragma unroll
        for (int k = 0; k < kScanIters; ++k) {
            const int i = wb + k * 32 + lane;
            c[k] = i < r1 ? __ldg(tileCount + i) : 0u;
            bigBits |= (i < r1 && __ldg(tileHasBig + i) != 0u) ? (1u << k) : 0u;     // loaded with the counts, not later
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
        uint32_t orest = (uint32_t)(min(wb, r1) - tile0) - oh - om;      // earlier tiles that are neither
        if (wb >= r1) continue;                            // (uniform per warp; no barrier below)
        // Split bins: the launch-order entries of their parts beyond the first are reserved with ONE atomic per warp
        // and chunk (a reservation per iteration below put 16 dependent L2 round trips into this latency-bound kernel).
        uint32_t ebase = 0u;
        if (SPLIT && splitMin != 0u) {
            uint32_t mine = 0u;
#pragma unroll
            for (int k = 0; k < kScanIters; ++k)
                if (c[k] >= splitMin && c[k] > partSize) mine += (c[k] + partSize - 1u) / partSize - 1u;     // c[k] == 0 past r1
            uint32_t inc = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(full, inc, o);
                if (lane >= o) inc += t;
            }
            const uint32_t tot = __shfl_sync(full, inc, 31);
            if (tot) {                                     // uniform over the warp
                if (lane == 31) ebase = atomicAdd(&cnt->nExtra, tot);
                ebase = __shfl_sync(full, ebase, 31) + inc - mine;
            }
        }
#pragma unroll
        for (int k = 0; k < kScanIters; ++k) {
            const int i = wb + k * 32 + lane;
            uint32_t incl = c[k];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(full, incl, o);
                if (lane >= o) incl += t;
            }
            const bool valid = i < r1;
            const bool hv = valid && c[k] >= kHeavyBin, md = valid && c[k] >= kMediumBin && !hv;
            const unsigned mh = __ballot_sync(full, hv), mm = __ballot_sync(full, md), mv = __ballot_sync(full, valid);
            const unsigned lt = (1u << lane) - 1u;
            const uint32_t excl = run + incl - c[k];
            uint32_t pos = 0u;
            if (hv) pos = oh + __popc(mh & lt);
            else if (md) pos = th + om + __popc(mm & lt);
            else pos = th + tm + orest + __popc(mv & ~mh & ~mm & lt);
            // Split bins: the tile's own entry becomes part 0; the index of its first extra entry doubles as the id
            // of its key slice.
            // bit 31: a triangle that spans many tiles is in this bin (they are filled first: hi-Z has its occluders)
            const uint32_t big = (bigBits >> k & 1u) ? 0x80000000u : 0u;
            uint32_t w = big, firstEnd = excl + c[k];
            if (SPLIT && splitMin != 0u && c[k] >= splitMin && c[k] > partSize) {
                const uint32_t extra = (c[k] + partSize - 1u) / partSize - 1u;
                w = (extra + 1u) | (ebase << 16) | big;    // parts (<= kMaxExtra + 1) | slice (< kMaxExtra) | large-triangle flag
                firstEnd = excl + partSize;
                SRB_ASSERT(ebase + extra <= maxExtra && extra + 1u <= 0xffffu);
                for (uint32_t q = 1; q <= extra; ++q)
                    order[(uint32_t)nTiles + ebase + q - 1u] =
                        make_uint4((uint32_t)i, excl + q * partSize, min(excl + (q + 1u) * partSize, excl + c[k]), w);
                ebase += extra;
            }
            if (valid) {
                tileStart[i] = excl;
                tileCursor[i] = excl;
                order[pos] = make_uint4((uint32_t)i, excl, firstEnd, w);      // tile, bin start, bin end, split info
            }
            run += __shfl_sync(full, incl, 31);
            oh += __popc(mh); om += __popc(mm); orest += __popc(mv & ~mh & ~mm);
        }
    }
    if (rank == 0 && threadIdx.x == 0) {
        cnt->nPairs = tp;
        if (SPLIT) cnt->nSplitEntries = tsp;
        if (tp > pairCap) atomicOr(&cnt->overflow, 1u);
    }
    cluster.sync();                                     // every CTA's reservations (nExtra) are in
    if (rank == 0 && threadIdx.x < sizeof(Counters) / 4) {
        hostCnt[threadIdx.x] = reinterpret_cast<volatile uint32_t*>(cnt)[threadIdx.x];
        __threadfence_system();
    }
}

// ---------------------------------------------------------------------------
// Fill pass: grid-stride over the compact accepted-triangle list; bins receive
// seq ids in arbitrary order (the visibility path is order-independent; the
// ordered path sorts each bin by seq before use).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void fill_small_body(const TriRef* __restrict__ refs, const Counters* __restrict__ cnt,
                                                const FrameParams& fp, uint32_t* __restrict__ tileCursor,
                                                uint32_t* __restrict__ bins, TileAgg& agg, const uint32_t block,
                                                const uint32_t nBlocks) {
    const uint32_t nRefs = cnt->nRefs;
    const uint32_t stride = nBlocks * blockDim.x;
    for (uint32_t base = block * blockDim.x; base < nRefs; base += stride) {     // uniform over the CTA
        agg_clear(agg);
        __syncthreads();
        const uint32_t i = base + threadIdx.x;
        TriRef rf;
        rf.seq = 0; rf.tx = 0; rf.ty = 0; rf._pad = 0;
        if (i < nRefs) rf = refs[i];
        const int nt = i < nRefs ? tile_range_count(rf.tx, rf.ty) : 0;
        const int nSmall = nt <= kSmallTiles ? nt : 0;              // large ones: bin_big_body<1>
        uint32_t where[kSmallTiles];
#pragma unroll
        for (int j = 0; j < kSmallTiles; ++j)
            if (j < nSmall) where[j] = agg_insert(agg, small_tile(rf.tx, rf.ty, nSmall, j, fp.tilesX));
        __syncthreads();
        agg_flush<1>(agg, tileCursor);
        __syncthreads();
#pragma unroll
        for (int j = 0; j < kSmallTiles; ++j)
            if (j < nSmall) {
                const uint32_t slot = agg.val[where[j] & 0xffffu] + (where[j] >> 16);
                if (slot < fp.pairCap) bins[slot] = rf.seq;
            }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
fill_bins_kernel(const TriRef* __restrict__ refs, const Counters* __restrict__ cnt, FrameParams fp,
                 uint32_t* __restrict__ tileCursor, uint32_t* __restrict__ bins) {
    __shared__ TileAgg agg;
    fill_small_body(refs, cnt, fp, tileCursor, bins, agg, blockIdx.x, gridDim.x);
}

// Both fill passes in one launch, for frames whose accepted list fits one wave: the two are independent and
// each is a chain of atomic round trips that leaves the SMs mostly idle (city4k: 24 us as two launches,
// 20 us as one).  The large-triangle CTAs come first so that both roles are resident together.  Frames with
// millions of small triangles keep the lean fill kernel at full occupancy (this one needs twice the registers).
__global__ void __launch_bounds__(256, 4)
fill_merged_kernel(const TriRef* __restrict__ refs, const BigEntry* __restrict__ bigQ, const Counters* __restrict__ cnt,
                   FrameParams fp, uint32_t* __restrict__ tileCursor, uint32_t* __restrict__ bins, const uint32_t nBigBlocks) {
    __shared__ TileAgg agg;
    if (blockIdx.x < nBigBlocks) bin_big_body<1>(bigQ, cnt, fp, tileCursor, bins, blockIdx.x, nBigBlocks);
    else fill_small_body(refs, cnt, fp, tileCursor, bins, agg, blockIdx.x - nBigBlocks, gridDim.x - nBigBlocks);
}

}  // namespace srb
