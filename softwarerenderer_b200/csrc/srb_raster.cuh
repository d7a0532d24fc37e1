// srb_raster.cuh -- the two tile rasterizers.
//
//  * ordered_tile_kernel: exact submission-order semantics for ANY state
//    (all depth tests, all blend modes, alpha <= 0 discards, BlendMode.None
//    row break).  One CTA per 16x16 tile, one thread per pixel, bin sorted by
//    submission sequence, colour+depth of the pixel live in registers.
//  * vis_tile_kernel: the fast path for frames whose result is provably
//    order-independent: per-pixel winner = extreme depth, ties
//    by submission order, kept as a 64-bit key in shared memory; coverage is
//    walked by (triangle,row) work items; the winner is shaded once and the
//    tile is written with 128-bit stores.  The clear is fused into tile init.
#pragma once
#include "srb_shade.cuh"

namespace srb {

// Edge values at pixel (x,y) of the tile whose clipped bbox starts at
// (startX,startY): direct formula at the start pixel, then dy rounded adds of
// b and dx rounded adds of a -- the reference's accumulation order
// (Rasterizer.cs:481-483, 527-534).
__device__ __forceinline__ void edge_chain(const TriRec& r, const TriEq& e, int startX, int startY, int x, int y,
                                           float& w0, float& w1, float& w2) {
    row_base(r, e, startX, startY, w0, w1, w2);
    // (unroll 4: the loop counter, compare and branch are half of a step otherwise; raster 115.4 -> 113.2 us.  A fixed
    // 15-step loop under a predicate measured the same.)
#pragma unroll 4
    for (int i = startY; i < y; ++i) { w0 = add(w0, e.b12); w1 = add(w1, e.b20); w2 = add(w2, e.b01); }
#pragma unroll 4
    for (int i = startX; i < x; ++i) { w0 = add(w0, e.a12); w1 = add(w1, e.a20); w2 = add(w2, e.a01); }
}

// ---------------------------------------------------------------------------
// Bin sort by submission sequence (ordered path only).
// ---------------------------------------------------------------------------
constexpr int kSortSmem = 4096;
__global__ void __launch_bounds__(256)
sort_bins_kernel(const uint32_t* __restrict__ tileStart, const uint32_t* __restrict__ tileEnd,
                 uint32_t* __restrict__ bins, uint32_t* __restrict__ scratch, int tileBase,
                 const Counters* __restrict__ cnt) {
    __shared__ uint32_t sm[kSortSmem];
    if (cnt->overflow) return;
    int tile = tileBase + blockIdx.x;
    uint32_t s = tileStart[tile], e = tileEnd[tile];
    int n = (int)(e - s);
    if (n <= 1) return;
    uint32_t* b = bins + s;
    if (n <= kSortSmem) {
        int P = 1;
        while (P < n) P <<= 1;
        for (int i = threadIdx.x; i < P; i += blockDim.x) sm[i] = i < n ? b[i] : 0xffffffffu;
        __syncthreads();
        for (int k = 2; k <= P; k <<= 1) {
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = threadIdx.x; i < P; i += blockDim.x) {
                    int ixj = i ^ j;
                    if (ixj > i) {
                        uint32_t a = sm[i], c = sm[ixj];
                        bool up = (i & k) == 0;
                        if ((a > c) == up) { sm[i] = c; sm[ixj] = a; }
                    }
                }
                __syncthreads();
            }
        }
        for (int i = threadIdx.x; i < n; i += blockDim.x) b[i] = sm[i];
    } else {
        // Rank sort through a scratch copy (sequence ids are unique).
        uint32_t* t = scratch + s;
        for (int i = threadIdx.x; i < n; i += blockDim.x) t[i] = b[i];
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            uint32_t v = t[i];
            int rank = 0;
            for (int j = 0; j < n; ++j) rank += t[j] < v;
            b[rank] = v;
        }
    }
}

// ---------------------------------------------------------------------------
// DebugMode.Wireframe (Rasterizer.cs:419-425): one pixel against the three DrawLine calls of a triangle
// (Rasterizer.cs:232-340), in call order (s0,s1), (s1,s2), (s2,s0).  Every call receives the triangle's
// depths[] and outputs[] arrays and reads elements 0 and 1 of them, whichever edge it draws -- kept as is.
// Line fragments differ from filled ones: pixel centres at +0.5, depth = 1/lerp(d0,d1,t) (positive),
// written when alpha != 0, and no BlendMode.None row break.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void wire_pixel(const RasterArgs& A, const FrameParams& fp, const SrbUniforms& U, const TriRec& r,
                                        const uint32_t seq, const int x, const int y, const int depthTest,
                                        const int blendMode, float4& col, float& dep) {
    const float cx = add((float)x, 0.5f), cy = add((float)y, 0.5f);
#pragma unroll 1
    for (int k = 0; k < 3; ++k) {
        const float p0x = k == 0 ? r.s0x : k == 1 ? r.s1x : r.s2x, p0y = k == 0 ? r.s0y : k == 1 ? r.s1y : r.s2y;
        const float p1x = k == 0 ? r.s1x : k == 1 ? r.s2x : r.s0x, p1y = k == 0 ? r.s1y : k == 1 ? r.s2y : r.s0y;
        const int minX = sat_int(net_max(net_min(p0x, p1x), 0.0f));                 // :242-245
        const int maxX = sat_int(net_min(net_max(p0x, p1x), (float)(fp.W - 1)));
        const int minY = sat_int(net_max(net_min(p0y, p1y), 0.0f));
        const int maxY = sat_int(net_min(net_max(p0y, p1y), (float)(fp.H - 1)));
        if (x < minX || x > maxX || y < minY || y > maxY) continue;
        const float dx = sub(p1x, p0x), dy = sub(p1y, p0y);                         // :257-259
        const float lenSq = add(mul(dx, dx), mul(dy, dy));
        const float px = sub(cx, p0x), py = sub(cy, p0y);                           // :296-297
        float t = 0.0f;
        if (lenSq > 0.0f) t = fdiv(add(mul(px, dx), mul(py, dy)), lenSq);           // :299-301
        t = net_max(0.0f, net_min(1.0f, t));                                        // :303
        const float distX = sub(cx, add(p0x, mul(t, dx))), distY = sub(cy, add(p0y, mul(t, dy)));
        const float distSq = add(mul(distX, distX), mul(distY, distY));
        if (!(distSq <= 0.25f)) continue;                                           // :312-313
        const float omt = sub(1.0f, t);
        const float depth = fdiv(1.0f, add(mul(r.d0, omt), mul(r.d1, t)));          // :315
        if (!depth_pass(depthTest, depth, dep)) continue;                           // :318-319
        int dtmp;
        const float4 c = shade_fragment<true>(A, fp, U, r, seq, omt, t, 0.0f, dtmp);
        if (c.w != 0.0f) {                                                          // :325
            col = blend_px(c, col, blendMode);
            if (depthTest != SRB_DEPTH_DISABLED) dep = depth;                       // :331-332
        }
    }
}

// ---------------------------------------------------------------------------
// Ordered tile rasterizer: pixel loops of RasterizeTriangle (Rasterizer.cs:
// 466-536) replayed per pixel in submission order.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ordered_tile_kernel(RasterArgs A, FrameParams fp, SrbUniforms U) {
    const unsigned full = 0xffffffffu;
    if (A.cnt->overflow) return;
    const int tile = fp.bandRow0 * fp.tilesX + blockIdx.x;
    const int tX = tile % fp.tilesX, tY = tile / fp.tilesX;
    const int tx0 = tX * kTile, ty0 = tY * kTile;
    const int px = threadIdx.x & 15, py = threadIdx.x >> 4;
    const int x = tx0 + px, y = ty0 + py;
    const bool inb = x < fp.W && y < fp.H;
    const size_t pidx = (size_t)y * fp.W + x;
    float4 col = make_float4(fp.clearColor[0], fp.clearColor[1], fp.clearColor[2], fp.clearColor[3]);
    float dep = kClearDepth;
    if (inb && fp.loadColor) col = A.color[pidx];
    if (inb && fp.loadDepth) dep = A.depth[pidx];
    const int tileEndX = min(tx0 + kTile - 1, fp.W - 1), tileEndY = min(ty0 + kTile - 1, fp.H - 1);
    const uint32_t bs = A.tileStart[tile], be = A.tileEnd[tile];
    for (uint32_t ei = bs; ei < be; ++ei) {
        const uint32_t seq = __ldg(A.bins + ei);
        const TriRec r = load_rec(A.recs, seq);
        const TriEq e = tri_eq(r);
        const int startX = max(e.minX, tx0), endX = min(e.maxX, tileEndX);
        const int startY = max(e.minY, ty0), endY = min(e.maxY, tileEndY);
        if (startX > endX || startY > endY) continue;            // uniform over the CTA
        const int di = (r.meta & 0x80000000u) ? A.clipAttrs[r.meta & 0x7fffffffu].draw : (int)r.meta;
        const int depthTest = A.draws[di].depth_test, blendMode = A.draws[di].blend;
        if (fp.wireframe) {                                         // uniform over the frame
            if (inb && x >= startX && x <= endX && y >= startY && y <= endY)    // the lines' boxes lie inside the triangle's
                wire_pixel(A, fp, U, r, seq, x, y, depthTest, blendMode, col, dep);
            continue;
        }
        bool pass = false, writeOk = false;
        float depth = 0.0f;
        float4 c = make_float4(0, 0, 0, 0);
        if (inb && x >= startX && x <= endX && y >= startY && y <= endY) {
            float w0, w1, w2;
            edge_chain(r, e, startX, startY, x, y, w0, w1, w2);
            if (inside3(w0, w1, w2)) {
                float w0f, w1f, w2f;
                depth = frag_depth(r.d0, r.d1, r.d2, e.invArea, w0, w1, w2, w0f, w1f, w2f);
                if (depth_pass(depthTest, depth, dep)) {
                    pass = true;
                    int dtmp;
                    c = shade_fragment(A, fp, U, r, seq, w0f, w1f, w2f, dtmp);
                    writeOk = c.w > 0.0f;                                      // :511
                }
            }
        }
        // BlendMode.None: the first depth-passing fragment of a tile row with
        // alpha <= 0 (or NaN) ends that row for this triangle (:520-523).
        bool skipped = false;
        if (blendMode == SRB_BLEND_NONE) {
            unsigned brk = __ballot_sync(full, pass && !writeOk);
            unsigned rowBits = (brk >> (16 * ((threadIdx.x & 31) >> 4))) & 0xffffu;
            if (rowBits) skipped = px > (__ffs(rowBits) - 1);
        }
        if (pass && writeOk && !skipped) {
            col = blend_px(c, col, blendMode);
            if (depthTest != SRB_DEPTH_DISABLED) dep = depth;
        }
    }
    if (inb) {
        A.colorOut[pidx] = col;
        A.depth[pidx] = dep;
        if (A.color8Out) A.color8Out[pidx] = pack_unorm8(col);
    }
}

// ---------------------------------------------------------------------------
// Visibility tile rasterizer.
// MODE: 0 LessEqual (max depth, later wins ties), 1 Less (max, earlier wins),
//       2 GreaterEqual (min, later wins), 3 Greater (min, earlier wins).
// ---------------------------------------------------------------------------
#ifndef SRB_VIS_THREADS
#define SRB_VIS_THREADS 128
#endif
constexpr int kVisThreads = SRB_VIS_THREADS;
constexpr int kVisWarps = kVisThreads / 32;
constexpr int kVisFields = 15;
constexpr int kKeyPitch = kTile + 1;
#ifndef SRB_HIZ_FIRST
#define SRB_HIZ_FIRST 32
#endif
constexpr uint32_t kHiZFirst = SRB_HIZ_FIRST;          // entries of a long bin's first chunk (0 disables hi-Z chunking)     // 17 x 8 B = 34 banks: consecutive rows are two banks apart

// Walk one row of one triangle inside a tile: replay the reference's x accumulation
// (Rasterizer.cs:485-535) and raise the per-pixel winner keys.
// (Measured and rejected on city4k: a min3/max3 form of the inclusive test selected per row by a
// finiteness check, and walking small footprints in the loading thread -- both slower.)
template <int MODE>
__device__ __forceinline__ void vis_walk_row(float w0, float w1, float w2, const float a12, const float a20,
                                             const float a01, const float invArea, const float d0, const float d1,
                                             const float d2, const uint32_t seq, unsigned long long* krow,
                                             const int sx, const int ex) {
    // (Measured and rejected: separating "advance to the next covered pixel" from the depth/key part
    // so that lanes would run the expensive part together -- 139 -> 196 us, the lanes serialised.)
    // (Measured and rejected: a straight-line body that always computes depth and key -- 139 -> 154 us;
    // whole warps of uncovered pixels are common enough for the branch to pay.)
    SRB_ASSERT(sx >= 0 && ex < kTile);
    for (int xx = sx; xx <= ex; ++xx) {
        // (Measured and rejected: deciding by sign bits, (s0^s1)|(s0^s2) >= 0, with the exact compares only when
        // an edge value is zero -- half the ALU-pipe work but one more branch level: 119 -> 141 us.)
        if (inside3_minmax(w0, w1, w2)) {                // 113 -> 112 us; a NaN edge value falls out as a NaN depth below
            float w0f, w1f, w2f;
            const float depth = frag_depth(d0, d1, d2, invArea, w0, w1, w2, w0f, w1f, w2f);
            // NaN never passes a depth test: it becomes key 0, which loses to everything.  No pre-check of the
            // stored key: atomicMax itself reads and compares before its CAS.  (Measured: the branchy form with
            // a pre-check 119 us, this 117 us.  Rejected: a hand-written CAS loop reusing the loaded value,
            // 123 -> 260 us -- the compiler's ATOMS.CAST.SPIN sequence is far better.)
            const unsigned long long key = make_key<MODE>(depth, seq);
#ifdef SRB_EXP_NO_ATOMIC                         /* diagnostic builds only (tools/tile_breakdown.sh): wrong images */
            krow[xx] = depth == depth ? key : 0ull;
#else
            // (Measured and rejected, round 2: in the second and later chunks of a bin, a plain load of the stored key
            // first and no atomic when the fragment loses -- raster 111.2 -> 113.4 us, stress8k 1106 -> 1143.)
            atomicMax(krow + xx, depth == depth ? key : 0ull);
#endif
        }
        w0 = add(w0, a12); w1 = add(w1, a20); w2 = add(w2, a01);
    }
}

struct VisSmem {
    unsigned long long keys[kTile * kKeyPitch]; // rows padded: lanes walk DIFFERENT rows at the same x, and an
                                                // unpadded 128-byte row pitch would put them all on the same two banks
    float S[kVisFields][kVisThreads];      // raster constants of the chunk's triangles (SoA)
    int warpRows[kVisWarps];
    uint32_t warpMin[kVisWarps];           // hi-Z: per-warp minimum of the stored depth keys
    int last;                              // split bins: this CTA merged last and resolves the tile
    unsigned short items[kVisThreads * kTile];   // (triangle slot | row << 8) per work item of the chunk
#ifdef SRB_TILE_TIMING
    unsigned long long phase[2];           // diagnostic: time after the first chunk's load+scan, time before resolve
#endif
};

template <int MODE, bool OUT8>
__device__ __forceinline__ void vis_resolve(const unsigned long long* keys, const unsigned long long ckey,
                                            const int tx0, const int ty0, const float4 clearCol, const RasterArgs& A,
                                            const FrameParams& fp, const SrbUniforms& U) {
    // (Measured and rejected, round 2: "span resolve" -- the winner's edge values at a pixel are the end of a chain of
    // up to 15 + 15 rounded adds, replayed below for every pixel; pixels of one row that share a winner share the chain,
    // so one thread per SPAN (maximal run of equal winners in a row; found by ballot, compacted into shared memory)
    // walked it once exactly as the coverage pass does and left (w0, w1, w2, invArea) per pixel in shared memory for a
    // pixel-parallel shading pass that then needs only the 32-byte shading half of the record.  Same bits (37 GPU
    // parity tests), fewer instructions, and slower: raster city4k 102.6 -> 108.1 us, gun1080 25.4 -> 27.0, stress8k
    // 1111 -> 1158 (falling back to the replay above 32 / 96 / 256 spans made no difference on city4k).  Three more
    // CTA barriers and a serial span walk on a few threads cost more than 8 warps replaying short chains in lockstep.
    // profiles/r02zy_ab_span_resolve.txt)
    for (int p = threadIdx.x; p < kTile * kTile; p += kVisThreads) {
        const int px = p & 15, py = p >> 4;
        const int x = tx0 + px, y = ty0 + py;
        if (x >= fp.W || y >= fp.H) continue;
        const unsigned long long key = keys[py * kKeyPitch + px];
        float4 outc = clearCol;
        float outd = kClearDepth;
#ifdef SRB_EXP_SKIP_SHADE
        if (false) {
#else
        if (key != ckey) {
#endif
            uint32_t hi = (uint32_t)(key >> 32);
            const uint32_t seq = key_seq<MODE>(key);
            if (MODE >= 2) hi = ~hi;
            const TriRec r = load_rec(A.recs, seq);
            const TriEq e = tri_eq(r);
            const int startX = max(e.minX, tx0), startY = max(e.minY, ty0);
            float w0, w1, w2, w0f, w1f, w2f;
            edge_chain(r, e, startX, startY, x, y, w0, w1, w2);
            frag_depth(r.d0, r.d1, r.d2, e.invArea, w0, w1, w2, w0f, w1f, w2f);
            int di;
            float4 c = shade_fragment(A, fp, U, r, seq, w0f, w1f, w2f, di);
            // Eligibility guarantees alpha > 0 and alpha == 1 up to rounding, so
            // the blend destination is immaterial at 8 bits; use the clear colour.
            outc = blend_px(c, clearCol, A.draws[di].blend);
            outd = ord2f(hi);
        }
        const size_t pidx = (size_t)y * fp.W + x;
        __stcs(A.colorOut + pidx, outc);
        __stcs(A.depth + pidx, outd);
        if (OUT8) __stcs(A.color8Out + pidx, pack_unorm8(outc));
    }
}


// One 16x16 tile, whole CTA.  Ends with every thread past its last shared-memory access
// only after the caller's __syncthreads() (heavy-list CTAs loop over several tiles).
template <int MODE, bool OUT8>
#ifdef SRB_EXP_CAP_BIN
#define SRB_BE be_
#else
#define SRB_BE be
#endif
__device__ __forceinline__ void vis_process_tile(const int tile, const uint32_t bs, const uint32_t SRB_BE,
                                                 const uint32_t parts, const uint32_t slice, const bool hasBig,
                                                 const RasterArgs& A, const FrameParams& fp, const SrbUniforms& U,
                                                 VisSmem& sm) {
    unsigned long long* keys = sm.keys;
    float (*S)[kVisThreads] = sm.S;
    int* warpRows = sm.warpRows;
    const unsigned full = 0xffffffffu;
    const int tX = tile % fp.tilesX, tY = tile / fp.tilesX;
    const int tx0 = tX * kTile, ty0 = tY * kTile;
    const int tileEndX = min(tx0 + kTile - 1, fp.W - 1), tileEndY = min(ty0 + kTile - 1, fp.H - 1);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float4 clearCol = make_float4(fp.clearColor[0], fp.clearColor[1], fp.clearColor[2], fp.clearColor[3]);

#ifdef SRB_EXP_ALL_EMPTY                         /* diagnostic builds only: every tile takes the clear-only path */
    if (true) {
#else
    if (bs == SRB_BE) {
#endif
#ifdef SRB_EXP_SKIP_EMPTY
        return;
#endif
        // Empty bin: the fused clear is all there is (pure store bandwidth).
        for (int p = tid; p < kTile * kTile; p += kVisThreads) {
            const int x = tx0 + (p & 15), y = ty0 + (p >> 4);
            if (x < fp.W && y < fp.H) {
                const size_t pidx = (size_t)y * fp.W + x;
                __stcs(A.colorOut + pidx, clearCol);
                __stcs(A.depth + pidx, kClearDepth);
                if (OUT8) __stcs(A.color8Out + pidx, pack_unorm8(clearCol));
            }
        }
        return;
    }

#ifdef SRB_EXP_CAP_BIN
    const uint32_t be = min(be_, bs + (uint32_t)SRB_EXP_CAP_BIN);
#endif
    // (Measured and rejected, round 2: bins of <= 8 / 16 / 32 / 64 triangles rasterized pixel-parallel -- every
    // thread takes its two pixels through all triangles, replaying the accumulation chain up to ITS pixel, best key in
    // registers; no key array, scan, items or atomics, one barrier.  Raster 115 -> 119 / 123 / 135 / 151 us, gun1080
    // 25.7 -> 26.7 / 32.6 / 42.3 / 47.1: the chain up to a pixel is dy + dx rounded adds where the row walk pays one
    // per pixel, and the kernel is bound by instruction issue, not by the latency the variant removes.
    // profiles/r02k_ab_light_tile_path.txt)
    const unsigned long long ckey = clear_key<MODE>();
    for (int i = tid; i < kTile * kKeyPitch; i += kVisThreads) keys[i] = ckey;

    // Hi-Z inside the tile (depth functions Less / LessEqual).  Between chunks the CTA takes the smallest depth stored
    // in the tile so far; a triangle whose best possible depth over its rectangle (tri_depth_upper_bound, proven
    // conservative against the reference's rounded arithmetic) is below it cannot win a pixel and is not walked.  The
    // fill pass puts the triangles that span many tiles FIRST into every bin -- near walls and ground, the usual
    // occluders -- and a long bin's first chunk is kept short so that the bound exists before the crowd behind them is
    // looked at: on the horizon of the city frame half to four fifths of a heavy bin's triangles are hidden.
    // Only bins that hold such a triangle take part (a flag from the binning passes): crowds of small triangles alone
    // never close a tile early, and the extra chunk cost the 10 M-triangle frame 10 % when every long bin paid it.
    constexpr bool kHiZmode = MODE < 2 && kHiZFirst > 0u;
    const bool kHiZ = kHiZmode && hasBig;
    uint32_t minHi = 0u;                                   // smallest stored depth key (high word) of the tile: 0 = unknown
    uint32_t chunk = (kHiZ && be - bs > kHiZFirst + kHiZFirst / 2) ? kHiZFirst : (uint32_t)kVisThreads;
    for (uint32_t cb = bs; cb < be; cb += chunk, chunk = kVisThreads) {
        // ---- one triangle per thread: derive its raster constants for this tile ----
        int nrows = 0;
        if ((uint32_t)tid < chunk && cb + tid < be) {
            const uint32_t seq = __ldg(A.bins + cb + tid);
            const TriRec r = load_rec_raster(A.recs, seq);
            const TriEq e = tri_eq(r);
            const int startX = max(e.minX, tx0), endX = min(e.maxX, tileEndX);
            const int startY = max(e.minY, ty0), endY = min(e.maxY, tileEndY);
            if (startX <= endX && startY <= endY) {
                nrows = endY - startY + 1;
                float w0, w1, w2;
                row_base(r, e, startX, startY, w0, w1, w2);
                if (kHiZ && minHi != 0u &&
                    f2ord(tri_depth_upper_bound(r, e, w0, w1, w2, startX, endX, startY, endY)) < minHi)
                    nrows = 0;                             // hidden behind what the tile already holds
                S[0][tid] = w0; S[1][tid] = w1; S[2][tid] = w2;
                S[3][tid] = e.a12; S[4][tid] = e.a20; S[5][tid] = e.a01;
                S[6][tid] = e.b12; S[7][tid] = e.b20; S[8][tid] = e.b01;
                S[9][tid] = e.invArea;
                S[10][tid] = r.d0; S[11][tid] = r.d1; S[12][tid] = r.d2;
                S[13][tid] = __uint_as_float((uint32_t)(startX - tx0) | ((uint32_t)(endX - tx0) << 8) |
                                             ((uint32_t)(startY - ty0) << 16));
                S[14][tid] = __uint_as_float(seq);
            }
        }
        // ---- CTA-wide exclusive scan of the row counts ----
        int incl = nrows;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(full, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) warpRows[wid] = incl;
        __syncthreads();
        int woff = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kVisWarps; ++w) {
            const int v = warpRows[w];
            if (w < wid) woff += v;
            total += v;
        }
        {   // expand the work items: cheaper than a search per item (123 -> 121 us)
            unsigned short* it = sm.items + (woff + incl - nrows);
            SRB_ASSERT(nrows >= 0 && nrows <= kTile && woff + incl <= kVisThreads * kTile);
            for (int r = 0; r < nrows; ++r) it[r] = (unsigned short)(tid | (r << 8));
        }
        __syncthreads();
#ifdef SRB_TILE_TIMING
        if (tid == 0 && cb == bs) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(sm.phase[0]));
#endif
        // ---- (triangle,row) work items spread over the whole CTA.  Consecutive lanes take consecutive rows
        // of the same triangle, which keeps a warp's row walks coherent.  (Measured and rejected: one
        // contiguous run of items per thread, reusing constants and the y accumulation -- 123 -> 190 us.) ----
        // (Measured, no gain: splitting each row over 2 or 4 threads when a chunk has few items.)
#ifdef SRB_EXP_NO_WALK                           /* diagnostic builds only */
        for (int i = tid; i < 0; i += kVisThreads) {
#else
        for (int i = tid; i < total; i += kVisThreads) {
#endif
            const int itv = sm.items[i];
            const int j = itv & 0xff, dy = itv >> 8;
            float w0 = S[0][j], w1 = S[1][j], w2 = S[2][j];
            const float a12 = S[3][j], a20 = S[4][j], a01 = S[5][j];
            const float b12 = S[6][j], b20 = S[7][j], b01 = S[8][j];
            const float invArea = S[9][j], d0 = S[10][j], d1 = S[11][j], d2 = S[12][j];
            const uint32_t pk = __float_as_uint(S[13][j]);
            const uint32_t seq = __float_as_uint(S[14][j]);
            const int sx = pk & 0xff, ex = (pk >> 8) & 0xff, row = (int)(pk >> 16) + dy;
            SRB_ASSERT(j < kVisThreads && row >= 0 && row < kTile);
#pragma unroll 4
            for (int k = 0; k < dy; ++k) { w0 = add(w0, b12); w1 = add(w1, b20); w2 = add(w2, b01); }
            vis_walk_row<MODE>(w0, w1, w2, a12, a20, a01, invArea, d0, d1, d2, seq, keys + row * kKeyPitch, sx, ex);
        }
        // (Measured, no effect: loading the next chunk's bin entries early and prefetching their records.)
        __syncthreads();
        if (kHiZ && cb + chunk < be) {                     // more chunks follow: the tile's smallest stored depth so far
            uint32_t m = 0xffffffffu;
            for (int p = tid; p < kTile * kTile; p += kVisThreads) {
                const int px = p & 15, py = p >> 4;
                if (tx0 + px < fp.W && ty0 + py < fp.H) m = min(m, (uint32_t)(keys[py * kKeyPitch + px] >> 32));
            }
            m = __reduce_min_sync(full, m);
            if (lane == 0) sm.warpMin[wid] = m;
            __syncthreads();
            m = sm.warpMin[0];
#pragma unroll
            for (int w = 1; w < kVisWarps; ++w) m = min(m, sm.warpMin[w]);
            minHi = m == (uint32_t)(ckey >> 32) ? 0u : m;  // an uncovered pixel holds the clear key: nothing is below it
        }
    }

#ifdef SRB_TILE_TIMING
    if (tid == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(sm.phase[1]));
#endif
    if (parts > 1u) {
        // One of several CTAs of a split bin: merge this part's keys into the bin's global slice; whoever arrives
        // last takes the merged keys back (leaving the slice and its counter zero for the next user) and resolves.
        SRB_ASSERT(slice < kMaxExtra && parts <= kMaxExtra + 1u);
        unsigned long long* gk = A.gkeys + (size_t)slice * (kTile * kTile);
        for (int i = tid; i < kTile * kTile; i += kVisThreads) {
            const unsigned long long k = keys[(i >> 4) * kKeyPitch + (i & 15)];
            if (k != ckey) atomicMax(gk + i, k);
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) sm.last = atomicAdd(A.sliceDone + slice, 1u) == parts - 1u;
        __syncthreads();
        if (!sm.last) return;
        __threadfence();
        for (int i = tid; i < kTile * kTile; i += kVisThreads) {
            const unsigned long long k = __ldcg(gk + i);             // written by other SMs' atomics: not through L1
            gk[i] = 0ull;
            keys[(i >> 4) * kKeyPitch + (i & 15)] = max(k, ckey);
        }
        if (tid == 0) A.sliceDone[slice] = 0u;
        __syncthreads();
    }
    // ---- resolve: shade each pixel's winner once, write colour + depth ----
    // (Measured and rejected, round 2 -- the one TMA / bulk-async experiment the path offers: the finished tile staged
    // in shared memory (the dead raster-constant array) and written back by the bulk-copy engine, 16 colour rows of
    // 256 B + 16 depth rows of 64 B per tile with cp.async.bulk.global.shared::cta (UBLKCP.G.S in SASS,
    // profiles/r02t_sass_bulk_store_experiment_UBLKCP.txt), instead of per-thread 128-bit streaming stores.  Same
    // images (87 GPU tests), slower everywhere: raster city4k 112.7 -> 129.9 us, gun1080 25.0 -> 27.1, stress8k
    // 1103 -> 1243 (profiles/r02t_ab_bulk_async_tile_store.txt).  The stores were never the cost -- they leave the
    // thread at issue --, while the staging adds a barrier, a proxy fence and 32 serial bulk issues to every tile.
    // The tile's INPUTS are gathers of 48-byte records by index, nothing a tensor map describes.)
    vis_resolve<MODE, OUT8>(keys, ckey, tx0, ty0, clearCol, A, fp, U);
}

#ifndef SRB_VIS_MINBLOCKS
#define SRB_VIS_MINBLOCKS 8
#endif
// OUT8: the frame is also stored at display precision into A.color8Out (RGBA8 band gather); a template parameter
// because a run-time test of the pointer cost the default path 2 us.
template <int MODE, bool OUT8>
__global__ void __launch_bounds__(kVisThreads, SRB_VIS_MINBLOCKS)
vis_tile_kernel(RasterArgs A, FrameParams fp, SrbUniforms U) {
    __shared__ VisSmem sm;
    // (Measured and rejected: a persistent grid of smCount x 8 CTAs taking launch-order entries from a device-side
    // queue, each CTA fetching its next tile's order entry and first bin entries and prefetching their records
    // while the current tile is shaded -- raster 113 -> 124 us, and resident CTAs that never retire keep the
    // pipelined geometry kernels of the next frame off the SMs: city4k 5,470 -> 4,850 frames/s.)
    // (Measured and rejected as well: M tiles per CTA at a fixed stride through the launch order, look-ahead fetches
    // included, M = 2..8: raster 115 -> 117..124 us.)
    // One load gives the tile and its bin range.  (Measured and rejected: clearing the frame on a second
    // stream beside the geometry stages -- at full store bandwidth it slows them by its own duration.)
    // The extra entries of split bins (heavy work, behind the one-per-tile entries in memory) are launched first.
    const bool extra = blockIdx.x < A.extraCtas;
    // (Measured and rejected, round 2: the order entry carrying the tile as column | row << 16 to spare every CTA the
    // division below -- the tile kernel did not move, the scan kernel paid 2.4 us for the division.
    // profiles/r02r_ab_txy_precheck.txt)
    const uint4 o = __ldg(A.order + (extra ? (uint32_t)A.bandTiles + blockIdx.x : blockIdx.x - A.extraCtas));
    if (A.cnt->overflow) return;
    if (extra && blockIdx.x >= A.cnt->nExtra) return;       // unused slot (stale entry of an earlier frame)
    SRB_ASSERT((int)o.x >= fp.bandRow0 * fp.tilesX && (int)o.x < fp.bandRow1 * fp.tilesX && o.y <= o.z && o.z <= fp.pairCap);
#ifdef SRB_TILE_TIMING
    unsigned long long t0, t1;
    unsigned smid;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
#endif
    vis_process_tile<MODE, OUT8>((int)o.x, o.y, o.z, o.w & 0xffffu, (o.w >> 16) & 0x7fffu, (o.w >> 31) != 0u, A, fp, U, sm);
#ifdef SRB_TILE_TIMING
    __syncthreads();
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
    if (threadIdx.x == 0 && A.dbg) {
        const unsigned long long pa = o.y == o.z ? 0ull : min(sm.phase[0] - t0, 0xfffffull);
        const unsigned long long pb = o.y == o.z ? 0ull : min(sm.phase[1] - t0, 0xfffffull);
        A.dbg[blockIdx.x] = make_ulonglong4(t0, t1, ((unsigned long long)o.x << 32) | (o.z - o.y), smid | (pa << 8) | (pb << 28));
    }
#endif
}

// ---------------------------------------------------------------------------
// Frame barrier between the band ranks of a multi-GPU frame (one rank per GPU), in peer memory.  Every rank owns a
// row of flags, one per rank; barrier k of rank r stores k into flag r of every rank's row (peer stores over NVLink,
// after a system-scope fence: this rank's tile kernel, which precedes the barrier on the stream, stored its band
// into rank 0's frame the same way), then waits until every flag of its own row has reached `waitFor` -- k itself
// when the caller reads the assembled frame next, k - 1 when the ranks only have to stay within a frame of each
// other.  One 32-thread launch per frame instead of a 4-byte NCCL all-reduce.  The wait is bounded: after
// kBarrierTimeoutNs the kernel gives up and raises *timedOut (mapped host memory), so a missing rank becomes an
// error code, never a hung device.
// ---------------------------------------------------------------------------
constexpr int kMaxRanks = 16;
constexpr int kFlagStride = 32;                       // uint32 per flag: every flag on a 128-byte line of its own
constexpr unsigned long long kBarrierTimeoutNs = 4000000000ull;
struct BarrierRows { uint32_t* row[kMaxRanks]; };
__global__ void frame_barrier_kernel(BarrierRows rows, int rank, int world, uint32_t k, uint32_t waitFor,
                                     uint32_t* __restrict__ timedOut) {
    const int t = threadIdx.x;
    __threadfence_system();
    if (t < world) *reinterpret_cast<volatile uint32_t*>(rows.row[t] + rank * kFlagStride) = k;
    __threadfence_system();
    if (t < world && waitFor != 0u) {
        const volatile uint32_t* f = rows.row[rank] + t * kFlagStride;
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while ((int32_t)(*f - waitFor) < 0) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > kBarrierTimeoutNs) { *timedOut = 1u; break; }
        }
    }
    __threadfence_system();
}

__global__ void texture_sample_kernel(const uchar4* __restrict__ tex, int tw, int th, const float2* __restrict__ uv,
                                      int n, int filter, float4* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float2 t = uv[i];
    out[i] = filter == 1 ? tex_bilinear(tex, tw, th, t.x, t.y) : tex_nearest(tex, tw, th, t.x, t.y);
}

// Present-format step (MainWindow.cs:234-250 converts the float buffer for the
// GL upload): float4 -> RGBA8, clamp to [0,1], round to nearest.
__global__ void pack_rgba8_kernel(const float4* __restrict__ color, uchar4* __restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 c = color[i];
    auto q = [](float v) { v = fminf(fmaxf(v, 0.0f), 1.0f); return (unsigned char)__float2int_rn(v * 255.0f); };
    out[i] = make_uchar4(q(c.x), q(c.y), q(c.z), q(c.w));
}

// The same step at display precision: the reference draws its RGBA32F texture into the window's 8-bit
// framebuffer (MainWindow.cs:251-253), i.e. clamp to [0,1] and round to the nearest of 255 steps -- the format
// of the parity gate.  Four pixels per thread: 4 x 128-bit loads, one 128-bit store.
__global__ void present_rgba8_kernel(const float4* __restrict__ color, uint32_t* __restrict__ out, size_t n) {
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = q * 4;
    if (i + 3 < n) {
        const float4 a = color[i], b = color[i + 1], c = color[i + 2], d = color[i + 3];
        __stcs(reinterpret_cast<uint4*>(out) + q, make_uint4(pack_unorm8(a), pack_unorm8(b), pack_unorm8(c), pack_unorm8(d)));
    } else {
        for (size_t k = i; k < n; ++k) out[k] = pack_unorm8(color[k]);
    }
}

// MainWindow.OnRender's flatten (MainWindow.cs:232-238): Vector4 colour -> packed Vector3, the layout the
// reference hands to TexSubImage2D(Rgb, Float).  Four pixels per thread: 4 x 128-bit loads, 3 x 128-bit stores.
__global__ void present_rgb_kernel(const float4* __restrict__ color, float* __restrict__ out, size_t n) {
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t i = q * 4;
    if (i + 3 < n) {
        const float4 a = color[i], b = color[i + 1], c = color[i + 2], d = color[i + 3];
        float4* o = reinterpret_cast<float4*>(out) + q * 3;
        __stcs(o, make_float4(a.x, a.y, a.z, b.x));
        __stcs(o + 1, make_float4(b.y, b.z, c.x, c.y));
        __stcs(o + 2, make_float4(c.z, d.x, d.y, d.z));
    } else {
        for (size_t k = i; k < n; ++k) {
            const float4 a = color[k];
            out[3 * k] = a.x; out[3 * k + 1] = a.y; out[3 * k + 2] = a.z;
        }
    }
}

}  // namespace srb
