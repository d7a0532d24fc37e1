// srb_device.cuh -- device-side data layout and exact-arithmetic helpers.
//
// Everything that feeds coverage or depth is written with explicit
// round-to-nearest intrinsics (__fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn), which
// ptxas never contracts into FMA, so the results are the same IEEE single
// operations RyuJIT emits for the reference's scalar C# (SURVEY.md A.5-A.7).
// The TU is additionally compiled with -fmad=false.  Fused operations appear
// only where System.Numerics fuses (numerics_fma switch).
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/srb200.h"

// -DSRB_CHECKED: bounds checks of every computed index (bins, records, work items, key rows, key slices, vertex and
// triangle indices) as device asserts.  compute-sanitizer is closed on this pool, so the whole GPU test suite is run
// against a checked build instead (tools/checked_suite.sh); a release build compiles the checks away.
#ifdef SRB_CHECKED
#include <assert.h>
#define SRB_ASSERT(c) assert(c)
#else
#define SRB_ASSERT(c) ((void)0)
#endif

namespace srb {

constexpr int kTile = 16;                 // Rasterizer.cs:53
constexpr float kEps = 1e-6f;             // Rasterizer.cs:52
constexpr float kClearDepth = -3.402823466e+38f;   // float.MinValue, MainWindow.cs:434

// ---- per-frame draw record (device copy of one RenderMesh call) ----
struct DrawRec {
    float model[16];
    float view[16];
    float proj[16];
    float sphere[4];
    const float4*   vtx;      // mesh vertices, 3 float4 per vertex (48 B, Shaders.cs:10-24)
    const uint16_t* idx;
    const uchar4*   tex;      // nullptr = untextured
    int32_t n_vertices;
    int32_t n_triangles;
    int32_t vbase;            // first slot in the transformed-vertex arrays
    int32_t tbase;            // first global triangle id
    int32_t tex_w, tex_h;
    int32_t cull, depth_test, blend;
    uint32_t flags;
    int32_t visible;          // written by the frustum kernel
    int32_t _pad;
    const float4* tblockBoxes;    // object-space bounding box (centre, half extents: 2 float4) of every run of kGeomBlock triangles (mesh upload)
    const int2* tblockVRange;     // lowest / highest vertex index used by every run of kGeomBlock triangles (mesh upload)
    float4 planes[5];         // written by the frustum kernel: object-space half-spaces left/right/top/bottom of the
                              // screen (top/bottom: of this context's band), nx,ny,nz,d normalised; inside: n.p + d >= 0.
                              // planes[4] is the half-space clip w > 0: the four screen planes describe where a vertex
                              // lands on screen only for w > 0 (for w < 0 the NDC position flips sides), so a block is
                              // culled only when its box also lies entirely inside this one.
};
static_assert(sizeof(DrawRec) == 384, "DrawRec layout");

// ---- triangle setup record, slot = 2*global_triangle + fan ----
// s = screen coords of outputs {v2,v1,v0} (Rasterizer.cs:367,383-386),
// d = depths (:388), cz/cw = clip z/w of the same three corners, invArea = 1f/area (:427),
// bbox = the clamped integer bounding box (:432-442).
// In memory the record is split: 48 B the coverage pass reads (RecStore::R: s, d, invArea,
// bbox) and 32 B only shading reads (RecStore::S: cz, cw, meta).
struct TriRec {
    float s0x, s0y, s1x, s1y;
    float s2x, s2y, d0, d1;
    float d2, invArea;
    uint32_t bx, by;          // minX | maxX << 16,  minY | maxY << 16
    float cz0, cz1, cz2, cw0;
    float cw1, cw2;
    uint32_t meta;            // bit31: clipped (low bits = side-buffer index); else draw id
};
struct RecStore {
    float4* R;                // 3 float4 per slot
    float4* S;                // 2 float4 per slot
    uint32_t fan1Base;        // first slot of the second fan triangles (= total triangle count of the frame)
};
#define SRB_ASSERT_SLOT(st, seq) SRB_ASSERT((seq >> 1) < (st).fan1Base)
// seq = 2*global_triangle + fan.  Fan 0 (every unclipped triangle) lives at slot global_triangle, so that a warp of
// the setup kernel writes one contiguous run of records and the records a tile gathers sit densely in their lines;
// the rare second triangle of a near-clipped fan lives at fan1Base + global_triangle.  (With slot = seq every other
// slot was a hole: 48-byte records at a 96-byte pitch, i.e. two half-used sectors per record in both directions.)
__device__ __forceinline__ size_t rec_slot(const RecStore& st, uint32_t seq) {
    return (size_t)(seq >> 1) + ((seq & 1u) ? (size_t)st.fan1Base : 0);
}
__device__ __forceinline__ void store_rec(const RecStore& st, uint32_t seq, const TriRec& r) {
    SRB_ASSERT_SLOT(st, seq);
    float4* q = st.R + 3 * rec_slot(st, seq);
    q[0] = make_float4(r.s0x, r.s0y, r.s1x, r.s1y);
    q[1] = make_float4(r.s2x, r.s2y, r.d0, r.d1);
    q[2] = make_float4(r.d2, r.invArea, __uint_as_float(r.bx), __uint_as_float(r.by));
    float4* t = st.S + 2 * rec_slot(st, seq);
    t[0] = make_float4(r.cz0, r.cz1, r.cz2, r.cw0);
    t[1] = make_float4(r.cw1, r.cw2, __uint_as_float(r.meta), 0.0f);
}
// Coverage part only (cz/cw/meta left unset).
__device__ __forceinline__ TriRec load_rec_raster(const RecStore& st, uint32_t seq) {
    SRB_ASSERT_SLOT(st, seq);
    const float4* q = st.R + 3 * rec_slot(st, seq);
    float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    TriRec r;
    r.s0x = a.x; r.s0y = a.y; r.s1x = a.z; r.s1y = a.w;
    r.s2x = b.x; r.s2y = b.y; r.d0 = b.z; r.d1 = b.w;
    r.d2 = c.x; r.invArea = c.y; r.bx = __float_as_uint(c.z); r.by = __float_as_uint(c.w);
    return r;
}
__device__ __forceinline__ TriRec load_rec(const RecStore& st, uint32_t seq) {
    TriRec r = load_rec_raster(st, seq);
    const float4* t = st.S + 2 * rec_slot(st, seq);
    float4 a = __ldg(t), b = __ldg(t + 1);
    r.cz0 = a.x; r.cz1 = a.y; r.cz2 = a.z; r.cw0 = a.w;
    r.cw1 = b.x; r.cw2 = b.y; r.meta = __float_as_uint(b.z);
    return r;
}

// Varyings of a near-clipped triangle (Shaders.Lerp outputs), 3 corners in output order.
struct __align__(16) ClipAttr {
    float uv[3][2];
    float col[3][4];
    float wn[3][3];
    float wpos[3][3];         // EXTENSION (point lights only): world positions, Lerp'd like the other varyings
    int32_t draw;
};

// Compact list entry for accepted triangles: seq + tile range.
struct __align__(16) TriRef {
    uint32_t seq;             // 2*global_triangle + fan
    uint32_t tx;              // tminx | tmaxx << 16
    uint32_t ty;              // tminy | tmaxy << 16
    uint32_t _pad;
};

// ---- .NET semantics helpers ----
__device__ __forceinline__ float net_max(float a, float b) {   // MathF.Max: NaN-propagating
    if (a != a) return a;
    if (b != b) return b;
    return a > b ? a : b;
}
__device__ __forceinline__ float net_min(float a, float b) {
    if (a != a) return a;
    if (b != b) return b;
    return a < b ? a : b;
}
__device__ __forceinline__ float net_clamp(float v, float lo, float hi) {   // Math.Clamp
    if (v < lo) return lo;
    if (v > hi) return hi;
    return v;
}
// .NET 9 float->int conversion saturates and maps NaN to 0: exactly cvt.rzi.s32.f32.
__device__ __forceinline__ int sat_int(float f) { return __float2int_rz(f); }

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }

// Vector4.Transform(v, M), row-vector convention, one output column.
template <bool FMA>
__device__ __forceinline__ float xform_col(float x, float y, float z, float w, const float* m, int c) {
    float acc = mul(m[c], x);
    if (FMA) {
        acc = __fmaf_rn(m[4 + c], y, acc);
        acc = __fmaf_rn(m[8 + c], z, acc);
        acc = __fmaf_rn(m[12 + c], w, acc);
    } else {
        acc = add(acc, mul(m[4 + c], y));
        acc = add(acc, mul(m[8 + c], z));
        acc = add(acc, mul(m[12 + c], w));
    }
    return acc;
}
template <bool FMA>
__device__ __forceinline__ float4 transform4(float4 v, const float* m) {
    return make_float4(xform_col<FMA>(v.x, v.y, v.z, v.w, m, 0), xform_col<FMA>(v.x, v.y, v.z, v.w, m, 1),
                       xform_col<FMA>(v.x, v.y, v.z, v.w, m, 2), xform_col<FMA>(v.x, v.y, v.z, v.w, m, 3));
}
// Vector3.TransformNormal(n, M), one output column.
template <bool FMA>
__device__ __forceinline__ float xform_normal_col(float x, float y, float z, const float* m, int c) {
    float acc = mul(m[c], x);
    if (FMA) {
        acc = __fmaf_rn(m[4 + c], y, acc);
        acc = __fmaf_rn(m[8 + c], z, acc);
    } else {
        acc = add(acc, mul(m[4 + c], y));
        acc = add(acc, mul(m[8 + c], z));
    }
    return acc;
}
__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz) {
    return add(add(mul(ax, bx), mul(ay, by)), mul(az, bz));
}
// Vector{2,3,4}.Lerp per component.
__device__ __forceinline__ float lerp1(float a, float b, float t, bool fma) {
    float omt = sub(1.0f, t);
    float bt = mul(b, t);
    return fma ? __fmaf_rn(a, omt, bt) : add(mul(a, omt), bt);
}

// Rasterizer.EdgeFunction (Rasterizer.cs:561-563).
__device__ __forceinline__ float edge_fn(float ax, float ay, float bx, float by, float cx, float cy) {
    return sub(mul(sub(cx, ax), sub(by, ay)), mul(sub(cy, ay), sub(bx, ax)));
}

// Monotone map float -> uint32 (greater float <=> greater uint); -0 canonicalised to +0.
__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f + 0.0f);
    return u ^ ((uint32_t)((int)u >> 31) | 0x80000000u);     // negative: ~u, else u | sign
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    return __uint_as_float(u);
}

// ---- per-pixel winner keys of the visibility path ----
// MODE: 0 LessEqual (max depth, later wins ties), 1 Less (max, earlier wins),
//       2 GreaterEqual (min, later wins), 3 Greater (min, earlier wins).
// A valid key is never 0 (low word >= 1 in the even modes, depth bits never all-zero with ~seq in the odd ones
// for any seq < 2^31 - 1): 0 is the "nothing here" value of the global key slices of split bins.
template <int MODE>
__device__ __forceinline__ unsigned long long make_key(float depth, uint32_t seq) {
    uint32_t hi = f2ord(depth);
    if (MODE >= 2) hi = ~hi;
    const uint32_t lo = (MODE & 1) ? (0xfffffffeu - seq) : (seq + 1u);
    return ((unsigned long long)hi << 32) | lo;
}
template <int MODE>
__device__ __forceinline__ unsigned long long clear_key() {
    uint32_t hi = f2ord(kClearDepth);
    if (MODE >= 2) hi = ~hi;
    const uint32_t lo = (MODE & 1) ? 0xffffffffu : 0u;
    return ((unsigned long long)hi << 32) | lo;
}
template <int MODE>
__device__ __forceinline__ uint32_t key_seq(unsigned long long key) {
    const uint32_t lo = (uint32_t)key;
    return (MODE & 1) ? (0xfffffffeu - lo) : (lo - 1u);
}

// GetDepthTestFunction (Rasterizer.cs:542-559), flipped compares included.
__device__ __forceinline__ bool depth_pass(int test, float nd, float od) {
    switch (test) {
        case SRB_DEPTH_LESS: return nd > od;
        case SRB_DEPTH_LESS_EQUAL: return nd >= od;
        case SRB_DEPTH_GREATER: return nd < od;
        case SRB_DEPTH_GREATER_EQUAL: return nd <= od;
        case SRB_DEPTH_EQUAL: return fabsf(sub(nd, od)) < kEps;
        case SRB_DEPTH_NOT_EQUAL: return fabsf(sub(nd, od)) >= kEps;
        default: return true;
    }
}

// Rasterizer.Blend (Rasterizer.cs:58-65).
__device__ __forceinline__ float4 blend_px(float4 s, float4 d, int mode) {
    switch (mode) {
        case SRB_BLEND_ALPHA: {
            float a = s.w, ia = sub(1.0f, s.w);
            return make_float4(add(mul(s.x, a), mul(d.x, ia)), add(mul(s.y, a), mul(d.y, ia)),
                               add(mul(s.z, a), mul(d.z, ia)), add(mul(s.w, a), mul(d.w, ia)));
        }
        case SRB_BLEND_ADDITIVE: {
            float x = add(s.x, d.x), y = add(s.y, d.y), z = add(s.z, d.z), w = add(s.w, d.w);
            return make_float4(x < 1.0f ? x : 1.0f, y < 1.0f ? y : 1.0f, z < 1.0f ? z : 1.0f, w < 1.0f ? w : 1.0f);
        }
        case SRB_BLEND_MULTIPLY:
            return make_float4(mul(s.x, d.x), mul(s.y, d.y), mul(s.z, d.z), mul(s.w, d.w));
        default: return s;
    }
}

// Per-triangle raster constants derived from a TriRec (Rasterizer.cs:427-447).
struct TriEq {
    float a01, b01, a12, b12, a20, b20;
    float invArea;
    int minX, maxX, minY, maxY;
};
__device__ __forceinline__ void tri_bbox(const TriRec& r, int W, int H, int& minX, int& maxX, int& minY, int& maxY) {
    float minXf = net_min(net_min(r.s0x, r.s1x), r.s2x);
    float maxXf = net_max(net_max(r.s0x, r.s1x), r.s2x);
    float minYf = net_min(net_min(r.s0y, r.s1y), r.s2y);
    float maxYf = net_max(net_max(r.s0y, r.s1y), r.s2y);
    minX = max(sat_int(floorf(minXf)), 0);            // :437-440
    maxX = min(sat_int(ceilf(maxXf)), W - 1);
    minY = max(sat_int(floorf(minYf)), 0);
    maxY = min(sat_int(ceilf(maxYf)), H - 1);
}
__device__ __forceinline__ TriEq tri_eq(const TriRec& r) {
    TriEq e;
    e.invArea = r.invArea;
    e.a01 = sub(r.s0y, r.s1y); e.b01 = sub(r.s1x, r.s0x);   // :445-447
    e.a12 = sub(r.s1y, r.s2y); e.b12 = sub(r.s2x, r.s1x);
    e.a20 = sub(r.s2y, r.s0y); e.b20 = sub(r.s0x, r.s2x);
    e.minX = (int)(r.bx & 0xffffu); e.maxX = (int)(r.bx >> 16);
    e.minY = (int)(r.by & 0xffffu); e.maxY = (int)(r.by >> 16);
    return e;
}
// Row bases at (startX, startY) by the direct formula (Rasterizer.cs:481-483).
__device__ __forceinline__ void row_base(const TriRec& r, const TriEq& e, int startX, int startY,
                                         float& w0, float& w1, float& w2) {
    float fx = (float)startX, fy = (float)startY;
    w0 = add(mul(e.a12, sub(fx, r.s1x)), mul(e.b12, sub(fy, r.s1y)));
    w1 = add(mul(e.a20, sub(fx, r.s2x)), mul(e.b20, sub(fy, r.s2y)));
    w2 = add(mul(e.a01, sub(fx, r.s0x)), mul(e.b01, sub(fy, r.s0y)));
}
// centerInside (Rasterizer.cs:493-494): inclusive on every edge, either sign.
__device__ __forceinline__ bool inside3(float w0, float w1, float w2) {
    // bitwise, not short-circuit: six predicated compares and no branches (NaN compares false either way)
    const bool ge = (w0 >= 0.0f) & (w1 >= 0.0f) & (w2 >= 0.0f);
    const bool le = (w0 <= 0.0f) & (w1 <= 0.0f) & (w2 <= 0.0f);
    return ge | le;
}
// The same decision for every input without a NaN, as two three-input min/max and two compares
// (FMNMX3 on sm_100).  min/max skip a NaN operand where the compares above would all be false, so a NaN
// edge value can be reported inside: the caller's interpolated depth is then NaN and must be rejected
// there (which is what the reference's all-false compares amount to).
__device__ __forceinline__ bool inside3_minmax(float w0, float w1, float w2) {
    return (fminf(fminf(w0, w1), w2) >= 0.0f) | (fmaxf(fmaxf(w0, w1), w2) <= 0.0f);
}
// depth at a covered pixel (Rasterizer.cs:498-502).
__device__ __forceinline__ float frag_depth(float d0, float d1, float d2, float invArea,
                                            float w0, float w1, float w2,
                                            float& w0f, float& w1f, float& w2f) {
    w0f = mul(w0, invArea); w1f = mul(w1, invArea); w2f = mul(w2, invArea);
    return add(add(mul(d0, w0f), mul(d1, w1f)), mul(d2, w2f));
}

// ---------------------------------------------------------------------------
// Conservative triangle-vs-tile rejection (used when binning triangles that span many tiles).
// Returns false only when NO pixel of [X0..X1]x[Y0..Y1] can pass the reference's inclusive test
// (Rasterizer.cs:493-494) under its rounded, accumulated edge values.
//
// Proof sketch.  Let W_i(x,y) = a_i*(x-p_ix) + b_i*(y-p_iy) in real arithmetic on the float
// constants a_i, b_i, p_i.  Every value the reference forms for a pixel of the region -- the
// direct row base (:481-483) and each partial sum of its <= 15 + 15 rounded adds (:527-534) --
// is W_i at a lattice point of the region up to accumulated rounding, so it is bounded by
// B_i = max|a_i*(x-p_ix)| + max|b_i*(y-p_iy)| and the <= 35 roundings on the way add at most
// 35*2^-24*B_i.  The corner values below are computed with 3 roundings (<= 3*2^-24*B_i) and
// W_i is affine, so its extremes over the region are at the corners.  With E_i = 2^-16*B_i
// (six times the sum of both error terms):
//   max_corners W_i < -E_i for some i   =>  the reference's w_i < 0 everywhere: "all >= 0" fails;
//   min_corners W_j > +E_j for some j   =>  "all <= 0" fails;
//   min_corners (W_0+W_1+W_2) > sum E   =>  the three cannot all be <= 0 ("all <= 0" fails),
//   max_corners (W_0+W_1+W_2) < -sum E  =>  "all >= 0" fails.
// Both alternatives of the test failing means no pixel of the region is covered.  Any NaN/Inf
// or huge magnitude makes every comparison false => the tile is kept.
// ---------------------------------------------------------------------------
__device__ __forceinline__ bool tile_may_cover(float s0x, float s0y, float s1x, float s1y, float s2x, float s2y,
                                               int X0, int X1, int Y0, int Y1) {
    const float fx0 = (float)X0, fx1 = (float)X1, fy0 = (float)Y0, fy1 = (float)Y1;
    const float a[3] = {sub(s1y, s2y), sub(s2y, s0y), sub(s0y, s1y)};     // a12, a20, a01
    const float b[3] = {sub(s2x, s1x), sub(s0x, s2x), sub(s1x, s0x)};     // b12, b20, b01
    const float px[3] = {s1x, s2x, s0x}, py[3] = {s1y, s2y, s0y};
    float sum[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    float sumE = 0.0f;
    bool someNeg = false, somePos = false, sane = true;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float ax0 = mul(a[i], sub(fx0, px[i])), ax1 = mul(a[i], sub(fx1, px[i]));
        const float by0 = mul(b[i], sub(fy0, py[i])), by1 = mul(b[i], sub(fy1, py[i]));
        const float w00 = add(ax0, by0), w10 = add(ax1, by0), w01 = add(ax0, by1), w11 = add(ax1, by1);
        const float B = add(fmaxf(fabsf(ax0), fabsf(ax1)), fmaxf(fabsf(by0), fabsf(by1)));
        const float E = mul(B, 1.52587890625e-05f);                       // 2^-16
        sane = sane && (B < 1e30f);                                       // false for NaN as well
        const float wmax = fmaxf(fmaxf(w00, w10), fmaxf(w01, w11));
        const float wmin = fminf(fminf(w00, w10), fminf(w01, w11));
        someNeg = someNeg || (wmax < -E);
        somePos = somePos || (wmin > E);
        sum[0] = add(sum[0], w00); sum[1] = add(sum[1], w10); sum[2] = add(sum[2], w01); sum[3] = add(sum[3], w11);
        sumE = add(sumE, E);
    }
    if (!sane) return true;
    const float smax = fmaxf(fmaxf(sum[0], sum[1]), fmaxf(sum[2], sum[3]));
    const float smin = fminf(fminf(sum[0], sum[1]), fminf(sum[2], sum[3]));
    const bool geImpossible = someNeg || (smax < -sumE);
    const bool leImpossible = somePos || (smin > sumE);
    return !(geImpossible && leImpossible);
}

// ---------------------------------------------------------------------------
// Occlusion bound (hi-Z inside a tile, depth functions Less / LessEqual: the LARGEST stored depth wins).
// Returns an upper bound of every depth value the reference can compute for this triangle at a COVERED pixel of the
// rectangle [X0..X1] x [Y0..Y1] (the tile-clipped bbox), or +inf when no bound can be given.  A triangle whose bound
// is below the smallest depth already stored in the tile cannot win a single pixel there and need not be walked.
//
// Proof sketch (notation of tile_may_cover above: W_i real affine edge functions on the float constants, B_i the
// magnitude bound, E_i = 2^-16 B_i, every edge value the reference forms for a pixel of the rectangle within
// 38 * 2^-24 B_i of W_i at that pixel).
//  * At a covered pixel the three computed edge values w_i have one sign (Rasterizer.cs:493-494), so the
//    normalised weights w_if = fl(w_i * invArea) have one sign too and sum |w_if| = |sum w_if|.
//  * s = w_0 + w_1 + w_2 (computed values, real sum) differs from S(x,y) = W_0 + W_1 + W_2 by less than sumE / 6.
//    S is affine, its extremes over the rectangle are at the corners; the corner values below carry a few more
//    roundings (again a small multiple of 2^-24 B), so s lies in [smin - sumE, smax + sumE].
//  * If that interval times invArea is entirely NEGATIVE, sum w_if < 0 at every covered pixel, hence every
//    w_if <= 0, and |sum w_if| >= |invArea| * m * (1 - 2^-23) with m = min(|smin - sumE|, |smax + sumE|).
//  * With all d_i >= 0: depth = fl(fl(d0 w0f + d1 w1f) + d2 w2f) is a sum of non-positive terms (no cancellation,
//    relative rounding error < 4 * 2^-24), so depth <= -dmin * |invArea| * m * (1 - 2^-20).
// Anything else -- a d_i that is negative or not finite, an interval that touches zero or has the other sign, NaN
// anywhere -- gives no bound (every comparison below is written so that NaN falls to "no bound").
// ---------------------------------------------------------------------------
__device__ __forceinline__ float tri_depth_upper_bound(const TriRec& r, const TriEq& e, const float w0, const float w1,
                                                       const float w2, const int X0, const int X1, const int Y0, const int Y1) {
    const float inf = __int_as_float(0x7f800000);
    const float dmin = fminf(fminf(r.d0, r.d1), r.d2), dmax = fmaxf(fmaxf(r.d0, r.d1), r.d2);
    if (!(dmin >= 0.0f) || !(dmax < 1e30f) || !(r.d0 == r.d0 && r.d1 == r.d1 && r.d2 == r.d2)) return inf;
    const float fx0 = (float)X0, fx1 = (float)X1, fy0 = (float)Y0, fy1 = (float)Y1;
    // B_i = max |a_i (x - p_ix)| + max |b_i (y - p_iy)| over the rectangle; edge i pivots at s1, s2, s0 (row_base)
    const float B0 = add(mul(fabsf(e.a12), fmaxf(fabsf(sub(fx0, r.s1x)), fabsf(sub(fx1, r.s1x)))),
                         mul(fabsf(e.b12), fmaxf(fabsf(sub(fy0, r.s1y)), fabsf(sub(fy1, r.s1y)))));
    const float B1 = add(mul(fabsf(e.a20), fmaxf(fabsf(sub(fx0, r.s2x)), fabsf(sub(fx1, r.s2x)))),
                         mul(fabsf(e.b20), fmaxf(fabsf(sub(fy0, r.s2y)), fabsf(sub(fy1, r.s2y)))));
    const float B2 = add(mul(fabsf(e.a01), fmaxf(fabsf(sub(fx0, r.s0x)), fabsf(sub(fx1, r.s0x)))),
                         mul(fabsf(e.b01), fmaxf(fabsf(sub(fy0, r.s0y)), fabsf(sub(fy1, r.s0y)))));
    const float sumE = mul(add(add(B0, B1), B2), 1.52587890625e-05f);          // 2^-16
    if (!(sumE < 1e30f)) return inf;
    // S at the four corners: w0 + w1 + w2 are the row-base values at (X0, Y0); S is affine with slopes sa, sb
    const float s00 = add(add(w0, w1), w2);
    const float sa = add(add(e.a12, e.a20), e.a01), sb = add(add(e.b12, e.b20), e.b01);
    const float dx = mul(sa, sub(fx1, fx0)), dy = mul(sb, sub(fy1, fy0));
    const float s10 = add(s00, dx), s01 = add(s00, dy), s11 = add(s10, dy);
    const float smax = fmaxf(fmaxf(s00, s10), fmaxf(s01, s11)), smin = fminf(fminf(s00, s10), fminf(s01, s11));
    const float lo = mul(sub(smin, sumE), e.invArea), hi = mul(add(smax, sumE), e.invArea);     // sum w_if lies between them
    if (!(lo < 0.0f && hi < 0.0f)) return inf;
    const float m = fminf(fabsf(lo), fabsf(hi));                               // |sum w_if| >= m (1 - 2^-23)
    const float b = mul(mul(dmin, m), 0.99999904632568359375f);                // 1 - 2^-20
    return b > 1e-30f ? -b : 0.0f;       // (below that, products may have gone denormal: only "depth <= 0" is claimed)
}

// ---------------------------------------------------------------------------
// Conservative "covers no pixel at all" test for triangles whose clamped pixel bbox is at most
// kSmallPx x kSmallPx: distant meshes produce many triangles that fall between pixel centres, and the
// reference walks them without ever passing its test.  Same bound as tile_may_cover with the bbox as
// the region, but evaluated per pixel: a pixel cannot pass when some edge value is provably < 0
// (W_i < -E_i) AND some edge value is provably > 0 (W_j > +E_j).  Returns false only when that holds
// for every pixel of the bbox; NaN/Inf/huge magnitudes keep the triangle.
// ---------------------------------------------------------------------------
constexpr int kSmallPx = 4;
__device__ __forceinline__ bool small_tri_may_cover(float s0x, float s0y, float s1x, float s1y, float s2x, float s2y,
                                                    int X0, int X1, int Y0, int Y1) {
    const float a[3] = {sub(s1y, s2y), sub(s2y, s0y), sub(s0y, s1y)};     // a12, a20, a01
    const float b[3] = {sub(s2x, s1x), sub(s0x, s2x), sub(s1x, s0x)};     // b12, b20, b01
    const float px[3] = {s1x, s2x, s0x}, py[3] = {s1y, s2y, s0y};
    float ax[3][kSmallPx], by[3][kSmallPx], E[3];
    bool sane = true;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float mx = 0.0f, my = 0.0f;
#pragma unroll
        for (int k = 0; k < kSmallPx; ++k) {
            ax[i][k] = mul(a[i], sub((float)min(X0 + k, X1), px[i]));
            by[i][k] = mul(b[i], sub((float)min(Y0 + k, Y1), py[i]));
            mx = fmaxf(mx, fabsf(ax[i][k]));
            my = fmaxf(my, fabsf(by[i][k]));
        }
        const float B = add(mx, my);
        sane = sane && (B < 1e30f);                                       // false for NaN as well
        E[i] = mul(B, 1.52587890625e-05f);                                // 2^-16
    }
    if (!sane) return true;
    bool may = false;
#pragma unroll
    for (int ky = 0; ky < kSmallPx; ++ky) {
#pragma unroll
        for (int kx = 0; kx < kSmallPx; ++kx) {                           // columns/rows past the bbox repeat its last one
            const float w0 = add(ax[0][kx], by[0][ky]), w1 = add(ax[1][kx], by[1][ky]), w2 = add(ax[2][kx], by[2][ky]);
            const bool geImpossible = (w0 < -E[0]) | (w1 < -E[1]) | (w2 < -E[2]);
            const bool leImpossible = (w0 > E[0]) | (w1 > E[1]) | (w2 > E[2]);
            may = may | !(geImpossible & leImpossible);
        }
    }
    return may;
}

}  // namespace srb
