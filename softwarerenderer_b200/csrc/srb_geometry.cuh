// srb_geometry.cuh -- frustum test, vertex stage, triangle setup + near clip,
// tile binning (count / scan / fill).  Included by srb_kernels.cu only.
#pragma once
#include "srb_device.cuh"

namespace srb {

// Frame-level parameters shared by all kernels.
struct FrameParams {
    int W, H;
    int tilesX, tilesY;
    int bandRow0, bandRow1;       // tile rows rendered by this context [row0,row1)
    int nDraws;
    int totalV, totalT;
    uint32_t pairCap;             // capacity of the bin array
    uint32_t clipCap;             // capacity of the ClipAttr side buffer
    uint32_t bigCap;              // capacity of the large-triangle queue
    float clearColor[4];
    int loadColor, loadDepth;     // 1 = keep existing framebuffer contents (no clear this frame)
    int depthMode;                // visibility path: the frame's common depth test
    int wireframe;                // Rasterizer.RenderDebugMode == Wireframe (Rasterizer.cs:419-425): ordered path only
};

// Device-side counters of one frame.
struct Counters {
    uint32_t nRefs;               // accepted triangles (compact list length)   } one 64-bit word: the setup kernel
    uint32_t nBig;                // entries in the large-triangle queue         } reserves both with ONE atomic per CTA
    uint32_t nPairs;              // total (triangle,tile) pairs (from the scan)
    uint32_t nClipped;            // ClipAttr entries used
    uint32_t nClippedTris;        // triangles that entered the clipper
    uint32_t overflow;            // bit0: bins, bit1: clip side buffer, bit2: large-triangle queue
    uint32_t nVisibleDraws;
    uint32_t _pad0, _pad00;
    // Statistics only, on a cache line of their own: nRefs / nBig above are on the critical path of every
    // setup warp (atomics with a return value), and one more counter on their line cost setup 5 us.
    uint32_t _gap[23];
    uint32_t nNoCover;            // small triangles dropped because they provably cover no pixel
    uint32_t nVLive, nTLive;      // live entries of the compacted vertex / triangle block tables (frustum kernel)
    uint32_t nExtra;              // launch-order entries of split bins beyond the one-per-tile entries (scan kernel)
    uint32_t nSplitEntries;       // bin entries of the bins that were split (scan kernel): sizes the next frame's extra CTAs
    uint32_t _pad1[27];
};
static_assert(sizeof(Counters) == 256, "Counters layout");

// ---------------------------------------------------------------------------
// FrustumCuller.IsSphereInFrustum (FrustumCuller.cs:153-224, Plane ctor :19-28)
// ---------------------------------------------------------------------------
template <bool FMA>
__device__ bool sphere_in_frustum(const float* sphere, const float* M, const float* V, const float* P) {
    float4 wc = transform4<FMA>(make_float4(sphere[0], sphere[1], sphere[2], 1.0f), M);
    float s0 = __fsqrt_rn(add(add(mul(M[0], M[0]), mul(M[1], M[1])), mul(M[2], M[2])));
    float s1 = __fsqrt_rn(add(add(mul(M[4], M[4]), mul(M[5], M[5])), mul(M[6], M[6])));
    float s2 = __fsqrt_rn(add(add(mul(M[8], M[8]), mul(M[9], M[9])), mul(M[10], M[10])));
    float r = mul(sphere[3], net_max(net_max(s0, s1), s2));
    float VP[16];                                    // Matrix4x4.Multiply(view, projection)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float4 row = transform4<FMA>(make_float4(V[4 * i], V[4 * i + 1], V[4 * i + 2], V[4 * i + 3]), P);
        VP[4 * i] = row.x; VP[4 * i + 1] = row.y; VP[4 * i + 2] = row.z; VP[4 * i + 3] = row.w;
    }
    bool ok = true;
    // Left, Right, Top, Bottom, Near, Far (:211-216): column k with sign +/-.
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            float sg = s == 0 ? 1.0f : -1.0f;
            float x = add(VP[3], mul(sg, VP[k])), y = add(VP[7], mul(sg, VP[4 + k]));
            float z = add(VP[11], mul(sg, VP[8 + k])), d = add(VP[15], mul(sg, VP[12 + k]));
            float mag = __fsqrt_rn(add(add(mul(x, x), mul(y, y)), mul(z, z)));     // NormalizePlane
            float nx = fdiv(x, mag), ny = fdiv(y, mag), nz = fdiv(z, mag), dist = fdiv(d, mag);
            float len = __fsqrt_rn(dot3(nx, ny, nz, nx, ny, nz));                  // Plane ctor :26
            nx = fdiv(nx, len); ny = fdiv(ny, len); nz = fdiv(nz, len);
            float sd = add(dot3(nx, ny, nz, wc.x, wc.y, wc.z), dist);
            ok = ok && (sd > -r);
        }
    }
    return ok;
}

// Sort-first bands (multi-GPU): a context that renders only tile rows [bandRow0, bandRow1) can skip a draw whose
// bounding sphere lies entirely above or below its band -- every triangle of it would fail the band clamp in
// setup anyway (coverage is only ever evaluated inside a triangle's pixel bbox), so the image is unchanged and
// the vertex / setup work of a rank shrinks with its band.  Same sphere-vs-plane form as the culler, the
// +-1 of the top/bottom planes replaced by the band's NDC bounds.  Deliberately loose: one pixel row of slack
// on either side, 1 % on the radius and a relative epsilon on the distance, because a wrongly kept draw costs
// nothing but time while a wrongly dropped one would be a wrong image.
__device__ bool sphere_misses_band(const float* sphere, const float* M, const float* V, const float* P,
                                   int H, int bandRow0, int bandRow1, int tilesY) {
    if (bandRow0 <= 0 && bandRow1 >= tilesY) return false;
    const float4 wc = transform4<false>(make_float4(sphere[0], sphere[1], sphere[2], 1.0f), M);
    const float s0 = sqrtf(M[0] * M[0] + M[1] * M[1] + M[2] * M[2]);
    const float s1 = sqrtf(M[4] * M[4] + M[5] * M[5] + M[6] * M[6]);
    const float s2 = sqrtf(M[8] * M[8] + M[9] * M[9] + M[10] * M[10]);
    const float r = sphere[3] * fmaxf(fmaxf(s0, s1), s2);
    float VP[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float4 row = transform4<false>(make_float4(V[4 * i], V[4 * i + 1], V[4 * i + 2], V[4 * i + 3]), P);
        VP[4 * i] = row.x; VP[4 * i + 1] = row.y; VP[4 * i + 2] = row.z; VP[4 * i + 3] = row.w;
    }
    {   // The band planes below say where a point lands on screen only for clip w > 0 (the reference's clip test is
        // z >= NearClip*w, which a non-standard projection can pass with w < 0, and NDC then flips sides): cull
        // only when the whole sphere has w > 0.
        const float a = VP[3], b = VP[7], c = VP[11], d = VP[15];
        const float mag = sqrtf(a * a + b * b + c * c);
        const float w = a * wc.x + b * wc.y + c * wc.z + d;
        const float slack = (1.01f * r + 1e-3f * (fabsf(wc.x) + fabsf(wc.y) + fabsf(wc.z)) + 1e-4f) * mag + 1e-3f * fabsf(d);
        if (!(w > slack)) return false;                        // also for NaN
    }
    bool miss = false;
#pragma unroll
    for (int side = 0; side < 2; ++side) {
        if (side == 0 && bandRow0 <= 0) continue;              // nothing above the first row
        if (side == 1 && bandRow1 >= tilesY) continue;
        // pixel row y  <->  ndc_y = 1 - 2 y / H  (setup_tri); inside the band: ndc_y <= nyTop and ndc_y >= nyBot
        const float yPix = side == 0 ? (float)(bandRow0 * kTile - 1) : (float)(bandRow1 * kTile + 1);
        const float ny = 1.0f - 2.0f * yPix / (float)H;
        const float sg = side == 0 ? 1.0f : -1.0f;             // inside: sg * (ny * w - y) >= 0
        const float a = sg * (ny * VP[3] - VP[1]), b = sg * (ny * VP[7] - VP[5]);
        const float c = sg * (ny * VP[11] - VP[9]), d = sg * (ny * VP[15] - VP[13]);
        const float mag = sqrtf(a * a + b * b + c * c);
        if (!(mag > 0.0f) || !isfinite(mag)) continue;
        const float sd = (a * wc.x + b * wc.y + c * wc.z + d) / mag;
        const float slack = 1.01f * r + 1e-3f * (fabsf(d) / mag + fabsf(wc.x) + fabsf(wc.y) + fabsf(wc.z)) + 1e-4f;
        if (sd < -slack) miss = true;                          // false for NaN
    }
    return miss;
}

// Block culling.  The frustum kernel leaves four object-space half-spaces per draw (screen left / right, band top
// / bottom, each pushed out by one pixel); a 128-triangle block whose object-space bounding
// box (computed at mesh upload) lies entirely outside one of them is dropped from a band context's block table.  Every triangle of such a block has an empty
// clamped pixel bbox (Rasterizer.cs:432-442) or misses the band, i.e. the reference walks no pixel of it, so the
// image is unchanged.  The comparison is deliberately loose (1 % of the box's extent along the plane normal plus a relative epsilon): rounding
// in the vertex transform is five orders of magnitude below that.
__device__ void draw_planes(DrawRec& d, int W, int H, int bandRow0, int bandRow1, int tilesY) {
    float MVP[16], VP[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float4 row = transform4<false>(make_float4(d.view[4 * i], d.view[4 * i + 1], d.view[4 * i + 2], d.view[4 * i + 3]), d.proj);
        VP[4 * i] = row.x; VP[4 * i + 1] = row.y; VP[4 * i + 2] = row.z; VP[4 * i + 3] = row.w;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float4 row = transform4<false>(make_float4(d.model[4 * i], d.model[4 * i + 1], d.model[4 * i + 2], d.model[4 * i + 3]), VP);
        MVP[4 * i] = row.x; MVP[4 * i + 1] = row.y; MVP[4 * i + 2] = row.z; MVP[4 * i + 3] = row.w;
    }
    const float yTop = bandRow0 > 0 ? (float)(bandRow0 * kTile - 1) : -1.0f;
    const float yBot = bandRow1 < tilesY ? (float)(bandRow1 * kTile + 1) : (float)(H + 1);
    // pixel x <-> ndc_x = 2 x / W - 1, pixel y <-> ndc_y = 1 - 2 y / H (setup_tri)
    const float bound[4] = {2.0f * -1.0f / (float)W - 1.0f, 2.0f * (float)(W + 1) / (float)W - 1.0f,
                            1.0f - 2.0f * yTop / (float)H, 1.0f - 2.0f * yBot / (float)H};
    const int col[4] = {0, 0, 1, 1};
    const float sg[4] = {1.0f, -1.0f, -1.0f, 1.0f};      // inside: clip_x >= b w | clip_x <= b w | clip_y <= b w | clip_y >= b w
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float a = sg[k] * (MVP[col[k]] - bound[k] * MVP[3]), b = sg[k] * (MVP[4 + col[k]] - bound[k] * MVP[7]);
        const float c = sg[k] * (MVP[8 + col[k]] - bound[k] * MVP[11]), e = sg[k] * (MVP[12 + col[k]] - bound[k] * MVP[15]);
        const float mag = sqrtf(a * a + b * b + c * c);
        if (mag > 0.0f && isfinite(mag) && isfinite(e)) d.planes[k] = make_float4(a / mag, b / mag, c / mag, e / mag);
        else d.planes[k] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);          // degenerate: never culls
    }
    {   // clip w > 0 as an object-space half-space (column 4 of MVP)
        const float a = MVP[3], b = MVP[7], c = MVP[11], e = MVP[15];
        const float mag = sqrtf(a * a + b * b + c * c);
        if (mag > 0.0f && isfinite(mag) && isfinite(e)) d.planes[4] = make_float4(a / mag, b / mag, c / mag, e / mag);
        else d.planes[4] = make_float4(0.0f, 0.0f, 0.0f, e > 0.0f ? 1.0e30f : -1.0f);   // w is the same everywhere (affine projection)
    }
}
__device__ __forceinline__ bool block_outside(const DrawRec& d, int block) {
    const float4 c = __ldg(d.tblockBoxes + 2 * block), h = __ldg(d.tblockBoxes + 2 * block + 1);
    // signed distance of the box's centre to a (normalised) plane and the box's extent along the plane's normal
    auto dist = [&](const float4 p, float& ext) {
        ext = fabsf(p.x) * h.x + fabsf(p.y) * h.y + fabsf(p.z) * h.z;
        return p.x * c.x + p.y * c.y + p.z * c.z + p.w;
    };
    auto slack = [&](const float4 p, float ext) {
        return 0.01f * ext + 1e-3f * (fabsf(p.w) + fabsf(c.x) + fabsf(c.y) + fabsf(c.z)) + 1e-4f;
    };
    {   // every vertex of the block must have clip w > 0 for the screen planes to mean anything (see DrawRec::planes)
        float ext;
        const float sd = dist(d.planes[4], ext);
        if (!(sd - ext > slack(d.planes[4], ext))) return false;        // also for NaN / the never-culled box
    }
    bool out = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float ext;
        const float sd = dist(d.planes[k], ext);
        out = out || (sd + ext < -slack(d.planes[k], ext));             // false for NaN
    }
    return out;
}

// (Measured, no effect on the 12 us stage: eight lanes per draw with one frustum plane each.)
// Also compacts the block tables: every draw that survives appends its runs of kGeomBlock vertices / triangles to the
// live tables, so that the vertex and setup kernels spend no CTA on culled draws (city4k: 63 % of the blocks).
__global__ void frustum_draws_kernel(DrawRec* draws, int nDraws, int fma, Counters* cnt, int W, int H, int bandRow0,
                                     int bandRow1, int tilesY, int2* __restrict__ vLive, int2* __restrict__ tLive) {
    // One warp per draw: every lane repeats the (broadcast-load) sphere tests, lane 0 owns the writes and the
    // reservations, and the 32 lanes write the draw's table entries together (one thread per draw writing them in
    // series cost 3 us on a single 50k-triangle draw).
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= nDraws) return;                                   // uniform over the warp
    DrawRec& d = draws[i];
    int vis = 1;
    if (d.flags & SRB_DRAW_FRUSTUM_CULL) {
        vis = fma ? sphere_in_frustum<true>(d.sphere, d.model, d.view, d.proj)
                  : sphere_in_frustum<false>(d.sphere, d.model, d.view, d.proj);
        if (vis && lane == 0) atomicAdd(&cnt->nVisibleDraws, 1u);   // the reference's count, whatever the band
        if (vis && sphere_misses_band(d.sphere, d.model, d.view, d.proj, H, bandRow0, bandRow1, tilesY)) vis = 0;
    } else if (lane == 0) {
        atomicAdd(&cnt->nVisibleDraws, 1u);
    }
    // (Measured on one GPU, full frame: block culling gains nothing -- few blocks of a draw that passed the sphere
    // test are off screen -- and the planes cost the frustum kernel 2.8 us; it is a band (multi-GPU) feature.)
    const int nvb = (d.n_vertices + 127) / 128, ntb = (d.n_triangles + 127) / 128;
    const bool banded = bandRow0 > 0 || bandRow1 < tilesY;
    uint32_t vo = 0, to = 0;
    __syncwarp();                                              // every lane has read d before lane 0 writes to it
    if (lane == 0) {
        d.visible = vis;
        if (vis && banded) draw_planes(d, W, H, bandRow0, bandRow1, tilesY);
        if (vis && !banded) { vo = atomicAdd(&cnt->nVLive, (uint32_t)nvb); to = atomicAdd(&cnt->nTLive, (uint32_t)ntb); }
    }
    if (!vis || banded) return;                                // a band context's tables are built by band_blocks_kernel
    vo = __shfl_sync(0xffffffffu, vo, 0);
    to = __shfl_sync(0xffffffffu, to, 0);
    for (int k = lane; k < nvb; k += 32) vLive[vo + k] = make_int2(i, k * 128);
    for (int k = lane; k < ntb; k += 32) tLive[to + k] = make_int2(i, k * 128);
}

// Band contexts (multi-GPU): the live tables hold only what the band needs.  One CTA per visible draw tests every run
// of kGeomBlock triangles against the band's half-spaces (block_outside) and appends the survivors to the triangle
// table; every survivor marks the runs of kGeomBlock VERTICES its triangles index (lowest..highest index of the run,
// recorded at mesh upload), and only marked vertex runs enter the vertex table.  Before, the vertex stage of a band
// rank transformed every vertex of every visible draw (stress8k: 84 us on each of 8 ranks) and three quarters of the
// setup kernel's CTAs loaded a table entry, a draw record and a sphere only to find their block off band.
constexpr int kMaxVBlocks = 65536 / 128;                       // srb_mesh_create: at most 65536 vertices (ushort indices)
__global__ void __launch_bounds__(128)
band_blocks_kernel(const DrawRec* __restrict__ draws, Counters* cnt, int2* __restrict__ vLive, int2* __restrict__ tLive) {
    __shared__ uint32_t marks[kMaxVBlocks / 32];
    const int di = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const DrawRec& d = draws[di];
    if (!d.visible) return;                                    // uniform over the CTA
    if (tid < kMaxVBlocks / 32) marks[tid] = 0u;
    __syncthreads();
    const int ntb = (d.n_triangles + 127) / 128, nvb = (d.n_vertices + 127) / 128;
    for (int base = 0; base < ntb; base += 128) {
        const int k = base + tid;
        const bool keep = k < ntb && !block_outside(d, k);
        if (keep) {
            const int2 vr = __ldg(d.tblockVRange + k);
            for (int vb = vr.x >> 7; vb <= (vr.y >> 7); ++vb) atomicOr(&marks[vb >> 5], 1u << (vb & 31));
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (m) {                                               // uniform over the warp
            uint32_t wbase = 0;
            if (lane == 0) wbase = atomicAdd(&cnt->nTLive, (uint32_t)__popc(m));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (keep) tLive[wbase + __popc(m & ((1u << lane) - 1u))] = make_int2(di, k * 128);
        }
    }
    __syncthreads();
    if (tid < 32) {                                            // the first warp writes the marked vertex runs
        uint32_t w = tid < kMaxVBlocks / 32 ? marks[tid] : 0u;
        if (tid * 32 >= nvb) w = 0u;
        else if (tid * 32 + 32 > nvb) w &= (1u << (nvb - tid * 32)) - 1u;
        uint32_t incl = (uint32_t)__popc(w);
        const uint32_t mine = incl;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
        uint32_t vbase = 0;
        if (lane == 31 && tot) vbase = atomicAdd(&cnt->nVLive, tot);
        vbase = __shfl_sync(0xffffffffu, vbase, 31) + incl - mine;
        while (w) {
            const int b = __ffs(w) - 1;
            w &= w - 1u;
            vLive[vbase++] = make_int2(di, (tid * 32 + b) * 128);
        }
    }
}

__global__ void frustum_batch_kernel(const float* spheres, const float* models, const float* viewproj,
                                     int n, int fma, int32_t* out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* V = viewproj;
    const float* P = viewproj + 16;
    out[i] = fma ? sphere_in_frustum<true>(spheres + 4 * i, models + 16 * i, V, P)
                 : sphere_in_frustum<false>(spheres + 4 * i, models + 16 * i, V, P);
}

// Geometry kernels run one CTA per block-table entry: a run of <= kGeomBlock
// vertices (or triangles) of ONE draw, built on the host when the draw list is
// uploaded.  The draw lookup is a single load and a frustum-culled draw retires
// whole CTAs at once.  int2 = {draw index, first element within the draw}.
constexpr int kGeomBlock = 128;

// ---------------------------------------------------------------------------
// Vertex stage: Renderer.VertexShader (Renderer.cs:830-846) once per unique
// vertex of every visible draw.  48-B vertex read as three float4.
// ---------------------------------------------------------------------------
template <bool FMA>
__global__ void __launch_bounds__(kGeomBlock)
vertex_kernel(const DrawRec* __restrict__ draws, const int2* __restrict__ vblocks, const Counters* __restrict__ cnt,
              float4* __restrict__ clipOut, float4* __restrict__ wnOut) {
    // vblocks is the live table the frustum kernel of this frame compacted (visible draws only).  The grid is sized
    // from an earlier frame's live count; the stride loop makes any grid size correct.  The first entry is loaded
    // beside the live count, not after it (the table has a slot for every block of the scene, live or not): one L2
    // round trip less per CTA.
    int2 blk = vblocks[blockIdx.x];
    const uint32_t nLive = cnt->nVLive;
  for (uint32_t bi = blockIdx.x; bi < nLive; bi += gridDim.x, blk = vblocks[min(bi, nLive - 1)]) {
    const DrawRec& d = draws[blk.x];
    const int v = blk.y + threadIdx.x;
    if (v >= d.n_vertices) continue;
    const int g = d.vbase + v;
    float4 a = __ldg(d.vtx + 3 * v);          // pos.xyz, uv.x
    float4 b = __ldg(d.vtx + 3 * v + 1);      // uv.y, normal.xyz
    float4 world = transform4<FMA>(make_float4(a.x, a.y, a.z, 1.0f), d.model);
    float4 view = transform4<FMA>(world, d.view);
    float4 clip = transform4<FMA>(view, d.proj);
    float nx = xform_normal_col<FMA>(b.y, b.z, b.w, d.model, 0);
    float ny = xform_normal_col<FMA>(b.y, b.z, b.w, d.model, 1);
    float nz = xform_normal_col<FMA>(b.y, b.z, b.w, d.model, 2);
    float len = __fsqrt_rn(dot3(nx, ny, nz, nx, ny, nz));     // Vector3.Normalize = v / Length
    clipOut[g] = clip;
    wnOut[g] = make_float4(fdiv(nx, len), fdiv(ny, len), fdiv(nz, len), 0.0f);
  }
}

// ---------------------------------------------------------------------------
// Triangle setup
// ---------------------------------------------------------------------------
// DrawTriangle + head of RasterizeTriangle (Rasterizer.cs:365-398, 411-442) on
// corners already in output order {v2,v1,v0}.  Returns false on rejection.
__device__ __forceinline__ bool setup_tri(const float4 c[3], int W, int H, int cull, TriRec& r,
                                          int& minX, int& maxX, int& minY, int& maxY) {
    float sx[3], sy[3], dz[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float invW = fdiv(1.0f, c[i].w);
        float nx = mul(c[i].x, invW), ny = mul(c[i].y, invW), nz = mul(c[i].z, invW);
        if (isnan(nx) || isnan(ny) || isnan(nz) || isinf(nx) || isinf(ny) || isinf(nz)) return false;
        sx[i] = mul(add(mul(nx, 0.5f), 0.5f), (float)W);
        sy[i] = mul(sub(1.0f, add(mul(ny, 0.5f), 0.5f)), (float)H);
        dz[i] = mul(add(nz, 1.0f), 0.5f);
    }
    if (c[0].w == 0.0f || c[1].w == 0.0f || c[2].w == 0.0f) return false;
    float area = edge_fn(sx[0], sy[0], sx[1], sy[1], sx[2], sy[2]);
    if (area == 0.0f) return false;
    bool front = area < 0.0f;
    if ((cull == SRB_CULL_BACK && !front) || (cull == SRB_CULL_FRONT && front)) return false;
    r.s0x = sx[0]; r.s0y = sy[0]; r.s1x = sx[1]; r.s1y = sy[1]; r.s2x = sx[2]; r.s2y = sy[2];
    r.d0 = dz[0]; r.d1 = dz[1]; r.d2 = dz[2];
    r.cz0 = c[0].z; r.cz1 = c[1].z; r.cz2 = c[2].z;
    r.cw0 = c[0].w; r.cw1 = c[1].w; r.cw2 = c[2].w;
    r.invArea = fdiv(1.0f, area);                                 // :427
    tri_bbox(r, W, H, minX, maxX, minY, maxY);
    if (minX > maxX || minY > maxY) return false;                 // :442
    r.bx = (uint32_t)minX | ((uint32_t)maxX << 16);
    r.by = (uint32_t)minY | ((uint32_t)maxY << 16);
    return true;
}

// Full varyings of one corner, needed only inside the near clipper.
struct ClipVtx {
    float4 clip;
    float uv[2];
    float col[4];
    float wn[3];
    float wpos[3];       // EXTENSION (point lights): world position
};
__device__ __forceinline__ ClipVtx clipvtx_lerp(const ClipVtx& a, const ClipVtx& b, float t, bool fma) {
    ClipVtx o;                                       // Shaders.Lerp (Shaders.cs:49-95), interpolate=true
    o.clip = make_float4(lerp1(a.clip.x, b.clip.x, t, fma), lerp1(a.clip.y, b.clip.y, t, fma),
                         lerp1(a.clip.z, b.clip.z, t, fma), lerp1(a.clip.w, b.clip.w, t, fma));
#pragma unroll
    for (int k = 0; k < 2; ++k) o.uv[k] = lerp1(a.uv[k], b.uv[k], t, fma);
#pragma unroll
    for (int k = 0; k < 4; ++k) o.col[k] = lerp1(a.col[k], b.col[k], t, fma);
#pragma unroll
    for (int k = 0; k < 3; ++k) o.wn[k] = lerp1(a.wn[k], b.wn[k], t, fma);
#pragma unroll
    for (int k = 0; k < 3; ++k) o.wpos[k] = lerp1(a.wpos[k], b.wpos[k], t, fma);
    return o;
}

// Result of per-thread setup handed to the binning helpers.
// Screen-space geometry of an accepted triangle, kept for the conservative tile test applied
// when a triangle that spans many tiles is binned (tile_may_cover).
struct BinGeom {
    float s0x, s0y, s1x, s1y, s2x, s2y;
    uint32_t bx, by;       // clamped pixel bbox, packed as in TriRec
};
struct SetupOut {
    int n;                 // accepted triangles of this thread (0..2)
    int nNoCover;          // triangles dropped by small_tri_may_cover (statistics only)
    uint32_t seq[2];
    uint32_t tx[2], ty[2]; // packed tile ranges
    uint32_t bigMask;      // bit k: triangle k spans more than kSmallTiles tiles (goes to the large-triangle queue)
};
// Triangles that span more than kSmallTiles tiles are queued in shared memory and binned by the
// whole CTA (cta_bin_big).
constexpr int kSmallTiles = 4;
struct __align__(16) BigEntry {
    BinGeom g;
    uint32_t seq, tx, ty, _pad;
};
static_assert(sizeof(BigEntry) == 48, "BigEntry is read as three float4");
// Common tail of the accepted-triangle paths: band clamp, no-coverage test, record store.
// (Measured and rejected, round 2: rasterizing unclipped triangles of <= 4x4 or <= 8x8 px right here, one global
// 64-bit atomicMax per covered pixel into a per-pixel key buffer, no record / list / bin entry, winners rebuilt from
// the mesh in the resolve step.  Bit-exact, but slower everywhere: city4k 5,740 -> 5,260 / 4,830 frames/s, stress8k
// 499 -> 291 / 430 -- the setup warp serialises on its few small-triangle lanes, the global atomics cost more than
// the 80-byte records they replace, and rebuilding a winner is a chain of five dependent gathers per pixel.
// profiles/r02a_*_directpath_experiment.json.)
__device__ __forceinline__ void accept_tri(const TriRec& r, uint32_t seq, int minX, int maxX, int minY, int maxY,
                                           const FrameParams& fp, const RecStore& recs, SetupOut& out) {
    const int ty0 = max(minY / kTile, fp.bandRow0), ty1 = min(maxY / kTile, fp.bandRow1 - 1);
    if (ty0 > ty1) return;
#ifndef SRB_EXP_NO_SMALLCULL
    // Triangles of a few pixels that provably pass the reference's test nowhere are dropped here: they would
    // only be walked for nothing (about 30 % of the accepted triangles of the city scene).
    if (!fp.wireframe && maxX - minX < kSmallPx && maxY - minY < kSmallPx &&
        !small_tri_may_cover(r.s0x, r.s0y, r.s1x, r.s1y, r.s2x, r.s2y, minX, maxX, minY, maxY)) {
        ++out.nNoCover;
        return;
    }
#endif
    // (Measured and rejected: applying tile_may_cover to the 2..4 tiles of small triangles as well and handing
    // the fill pass a mask -- 9 % fewer pairs, but raster only 113.3 -> 112.7 us.)
    store_rec(recs, seq, r);
    const int k = out.n++;
    out.seq[k] = seq;
    out.tx[k] = (uint32_t)(minX / kTile) | ((uint32_t)(maxX / kTile) << 16);
    out.ty[k] = (uint32_t)ty0 | ((uint32_t)ty1 << 16);
    if ((maxX / kTile - minX / kTile + 1) * (ty1 - ty0 + 1) > kSmallTiles) out.bigMask |= 1u << k;
}

// Near-clipped triangle (Rasterizer.cs:95-160, 217-224).  Rare and divergent.
__device__ __noinline__ void setup_clipped(const DrawRec& d, int drawIndex, int vbase, const int i3[3],
                                           const float4* __restrict__ clipIn, const float4* __restrict__ wnIn,
                                           const FrameParams& fp, const SrbUniforms& U, uint32_t gtri,
                                           const RecStore& recs, ClipAttr* clipAttrs, Counters* cnt, SetupOut& out) {
    const bool fma = U.numerics_fma != 0;
    ClipVtx v[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        int vi = i3[k];
        float4 a = __ldg(d.vtx + 3 * vi), b = __ldg(d.vtx + 3 * vi + 1), c = __ldg(d.vtx + 3 * vi + 2);
        float4 wn = wnIn[vbase + vi];
        v[k].clip = clipIn[vbase + vi];
        v[k].uv[0] = a.w; v[k].uv[1] = b.x;
        v[k].col[0] = c.x; v[k].col[1] = c.y; v[k].col[2] = c.z; v[k].col[3] = c.w;
        v[k].wn[0] = wn.x; v[k].wn[1] = wn.y; v[k].wn[2] = wn.z;
        const float4 wp = fma ? transform4<true>(make_float4(a.x, a.y, a.z, 1.0f), d.model)
                              : transform4<false>(make_float4(a.x, a.y, a.z, 1.0f), d.model);      // the vertex shader's `world`
        v[k].wpos[0] = wp.x; v[k].wpos[1] = wp.y; v[k].wpos[2] = wp.z;
    }
    atomicAdd(&cnt->nClippedTris, 1u);
    ClipVtx poly[4];
    int n = 0;
    const float nc = U.near_clip;
    for (int i = 0; i < 3; ++i) {
        const ClipVtx& cur = v[i];
        const ClipVtx& nxt = v[(i + 1) % 3];
        bool curIn = cur.clip.z >= mul(nc, cur.clip.w);
        bool nxtIn = nxt.clip.z >= mul(nc, nxt.clip.w);
        if (curIn && n < 4) poly[n++] = cur;
        if (curIn != nxtIn) {
            float z0 = cur.clip.z, w0 = cur.clip.w, z1 = nxt.clip.z, w1 = nxt.clip.w;
            float denom = sub(sub(z1, z0), mul(nc, sub(w1, w0)));
            float t;
            if (fabsf(denom) < kEps) {
                t = 0.5f;
            } else {
                t = fdiv(sub(z0, mul(nc, w0)), sub(mul(nc, sub(w1, w0)), sub(z1, z0)));
                t = net_clamp(t, 0.0f, 1.0f);
            }
            if (n < 4) poly[n++] = clipvtx_lerp(cur, nxt, t, fma);
        }
    }
    if (n < 3) return;
    for (int f = 0; f + 2 < n; ++f) {                 // fan (0, f+1, f+2)
        const ClipVtx* tri[3] = {&poly[0], &poly[f + 1], &poly[f + 2]};
        const ClipVtx* o[3] = {tri[2], tri[1], tri[0]};            // :367 order {v2,v1,v0}
        float4 c[3] = {o[0]->clip, o[1]->clip, o[2]->clip};
        TriRec r;
        int minX, maxX, minY, maxY;
        if (!setup_tri(c, fp.W, fp.H, d.cull, r, minX, maxX, minY, maxY)) continue;
        int ty0 = max(minY / kTile, fp.bandRow0), ty1 = min(maxY / kTile, fp.bandRow1 - 1);
        if (ty0 > ty1) continue;
        uint32_t slot = atomicAdd(&cnt->nClipped, 1u);
        if (slot >= fp.clipCap) { atomicOr(&cnt->overflow, 2u); continue; }
        ClipAttr ca;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            ca.uv[k][0] = o[k]->uv[0]; ca.uv[k][1] = o[k]->uv[1];
#pragma unroll
            for (int j = 0; j < 4; ++j) ca.col[k][j] = o[k]->col[j];
#pragma unroll
            for (int j = 0; j < 3; ++j) ca.wn[k][j] = o[k]->wn[j];
#pragma unroll
            for (int j = 0; j < 3; ++j) ca.wpos[k][j] = o[k]->wpos[j];
        }
        ca.draw = drawIndex;
        clipAttrs[slot] = ca;
        r.meta = 0x80000000u | slot;
        accept_tri(r, 2u * gtri + (uint32_t)f, minX, maxX, minY, maxY, fp, recs, out);
    }
}

}  // namespace srb
