// srb_physics.cuh -- Physics.Raycast (Physics.cs:19-179) for a batch of rays against one resident mesh.
//
// The reference transforms every vertex of the mesh to world space per call, then tests all triangles
// (Moller-Trumbore, Physics.cs:137-178) and keeps the closest hit.  Here one thread owns one triangle:
// it transforms its three corners (the same function of the same vertex the reference evaluates once per
// vertex, so the values are identical), keeps them in registers, and walks the rays of its chunk.  Hits
// are reduced per ray with a 64-bit key (distance bits, triangle index): the minimum is the closest hit,
// ties go to the lowest triangle index (the sequential order of the reference's loop).  A second kernel
// re-evaluates the winning triangle per ray for the barycentrics, the interpolated normal and the point.
#pragma once
#include "srb_device.cuh"

namespace srb {

struct Mat4 { float m[16]; };
struct RayHit {                 // == SrbRayHit (srb200.h)
    int32_t hit;
    float distance;
    float point[3];
    float normal[3];
    int32_t triangle;
    int32_t _pad[3];
};
static_assert(sizeof(RayHit) == 48, "RayHit layout");

__device__ __forceinline__ float3 cross3(float3 a, float3 b) {              // Vector3.Cross
    return make_float3(sub(mul(a.y, b.z), mul(a.z, b.y)), sub(mul(a.z, b.x), mul(a.x, b.z)), sub(mul(a.x, b.y), mul(a.y, b.x)));
}
__device__ __forceinline__ float dot3v(float3 a, float3 b) { return dot3(a.x, a.y, a.z, b.x, b.y, b.z); }
__device__ __forceinline__ float3 sub3(float3 a, float3 b) { return make_float3(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
__device__ __forceinline__ float3 normalize3v(float3 v) {                  // Vector3.Normalize: v / v.Length()
    const float len = __fsqrt_rn(dot3v(v, v));
    return make_float3(fdiv(v.x, len), fdiv(v.y, len), fdiv(v.z, len));
}

// RayIntersectsTriangle (Physics.cs:137-178) with the edges precomputed.
__device__ __forceinline__ bool ray_tri(const float3 ro, const float3 rd, const float3 v0, const float3 edge1,
                                        const float3 edge2, const int faceMask, float& distance, float& u, float& v) {
    const float Epsilon = 1e-8f;
    const float3 pvec = cross3(rd, edge2);
    const float det = dot3v(edge1, pvec);
    if ((faceMask & 1) && det < Epsilon) return false;                      // :156
    if ((faceMask & 2) && det > -Epsilon) return false;                     // :157
    if (fabsf(det) < Epsilon) return false;                                 // :158
    const float invDet = fdiv(1.0f, det);
    const float3 tvec = sub3(ro, v0);
    u = mul(dot3v(tvec, pvec), invDet);
    if (u < 0.0f || u > 1.0f) return false;
    const float3 qvec = cross3(tvec, edge1);
    v = mul(dot3v(rd, qvec), invDet);
    if (v < 0.0f || add(u, v) > 1.0f) return false;
    distance = mul(dot3v(edge2, qvec), invDet);
    if (distance < 0.0f) return false;
    return true;
}

template <bool FMA>
__device__ __forceinline__ float3 world_pos(const float4* __restrict__ vtx, int vi, const Mat4& M) {
    const float4 a = __ldg(vtx + 3 * vi);                                   // pos.xyz, uv.x
    const float4 p = transform4<FMA>(make_float4(a.x, a.y, a.z, 1.0f), M.m); // Vector3.Transform(position, model) (:46)
    return make_float3(p.x, p.y, p.z);
}

// rays: 6 floats each (origin, direction); normalised directions are written to dirs (:70).
__global__ void ray_prepare_kernel(const float* __restrict__ rays, int nRays, float4* __restrict__ dirs,
                                   unsigned long long* __restrict__ keys) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nRays) return;
    const float3 d = normalize3v(make_float3(rays[6 * r + 3], rays[6 * r + 4], rays[6 * r + 5]));
    dirs[r] = make_float4(d.x, d.y, d.z, 0.0f);
    keys[r] = ~0ull;
}

constexpr int kRayChunk = 64;      // rays per blockIdx.y
template <bool FMA>
__global__ void __launch_bounds__(256)
raycast_kernel(const float4* __restrict__ vtx, const uint16_t* __restrict__ idx, const int nTri, const Mat4 M,
               const float* __restrict__ rays, const float4* __restrict__ dirs, const int nRays, const int faceMask,
               unsigned long long* __restrict__ keys) {
    const unsigned full = 0xffffffffu;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = t < nTri;
    float3 v0 = make_float3(0, 0, 0), e1 = v0, e2 = v0;
    if (live) {
        const int i0 = __ldg(idx + 3 * t), i1 = __ldg(idx + 3 * t + 1), i2 = __ldg(idx + 3 * t + 2);
        v0 = world_pos<FMA>(vtx, i0, M);
        e1 = sub3(world_pos<FMA>(vtx, i1, M), v0);
        e2 = sub3(world_pos<FMA>(vtx, i2, M), v0);
    }
    const int r0 = blockIdx.y * kRayChunk, r1 = min(r0 + kRayChunk, nRays);
    for (int r = r0; r < r1; ++r) {
        const float3 ro = make_float3(__ldg(rays + 6 * r), __ldg(rays + 6 * r + 1), __ldg(rays + 6 * r + 2));
        const float4 d4 = __ldg(dirs + r);
        float dist = 0.0f, u, v;
        // NaN and >= float.MaxValue never pass `distance < local.Distance` (:104) against the MaxValue start
        const bool hit = live && ray_tri(ro, make_float3(d4.x, d4.y, d4.z), v0, e1, e2, faceMask, dist, u, v) &&
                         dist < 3.402823466e+38f;
        if (__ballot_sync(full, hit)) {
            unsigned long long key = hit ? (((unsigned long long)__float_as_uint(dist + 0.0f) << 32) | (unsigned)t) : ~0ull;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(full, key, o);
                key = other < key ? other : key;
            }
            if ((threadIdx.x & 31) == 0) atomicMin(keys + r, key);
        }
    }
}

// Per ray: winning triangle -> barycentrics, interpolated world normal, hit point (Physics.cs:96-107, 123-128).
template <bool FMA>
__global__ void ray_resolve_kernel(const float4* __restrict__ vtx, const uint16_t* __restrict__ idx, const Mat4 M,
                                   const Mat4 NM, const float* __restrict__ rays, const float4* __restrict__ dirs,
                                   const int nRays, const int faceMask, const unsigned long long* __restrict__ keys,
                                   RayHit* __restrict__ out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nRays) return;
    RayHit h;
    h.hit = 0; h.distance = 3.402823466e+38f; h.triangle = -1;              // float.MaxValue, Vector3.Zero (:66-68)
    h.point[0] = h.point[1] = h.point[2] = 0.0f;
    h.normal[0] = h.normal[1] = h.normal[2] = 0.0f;
    h._pad[0] = h._pad[1] = h._pad[2] = 0;
    const unsigned long long key = keys[r];
    if (key != ~0ull) {
        const int t = (int)(unsigned)key;
        const int i[3] = {__ldg(idx + 3 * t), __ldg(idx + 3 * t + 1), __ldg(idx + 3 * t + 2)};
        const float3 v0 = world_pos<FMA>(vtx, i[0], M);
        const float3 e1 = sub3(world_pos<FMA>(vtx, i[1], M), v0), e2 = sub3(world_pos<FMA>(vtx, i[2], M), v0);
        const float3 ro = make_float3(rays[6 * r], rays[6 * r + 1], rays[6 * r + 2]);
        const float4 d4 = dirs[r];
        const float3 rd = make_float3(d4.x, d4.y, d4.z);
        float dist = 0.0f, u = 0.0f, v = 0.0f;
        ray_tri(ro, rd, v0, e1, e2, faceMask, dist, u, v);
        const float b[3] = {sub(sub(1.0f, u), v), u, v};                    // :176
        float3 n[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float4 pb = __ldg(vtx + 3 * i[k] + 1);                    // uv.y, normal.xyz
            const float4 q = transform4<FMA>(make_float4(pb.y, pb.z, pb.w, 0.0f), NM.m);   // (:47)
            n[k] = normalize3v(make_float3(q.x, q.y, q.z));
        }
        const float3 nn = normalize3v(make_float3(add(add(mul(n[0].x, b[0]), mul(n[1].x, b[1])), mul(n[2].x, b[2])),
                                                  add(add(mul(n[0].y, b[0]), mul(n[1].y, b[1])), mul(n[2].y, b[2])),
                                                  add(add(mul(n[0].z, b[0]), mul(n[1].z, b[1])), mul(n[2].z, b[2]))));
        h.hit = 1; h.distance = dist; h.triangle = t;
        h.point[0] = add(ro.x, mul(rd.x, dist)); h.point[1] = add(ro.y, mul(rd.y, dist)); h.point[2] = add(ro.z, mul(rd.z, dist));
        h.normal[0] = nn.x; h.normal[1] = nn.y; h.normal[2] = nn.z;
    }
    out[r] = h;
}

}  // namespace srb
