// oracle.cpp -- CPU oracle for the SoftwareRenderer rasterization path.
//
// TEST INFRASTRUCTURE ONLY (see oracle.h).  PARITY UNPINNED by upstream: the
// reference has no tests or golden data and its C# cannot run in this image.
//
// Every function cites the reference lines whose semantics it restates.
// Build: g++ -O2 -ffp-contract=off -fno-fast-math -mfma (one IEEE rounding per
// scalar op, like RyuJIT scalar code; fused only where fmaf() is written).
#include "oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <climits>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace {

constexpr int   kTile = 16;        // Rasterizer.cs:53
constexpr float kEps  = 1e-6f;     // Rasterizer.cs:52

// ---- System.Numerics restatement (net9.0 BCL; see DESIGN.md "numerics") ----
struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };

// .NET 9 float->int: saturating, NaN -> 0.
inline int sat_int(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT_MAX;
    if (f <= -2147483648.0f) return INT_MIN;
    return (int)f;
}

// MathF.Max / MathF.Min: NaN-propagating.
inline float net_max(float a, float b) { if (a != a) return a; if (b != b) return b; return a > b ? a : b; }
inline float net_min(float a, float b) { if (a != a) return a; if (b != b) return b; return a < b ? a : b; }
// Math.Clamp(float): NaN passes through.
inline float net_clamp(float v, float lo, float hi) { if (v < lo) return lo; if (v > hi) return hi; return v; }

// Vector4.Transform(v, M): row-vector times matrix, accumulated row by row.
// fma=1: MultiplyAddEstimate chain (net9 on FMA hardware); fma=0: mul then add.
inline V4 transform4(V4 v, const float* m, bool fma) {
    float r[4];
    for (int c = 0; c < 4; ++c) {
        float acc = m[0 + c] * v.x;
        if (fma) {
            acc = fmaf(m[4 + c], v.y, acc);
            acc = fmaf(m[8 + c], v.z, acc);
            acc = fmaf(m[12 + c], v.w, acc);
        } else {
            acc = acc + m[4 + c] * v.y;
            acc = acc + m[8 + c] * v.z;
            acc = acc + m[12 + c] * v.w;
        }
        r[c] = acc;
    }
    return {r[0], r[1], r[2], r[3]};
}

// Vector3.TransformNormal(n, M): upper 3x3 only.
inline V3 transform_normal(V3 n, const float* m, bool fma) {
    float r[3];
    for (int c = 0; c < 3; ++c) {
        float acc = m[0 + c] * n.x;
        if (fma) {
            acc = fmaf(m[4 + c], n.y, acc);
            acc = fmaf(m[8 + c], n.z, acc);
        } else {
            acc = acc + m[4 + c] * n.y;
            acc = acc + m[8 + c] * n.z;
        }
        r[c] = acc;
    }
    return {r[0], r[1], r[2]};
}

inline float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }

// Vector3.Normalize: value / value.Length()
inline V3 normalize3(V3 v) {
    float len = sqrtf(dot3(v, v));
    return {v.x / len, v.y / len, v.z / len};
}

// Vector{2,3,4}.Lerp component: a*(1-t) + b*t, fused as fma(a, 1-t, b*t) under the switch.
inline float lerp1(float a, float b, float t, bool fma) {
    float omt = 1.0f - t;
    float bt = b * t;
    return fma ? fmaf(a, omt, bt) : a * omt + bt;
}

// ---- Shaders.VertexOutput restricted to what the stock shaders touch ----
struct VOut {
    V4 clip;      // ClipPosition
    V4 color;     // Color
    V2 uv;        // TexCoord
    V3 normal;    // Normal (object space, passed through)
    V2 screen;    // ScreenCoords
    V3 wn;        // Data["WorldNormal"]
    V3 wpos;      // EXTENSION (point lights only): world position, the vertex shader's `world` (Renderer.cs:832)
};

// Renderer.VertexShader (Renderer.cs:830-846).
inline VOut vertex_shader(const float* vin, const float* M, const float* V, const float* P, bool fma) {
    VOut o;
    V4 world = transform4({vin[0], vin[1], vin[2], 1.0f}, M, fma);
    V4 view  = transform4(world, V, fma);
    o.clip   = transform4(view, P, fma);
    V3 n     = {vin[5], vin[6], vin[7]};
    o.wn     = normalize3(transform_normal(n, M, fma));
    o.uv     = {vin[3], vin[4]};
    o.color  = {vin[8], vin[9], vin[10], vin[11]};
    o.normal = n;
    o.screen = {0.0f, 0.0f};
    o.wpos   = {world.x, world.y, world.z};
    return o;
}

// Shaders.Lerp (Shaders.cs:49-95) with interpolate=true; WorldNormal is
// Vector3.Lerp'ed and NOT renormalised; ScreenCoords stays default.
inline VOut vout_lerp(const VOut& a, const VOut& b, float t, bool fma) {
    VOut o;
    o.clip   = {lerp1(a.clip.x, b.clip.x, t, fma), lerp1(a.clip.y, b.clip.y, t, fma),
                lerp1(a.clip.z, b.clip.z, t, fma), lerp1(a.clip.w, b.clip.w, t, fma)};
    o.uv     = {lerp1(a.uv.x, b.uv.x, t, fma), lerp1(a.uv.y, b.uv.y, t, fma)};
    o.color  = {lerp1(a.color.x, b.color.x, t, fma), lerp1(a.color.y, b.color.y, t, fma),
                lerp1(a.color.z, b.color.z, t, fma), lerp1(a.color.w, b.color.w, t, fma)};
    o.normal = {lerp1(a.normal.x, b.normal.x, t, fma), lerp1(a.normal.y, b.normal.y, t, fma),
                lerp1(a.normal.z, b.normal.z, t, fma)};
    o.wn     = {lerp1(a.wn.x, b.wn.x, t, fma), lerp1(a.wn.y, b.wn.y, t, fma),
                lerp1(a.wn.z, b.wn.z, t, fma)};
    o.screen = {0.0f, 0.0f};
    o.wpos   = {lerp1(a.wpos.x, b.wpos.x, t, fma), lerp1(a.wpos.y, b.wpos.y, t, fma),
                lerp1(a.wpos.z, b.wpos.z, t, fma)};     // EXTENSION (point lights): Lerp'd like the other varyings
    return o;
}

// ClipTriangleAgainstNearPlane (Rasterizer.cs:95-160): Sutherland-Hodgman
// against z >= NearClip*w, then a fan.  Returns the number of triangles.
inline int clip_near(const VOut in[3], float nearClip, bool fma, VOut out[2][3]) {
    VOut poly[6];
    int n = 0;
    for (int i = 0; i < 3; ++i) {
        const VOut& cur = in[i];
        const VOut& nxt = in[(i + 1) % 3];
        bool curIn = cur.clip.z >= nearClip * cur.clip.w;
        bool nxtIn = nxt.clip.z >= nearClip * nxt.clip.w;
        if (curIn) poly[n++] = cur;
        if (curIn != nxtIn) {
            float z0 = cur.clip.z, w0 = cur.clip.w, z1 = nxt.clip.z, w1 = nxt.clip.w;
            float denom = (z1 - z0) - nearClip * (w1 - w0);
            float t;
            if (fabsf(denom) < kEps) {
                t = 0.5f;
            } else {
                t = (z0 - nearClip * w0) / (nearClip * (w1 - w0) - (z1 - z0));
                t = net_clamp(t, 0.0f, 1.0f);
            }
            poly[n++] = vout_lerp(cur, nxt, t, fma);
        }
    }
    if (n < 3) return 0;
    int nt = 0;
    for (int i = 1; i < n - 1 && nt < 2; ++i, ++nt) {
        out[nt][0] = poly[0];
        out[nt][1] = poly[i];
        out[nt][2] = poly[i + 1];
    }
    return nt;
}

// Texture.Sample (Texture.cs:42-63): nearest, repeat.
inline V4 tex_nearest(const uint8_t* px, int tw, int th, float uu, float vv) {
    float u = uu - (float)sat_int(uu);
    float v = vv - (float)sat_int(vv);
    u += (u < 0) ? 1.0f : 0.0f;
    v += (v < 0) ? 1.0f : 0.0f;
    int x = sat_int(u * (float)tw) % tw;
    int y = sat_int(v * (float)th) % th;
    if (x < 0) x += tw;
    if (y < 0) y += th;
    const uint8_t* p = px + 4 * ((size_t)y * tw + x);
    const float inv255 = 1.0f / 255.0f;
    return {p[0] * inv255, p[1] * inv255, p[2] * inv255, p[3] * inv255};
}

// EXTENSION (not in the reference; BASELINE.json config 5): bilinear, repeat
// wrap, texel centres at +0.5, fp32 lerp of the four x(1/255) texels.
inline V4 tex_bilinear(const uint8_t* px, int tw, int th, float uu, float vv) {
    float u = uu - floorf(uu);
    float v = vv - floorf(vv);
    float fx = u * (float)tw - 0.5f;
    float fy = v * (float)th - 0.5f;
    float x0f = floorf(fx), y0f = floorf(fy);
    float ax = fx - x0f, ay = fy - y0f;
    int x0 = sat_int(x0f) % tw; if (x0 < 0) x0 += tw;
    int y0 = sat_int(y0f) % th; if (y0 < 0) y0 += th;
    int x1 = (x0 + 1) % tw, y1 = (y0 + 1) % th;
    const float inv255 = 1.0f / 255.0f;
    float out[4];
    for (int c = 0; c < 4; ++c) {
        float t00 = px[4 * ((size_t)y0 * tw + x0) + c] * inv255;
        float t10 = px[4 * ((size_t)y0 * tw + x1) + c] * inv255;
        float t01 = px[4 * ((size_t)y1 * tw + x0) + c] * inv255;
        float t11 = px[4 * ((size_t)y1 * tw + x1) + c] * inv255;
        float top = t00 + (t10 - t00) * ax;
        float bot = t01 + (t11 - t01) * ax;
        out[c] = top + (bot - top) * ay;
    }
    return {out[0], out[1], out[2], out[3]};
}

// Renderer.FragmentShader (Renderer.cs:848-860).
inline V4 fragment_shader(V3 wn, V4 color, V2 uv, float clipZ, const OrcDraw& d, const OrcUniforms& U,
                          V3 wpos = {0.0f, 0.0f, 0.0f}) {
    bool fma = U.numerics_fma != 0;
    V3 negL = {-U.light_dir[0], -U.light_dir[1], -U.light_dir[2]};
    float diffuse = net_max(0.25f, dot3(wn, negL));
    V4 tex = {1.0f, 1.0f, 1.0f, 1.0f};
    if (d.tex_rgba8)
        tex = U.tex_filter == 1 ? tex_bilinear(d.tex_rgba8, d.tex_w, d.tex_h, uv.x, uv.y)
                                : tex_nearest(d.tex_rgba8, d.tex_w, d.tex_h, uv.x, uv.y);
    V4 base = {color.x * tex.x, color.y * tex.y, color.z * tex.z, color.w * tex.w};
    float fog = net_clamp((U.fog_end - clipZ) / (U.fog_end - U.fog_start), 0.0f, 1.0f);
    fog = fog * fog * (3.0f - 2.0f * fog);
    float k = 0.1f + 0.9f * diffuse;
    V4 lit = {base.x * k * U.light_color[0], base.y * k * U.light_color[1],
              base.z * k * U.light_color[2], base.w * k * U.light_color[3]};
    // EXTENSION: extra lights add Color*max(0, N.L)*base (no ambient floor).  Directional:
    // L = -Direction.  Point (Light.cs:9, 13-15): L = normalised (Position - fragment world position), scaled by
    // 1 / (AttenuationConstant + AttenuationLinear*d + AttenuationQuadratic*d*d).
    for (int i = 0; i < U.n_extra_lights && i < 4; ++i) {
        V3 nd = {-U.extra_dir[i][0], -U.extra_dir[i][1], -U.extra_dir[i][2]};
        float att = 1.0f;
        if (U.extra_dir[i][3] != 0.0f) {
            V3 L = {U.extra_dir[i][0] - wpos.x, U.extra_dir[i][1] - wpos.y, U.extra_dir[i][2] - wpos.z};
            float dist = sqrtf(dot3(L, L));
            float inv = 1.0f / dist;
            nd = {L.x * inv, L.y * inv, L.z * inv};
            att = 1.0f / ((U.extra_atten[i][0] + U.extra_atten[i][1] * dist) + U.extra_atten[i][2] * (dist * dist));
        }
        float di = net_max(0.0f, dot3(wn, nd)) * att;
        lit.x = lit.x + base.x * di * U.extra_color[i][0];
        lit.y = lit.y + base.y * di * U.extra_color[i][1];
        lit.z = lit.z + base.z * di * U.extra_color[i][2];
    }
    // Vector4.Lerp(FogColor, lit, fog)
    V4 fin = {lerp1(U.fog_color[0], lit.x, fog, fma), lerp1(U.fog_color[1], lit.y, fog, fma),
              lerp1(U.fog_color[2], lit.z, fog, fma), 0.0f};
    return {fin.x, fin.y, fin.z, base.w};
}

// Rasterizer.Blend (Rasterizer.cs:58-65).
inline V4 blend(V4 s, V4 d, int mode) {
    switch (mode) {
        case 1: {  // Alpha: src*src.W + dst*(1-src.W), all four channels
            float a = s.w, ia = 1.0f - s.w;
            return {s.x * a + d.x * ia, s.y * a + d.y * ia, s.z * a + d.z * ia, s.w * a + d.w * ia};
        }
        case 2: {  // Additive: Vector4.Min(src+dst, One)
            // Vector4.Min is a per-lane (x < y) ? x : y select on the sum.
            float x = s.x + d.x, y = s.y + d.y, z = s.z + d.z, w = s.w + d.w;
            return {x < 1.0f ? x : 1.0f, y < 1.0f ? y : 1.0f, z < 1.0f ? z : 1.0f, w < 1.0f ? w : 1.0f};
        }
        case 3: return {s.x * d.x, s.y * d.y, s.z * d.z, s.w * d.w};
        default: return s;
    }
}

// GetDepthTestFunction (Rasterizer.cs:542-559); note the flipped compares.
inline bool depth_pass(int test, float nd, float od) {
    switch (test) {
        case 1: return nd > od;                     // Less
        case 2: return nd >= od;                    // LessEqual
        case 3: return nd < od;                     // Greater
        case 4: return nd <= od;                    // GreaterEqual
        case 5: return fabsf(nd - od) < kEps;       // Equal
        case 6: return fabsf(nd - od) >= kEps;      // NotEqual
        default: return true;                       // Disabled, Always
    }
}

inline float edge_fn(V2 a, V2 b, V2 c) {            // Rasterizer.cs:561-563
    return (c.x - a.x) * (b.y - a.y) - (c.y - a.y) * (b.x - a.x);
}

}  // namespace

// ===========================================================================
// Window (MainWindow.cs:23-31, 378-436) and the triangle pipeline
// ===========================================================================
struct OrcWindow {
    int W, H;
    std::vector<V4> color;
    std::vector<float> depth;
    OrcStats stats;
};

namespace {

// Post-DrawTriangle state of one triangle (Rasterizer.cs:365-398, 411-452).
struct SetupTri {
    V2 s[3];          // screen coords of outputs {v2, v1, v0}
    float d[3];       // depths[i] = (ndc.z + 1) * 0.5
    float cz[3], cw[3];
    V4 col[3];
    V2 uv[3];
    V3 wn[3];
    V3 wpos[3];       // EXTENSION (point lights)
    float invArea;
    float a01, b01, a12, b12, a20, b20;
    int minX, maxX, minY, maxY;
    int draw;
};

// DrawTriangle + head of RasterizeTriangle (Rasterizer.cs:342-452).
// Returns false when the triangle is rejected before the pixel loops.
inline bool setup_triangle(const VOut& v0, const VOut& v1, const VOut& v2, int W, int H,
                           int cull, SetupTri& t) {
    const VOut* o[3] = {&v2, &v1, &v0};               // :367
    for (int i = 0; i < 3; ++i) {
        float invW = 1.0f / o[i]->clip.w;
        float nx = o[i]->clip.x * invW, ny = o[i]->clip.y * invW, nz = o[i]->clip.z * invW;
        if (std::isnan(nx) || std::isnan(ny) || std::isnan(nz) ||
            std::isinf(nx) || std::isinf(ny) || std::isinf(nz))
            return false;                             // :378-380
        t.s[i].x = (nx * 0.5f + 0.5f) * (float)W;     // :383-386
        t.s[i].y = (1.0f - (ny * 0.5f + 0.5f)) * (float)H;
        t.d[i]   = (nz + 1.0f) * 0.5f;                // :388
        t.cz[i] = o[i]->clip.z; t.cw[i] = o[i]->clip.w;
        t.col[i] = o[i]->color; t.uv[i] = o[i]->uv; t.wn[i] = o[i]->wn;
        t.wpos[i] = o[i]->wpos;
    }
    if (v0.clip.w == 0 || v1.clip.w == 0 || v2.clip.w == 0) return false;   // :393
    float area = edge_fn(t.s[0], t.s[1], t.s[2]);     // :396, :411
    if (area == 0) return false;
    bool front = area < 0;                            // :414
    if ((cull == 1 && !front) || (cull == 2 && front)) return false;
    t.invArea = 1.0f / area;                          // :427
    float minXf = net_min(net_min(t.s[0].x, t.s[1].x), t.s[2].x);
    float maxXf = net_max(net_max(t.s[0].x, t.s[1].x), t.s[2].x);
    float minYf = net_min(net_min(t.s[0].y, t.s[1].y), t.s[2].y);
    float maxYf = net_max(net_max(t.s[0].y, t.s[1].y), t.s[2].y);
    int a = sat_int(floorf(minXf)), b = sat_int(ceilf(maxXf));
    int c = sat_int(floorf(minYf)), e = sat_int(ceilf(maxYf));
    t.minX = a > 0 ? a : 0;                           // :437-440
    t.maxX = b < W - 1 ? b : W - 1;
    t.minY = c > 0 ? c : 0;
    t.maxY = e < H - 1 ? e : H - 1;
    if (t.minX > t.maxX || t.minY > t.maxY) return false;
    t.a01 = t.s[0].y - t.s[1].y; t.b01 = t.s[1].x - t.s[0].x;   // :445-447
    t.a12 = t.s[1].y - t.s[2].y; t.b12 = t.s[2].x - t.s[1].x;
    t.a20 = t.s[2].y - t.s[0].y; t.b20 = t.s[0].x - t.s[2].x;
    return true;
}

// Interpolate + InterpolateData (Rasterizer.cs:566-707), stock varyings only.
// Object Normal, ScreenCoords, Barycentric and clip x/y/w are unobservable
// through the stock fragment shader and are not computed.
inline void interpolate(const SetupTri& t, float w0, float w1, float w2,
                        V3& wn, V4& col, V2& uv, float& clipZ, V3* wpos = nullptr) {
    float ra = w0 / t.cw[0], rb = w1 / t.cw[1], rc = w2 / t.cw[2];   // :576-578
    float sum = ra + rb + rc;
    float w = 1.0f / sum;
    float wa = ra * w, wb = rb * w, wc = rc * w;
    clipZ = ((t.cz[0] * ra + t.cz[1] * rb) + t.cz[2] * rc) * w;       // :588-591
    uv.x = ((t.uv[0].x * ra + t.uv[1].x * rb) + t.uv[2].x * rc) * w;  // :593-596
    uv.y = ((t.uv[0].y * ra + t.uv[1].y * rb) + t.uv[2].y * rc) * w;
    col.x = ((t.col[0].x * ra + t.col[1].x * rb) + t.col[2].x * rc) * w;   // :615-618
    col.y = ((t.col[0].y * ra + t.col[1].y * rb) + t.col[2].y * rc) * w;
    col.z = ((t.col[0].z * ra + t.col[1].z * rb) + t.col[2].z * rc) * w;
    col.w = ((t.col[0].w * ra + t.col[1].w * rb) + t.col[2].w * rc) * w;
    V3 n = {(t.wn[0].x * wa + t.wn[1].x * wb) + t.wn[2].x * wc,      // :680-688
            (t.wn[0].y * wa + t.wn[1].y * wb) + t.wn[2].y * wc,
            (t.wn[0].z * wa + t.wn[1].z * wb) + t.wn[2].z * wc};
    float l2 = dot3(n, n);
    if (l2 > 1e-6f) {
        float inv = 1.0f / sqrtf(l2);
        n = {n.x * inv, n.y * inv, n.z * inv};
    }
    wn = n;
    if (wpos) {      // EXTENSION: world position, perspective-correct like the Vector3 varyings of :598-606
        wpos->x = ((t.wpos[0].x * ra + t.wpos[1].x * rb) + t.wpos[2].x * rc) * w;
        wpos->y = ((t.wpos[0].y * ra + t.wpos[1].y * rb) + t.wpos[2].y * rc) * w;
        wpos->z = ((t.wpos[0].z * ra + t.wpos[1].z * rb) + t.wpos[2].z * rc) * w;
    }
}

struct LocalStats { int64_t tested = 0, shaded = 0; };

// Pixel loops of RasterizeTriangle for one tile (Rasterizer.cs:466-536).
inline void raster_tile_fill(OrcWindow& win, const SetupTri& t, const OrcDraw& d, const OrcUniforms& U,
                        int tileX, int tileY, LocalStats& st) {
    const int W = win.W, H = win.H;
    int tsx = tileX * kTile, tsy = tileY * kTile;
    int tex_ = tsx + kTile - 1; if (tex_ > W - 1) tex_ = W - 1;
    int tey = tsy + kTile - 1; if (tey > H - 1) tey = H - 1;
    int startX = t.minX > tsx ? t.minX : tsx, endX = t.maxX < tex_ ? t.maxX : tex_;
    int startY = t.minY > tsy ? t.minY : tsy, endY = t.maxY < tey ? t.maxY : tey;
    if (startX > endX || startY > endY) return;
    float fx = (float)startX, fy = (float)startY;
    float w0Row = t.a12 * (fx - t.s[1].x) + t.b12 * (fy - t.s[1].y);   // :481-483
    float w1Row = t.a20 * (fx - t.s[2].x) + t.b20 * (fy - t.s[2].y);
    float w2Row = t.a01 * (fx - t.s[0].x) + t.b01 * (fy - t.s[0].y);
    const bool needsDepth = d.depth_test != 0;
    const bool canEarlyOut = d.blend == 0;
    for (int y = startY; y <= endY; ++y) {
        float w0 = w0Row, w1 = w1Row, w2 = w2Row;
        for (int x = startX; x <= endX; ++x) {
            bool inside = (w0 >= 0 && w1 >= 0 && w2 >= 0) || (w0 <= 0 && w1 <= 0 && w2 <= 0);
            if (inside) {
                ++st.tested;
                float w0f = w0 * t.invArea, w1f = w1 * t.invArea, w2f = w2 * t.invArea;
                float depth = t.d[0] * w0f + t.d[1] * w1f + t.d[2] * w2f;      // :502
                size_t idx = (size_t)y * W + x;
                float old = win.depth[idx];
                if (depth_pass(d.depth_test, depth, old)) {
                    ++st.shaded;
                    V3 wn; V4 col; V2 uv; float cz;
                    V3 wp = {0.0f, 0.0f, 0.0f};
                    interpolate(t, w0f, w1f, w2f, wn, col, uv, cz, &wp);
                    V4 c = fragment_shader(wn, col, uv, cz, d, U, wp);
                    if (c.w > 0.0f) {                                         // :511
                        win.color[idx] = blend(c, win.color[idx], d.blend);
                        if (needsDepth) win.depth[idx] = depth;
                    } else if (canEarlyOut) {
                        break;                                                // :520-523
                    }
                }
            }
            w0 += t.a12; w1 += t.a20; w2 += t.a01;                            // :527-529
        }
        w0Row += t.b12; w1Row += t.b20; w2Row += t.b01;                       // :532-534
    }
}


// DebugMode.Wireframe (Rasterizer.cs:419-425): the triangle is three DrawLine calls (Rasterizer.cs:232-340)
// over (s0,s1), (s1,s2), (s2,s0) -- every one of them with depths[0..1] and outputs[0..1] of the triangle,
// as the reference passes them.  This is the part of those three calls that falls into one tile; pixels are
// independent of each other, so running line after line per tile is the reference's per-pixel order.
inline void raster_tile_wire(OrcWindow& win, const SetupTri& t, const OrcDraw& d, const OrcUniforms& U,
                             int tileX, int tileY, LocalStats& st) {
    const int W = win.W, H = win.H;
    int tsx = tileX * kTile, tsy = tileY * kTile;
    int tex_ = tsx + kTile - 1; if (tex_ > W - 1) tex_ = W - 1;
    int tey = tsy + kTile - 1; if (tey > H - 1) tey = H - 1;
    SetupTri ends = t;                                   // Interpolate(outputs[0], outputs[1], outputs[0], 1-t, t, 0) (:314)
    ends.cz[2] = t.cz[0]; ends.cw[2] = t.cw[0]; ends.col[2] = t.col[0]; ends.uv[2] = t.uv[0]; ends.wn[2] = t.wn[0];
    ends.wpos[2] = t.wpos[0];
    for (int k = 0; k < 3; ++k) {
        const V2 p0 = t.s[k], p1 = t.s[(k + 1) % 3];
        int minX = sat_int(net_max(net_min(p0.x, p1.x), 0.0f));                    // :242-245
        int maxX = sat_int(net_min(net_max(p0.x, p1.x), (float)(W - 1)));
        int minY = sat_int(net_max(net_min(p0.y, p1.y), 0.0f));
        int maxY = sat_int(net_min(net_max(p0.y, p1.y), (float)(H - 1)));
        if (minX > maxX || minY > maxY) continue;                                  // :247
        int startX = minX > tsx ? minX : tsx, endX = maxX < tex_ ? maxX : tex_;    // :274-277
        int startY = minY > tsy ? minY : tsy, endY = maxY < tey ? maxY : tey;
        if (startX > endX || startY > endY) continue;
        float dx = p1.x - p0.x, dy = p1.y - p0.y;                                  // :257-259
        float lenSq = dx * dx + dy * dy;
        for (int y = startY; y <= endY; ++y) {
            for (int x = startX; x <= endX; ++x) {
                float cx = (float)x + 0.5f, cy = (float)y + 0.5f;
                float px = cx - p0.x, py = cy - p0.y;                              // :296-297
                float tt = 0.0f;
                if (lenSq > 0) tt = (px * dx + py * dy) / lenSq;                   // :299-301
                tt = net_max(0.0f, net_min(1.0f, tt));                             // :303
                float closestX = p0.x + tt * dx, closestY = p0.y + tt * dy;
                float distX = cx - closestX, distY = cy - closestY;
                float distSq = distX * distX + distY * distY;
                if (!(distSq <= 0.25f)) continue;                                  // :312-313
                ++st.tested;
                float omt = 1.0f - tt;
                float depth = 1.0f / (t.d[0] * omt + t.d[1] * tt);                 // :315
                size_t idx = (size_t)y * W + x;
                if (!depth_pass(d.depth_test, depth, win.depth[idx])) continue;    // :318-319
                ++st.shaded;
                V3 wn; V4 col; V2 uv; float cz;
                V3 wp = {0.0f, 0.0f, 0.0f};
                interpolate(ends, omt, tt, 0.0f, wn, col, uv, cz, &wp);
                V4 c = fragment_shader(wn, col, uv, cz, d, U, wp);
                if (c.w != 0.0f) {                                                 // :325 (W != 0, not > 0)
                    win.color[idx] = blend(c, win.color[idx], d.blend);
                    if (d.depth_test != 0) win.depth[idx] = depth;                 // :331-332
                }
            }
        }
    }
}

inline void raster_tile(OrcWindow& win, const SetupTri& t, const OrcDraw& d, const OrcUniforms& U,
                        int tileX, int tileY, LocalStats& st) {
    if (U.debug_mode == 1) raster_tile_wire(win, t, d, U, tileX, tileY, st);
    else raster_tile_fill(win, t, d, U, tileX, tileY, st);
}

// Body of the Parallel.For lambda in RenderMesh (Rasterizer.cs:200-229):
// shade three corners, behind tests, optional near clip, setup.
inline int process_triangle(const OrcDraw& d, const OrcUniforms& U, int tri, int W, int H,
                            int drawIndex, SetupTri out[2]) {
    bool fma = U.numerics_fma != 0;
    VOut v[3];
    for (int k = 0; k < 3; ++k) {
        int vi = d.indices[tri * 3 + k];
        v[k] = vertex_shader(d.vertices + (size_t)vi * 12, d.model, d.view, d.proj, fma);
    }
    bool b0 = v[0].clip.w <= 0, b1 = v[1].clip.w <= 0, b2 = v[2].clip.w <= 0;
    if (b0 && b1 && b2) return 0;
    int n = 0;
    if (b0 || b1 || b2) {
        VOut tris[2][3];
        int nt = clip_near(v, U.near_clip, fma, tris);
        for (int i = 0; i < nt; ++i)
            if (setup_triangle(tris[i][0], tris[i][1], tris[i][2], W, H, d.cull, out[n])) {
                out[n].draw = drawIndex; ++n;
            }
    } else if (setup_triangle(v[0], v[1], v[2], W, H, d.cull, out[0])) {
        out[0].draw = drawIndex; n = 1;
    }
    return n;
}

}  // namespace

extern "C" {

OrcWindow* orc_create(int width, int height) {
    if (width <= 0 || height <= 0) return nullptr;       // Rasterizer.cs:71-74
    OrcWindow* w = new OrcWindow();
    w->W = width; w->H = height;
    w->color.assign((size_t)width * height, V4{0, 0, 0, 0});
    w->depth.assign((size_t)width * height, 0.0f);
    std::memset(&w->stats, 0, sizeof(w->stats));
    return w;
}
void orc_destroy(OrcWindow* w) { delete w; }
int orc_width(const OrcWindow* w) { return w->W; }
int orc_height(const OrcWindow* w) { return w->H; }
float* orc_color(OrcWindow* w) { return &w->color[0].x; }
float* orc_depth(OrcWindow* w) { return w->depth.data(); }

void orc_clear_color(OrcWindow* w, const float rgba[4]) {
    V4 c = {rgba[0], rgba[1], rgba[2], rgba[3]};
    for (auto& p : w->color) p = c;
}
void orc_clear_depth(OrcWindow* w) {
    for (auto& p : w->depth) p = -3.402823466e+38f;     // float.MinValue
    std::memset(&w->stats, 0, sizeof(w->stats));
}
// Both clears on n_threads threads: the reference clears with Parallel.For over the pixels (MainWindow.cs:400-407,
// 429-436); the timed CPU baseline uses this form.
void orc_clear_frame_mt(OrcWindow* w, const float rgba[4], int n_threads) {
    if (n_threads < 1) n_threads = 1;
    const V4 c = {rgba[0], rgba[1], rgba[2], rgba[3]};
    const size_t n = w->color.size();
    auto part = [&](int t) {
        const size_t lo = n * (size_t)t / (size_t)n_threads, hi = n * (size_t)(t + 1) / (size_t)n_threads;
        for (size_t i = lo; i < hi; ++i) { w->color[i] = c; w->depth[i] = -3.402823466e+38f; }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < n_threads; ++t) th.emplace_back(part, t);
    part(0);
    for (auto& t : th) t.join();
    std::memset(&w->stats, 0, sizeof(w->stats));
}
void orc_get_stats(const OrcWindow* w, OrcStats* out) { *out = w->stats; }

void orc_render_mesh(OrcWindow* w, const OrcDraw* d, const OrcUniforms* u) {
    if (w->W <= 0 || w->H <= 0) return;                  // :176
    int ntri = d->n_indices / 3;
    LocalStats st;
    for (int tri = 0; tri < ntri; ++tri) {
        SetupTri ts[2];
        int n = process_triangle(*d, *u, tri, w->W, w->H, 0, ts);
        w->stats.triangles_in++;
        for (int i = 0; i < n; ++i) {
            const SetupTri& t = ts[i];
            w->stats.triangles_drawn++;
            int tminx = t.minX / kTile, tmaxx = t.maxX / kTile;      // :449-452
            int tminy = t.minY / kTile, tmaxy = t.maxY / kTile;
            for (int ty = tminy; ty <= tmaxy; ++ty)
                for (int tx = tminx; tx <= tmaxx; ++tx)
                    raster_tile(*w, t, *d, *u, tx, ty, st);
        }
    }
    w->stats.fragments_tested += st.tested;
    w->stats.fragments_shaded += st.shaded;
}

void orc_render_frame_mt(OrcWindow* w, const OrcDraw* draws, int n_draws, const OrcUniforms* u,
                         int n_threads, int tile_row0, int tile_row1) {
    if (w->W <= 0 || w->H <= 0 || n_draws <= 0) return;
    if (n_threads < 1) n_threads = 1;
    const int W = w->W, H = w->H;
    const int tilesY = (H + kTile - 1) / kTile;
    if (tile_row0 < 0) tile_row0 = 0;
    if (tile_row1 > tilesY) tile_row1 = tilesY;

    const bool trace = std::getenv("ORC_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count(); };
    const auto tA = now();
    // Phase A: triangle processing, parallel over contiguous chunks of the
    // global (draw, triangle) sequence; chunk order == submission order.
    std::vector<int64_t> triBase(n_draws + 1, 0);
    for (int i = 0; i < n_draws; ++i) triBase[i + 1] = triBase[i] + draws[i].n_indices / 3;
    const int64_t total = triBase[n_draws];
    const int nChunks = n_threads * 8;
    std::vector<std::vector<SetupTri>> chunks(nChunks);
    std::atomic<int> nextChunk{0};
    auto phaseA = [&]() {
        for (;;) {
            int c = nextChunk.fetch_add(1);
            if (c >= nChunks) break;
            int64_t lo = total * c / nChunks, hi = total * (c + 1) / nChunks;
            int di = 0;
            while (di < n_draws && triBase[di + 1] <= lo) ++di;
            auto& out = chunks[c];
            for (int64_t g = lo; g < hi; ++g) {
                while (triBase[di + 1] <= g) ++di;
                SetupTri ts[2];
                int n = process_triangle(draws[di], *u, (int)(g - triBase[di]), W, H, di, ts);
                for (int i = 0; i < n; ++i) out.push_back(ts[i]);
            }
        }
    };
    {
        std::vector<std::thread> th;
        for (int i = 1; i < n_threads; ++i) th.emplace_back(phaseA);
        phaseA();
        for (auto& t : th) t.join();
    }
    // Work units of phase B: runs of kSegTiles tiles of one tile row.  Every unit's list is in submission order
    // (chunks in order, triangles in order within a chunk), so every tile still sees its triangles in the
    // reference's serial order; units are handed out longest list first.  (Whole tile rows as units left most
    // threads idle while a few walked the horizon rows: 4.5x on 16-32 threads.)
    const auto tB = now();
    constexpr int kSegTiles = 8;
    const int tilesX = (W + kTile - 1) / kTile;
    const int segsX = (tilesX + kSegTiles - 1) / kSegTiles;
    std::vector<const SetupTri*> all;
    for (auto& c : chunks) for (auto& t : c) all.push_back(&t);
    std::vector<size_t> unitStart((size_t)tilesY * segsX, 0), unitEnd((size_t)tilesY * segsX, 0);
    std::vector<uint32_t> unitList;
    {
        // bucket build: first (in parallel) the tile ranges of all triangles as a compact array -- the setup records are
        // 250 bytes each --, then two serial passes over it
        struct Range { uint16_t x0, x1, y0, y1; };
        std::vector<Range> rg(all.size());
        auto in_threads = [&](auto&& fn) {
            std::vector<std::thread> th;
            for (int i = 1; i < n_threads; ++i) th.emplace_back(fn, i);
            fn(0);
            for (auto& t : th) t.join();
        };
        in_threads([&](int k) {
            const size_t lo = all.size() * (size_t)k / (size_t)n_threads, hi = all.size() * (size_t)(k + 1) / (size_t)n_threads;
            for (size_t i = lo; i < hi; ++i) {
                const SetupTri& t = *all[i];
                rg[i] = {(uint16_t)(t.minX / kTile), (uint16_t)(t.maxX / kTile), (uint16_t)(t.minY / kTile), (uint16_t)(t.maxY / kTile)};
            }
        });
        // count, prefix, fill over the compact array: one flat list, no reallocation.  Serial on purpose: half a million
        // visits take 2-3 ms, less than starting threads for them and keeping their counters off each other's lines.
        auto scan = [&](auto&& visit) {
            for (size_t i = 0; i < rg.size(); ++i) {
                const int a = std::max((int)rg[i].y0, tile_row0), b = std::min((int)rg[i].y1, tile_row1 - 1);
                const int s0 = rg[i].x0 / kSegTiles, s1 = rg[i].x1 / kSegTiles;
                for (int r = a; r <= b; ++r)
                    for (int sgi = s0; sgi <= s1; ++sgi) visit((size_t)r * segsX + sgi, (uint32_t)i);
            }
        };
        scan([&](size_t unit, uint32_t) { ++unitEnd[unit]; });
        size_t acc = 0;
        for (size_t k = 0; k < unitEnd.size(); ++k) { unitStart[k] = acc; acc += unitEnd[k]; unitEnd[k] = unitStart[k]; }
        unitList.resize(acc);
        scan([&](size_t unit, uint32_t i) { unitList[unitEnd[unit]++] = i; });
    }
    const auto tC = now();
    std::vector<uint32_t> orderU;
    for (size_t k = 0; k < unitStart.size(); ++k) if (unitEnd[k] > unitStart[k]) orderU.push_back((uint32_t)k);
    std::stable_sort(orderU.begin(), orderU.end(),
                     [&](uint32_t a, uint32_t b) { return unitEnd[a] - unitStart[a] > unitEnd[b] - unitStart[b]; });
    // Phase B: each thread takes whole units; a tile belongs to exactly one unit.
    std::atomic<size_t> nextUnit{0};
    std::atomic<int64_t> tested{0}, shaded{0};
    auto phaseB = [&]() {
        LocalStats st;
        for (;;) {
            const size_t k = nextUnit.fetch_add(1);
            if (k >= orderU.size()) break;
            const uint32_t unit = orderU[k];
            const int r = (int)(unit / segsX), sgi = (int)(unit % segsX);
            const int u0 = sgi * kSegTiles, u1 = std::min(u0 + kSegTiles, tilesX) - 1;
            for (size_t q = unitStart[unit]; q < unitEnd[unit]; ++q) {
                const SetupTri& t = *all[unitList[q]];
                const int tminx = std::max(t.minX / kTile, u0), tmaxx = std::min(t.maxX / kTile, u1);
                for (int tx = tminx; tx <= tmaxx; ++tx)
                    raster_tile(*w, t, draws[t.draw], *u, tx, r, st);
            }
        }
        tested += st.tested; shaded += st.shaded;
    };
    {
        std::vector<std::thread> th;
        for (int i = 1; i < n_threads; ++i) th.emplace_back(phaseB);
        phaseB();
        for (auto& t : th) t.join();
    }
    if (trace) std::fprintf(stderr, "orc mt: phase A %.1f ms, buckets %.1f ms, phase B %.1f ms (%zu units)\n", ms(tA, tB), ms(tB, tC), ms(tC, now()), orderU.size());
    w->stats.triangles_in += total;
    w->stats.triangles_drawn += (int64_t)all.size();
    w->stats.fragments_tested += tested.load();
    w->stats.fragments_shaded += shaded.load();
}

void orc_texture_sample(const uint8_t* rgba8, int tw, int th, float u, float v, int filter, float out[4]) {
    V4 r = filter == 1 ? tex_bilinear(rgba8, tw, th, u, v) : tex_nearest(rgba8, tw, th, u, v);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}

void orc_vertex_shader(const float vtx[12], const float model[16], const float view[16],
                       const float proj[16], int numerics_fma, float out_clip[4], float out_wn[3]) {
    VOut o = vertex_shader(vtx, model, view, proj, numerics_fma != 0);
    out_clip[0] = o.clip.x; out_clip[1] = o.clip.y; out_clip[2] = o.clip.z; out_clip[3] = o.clip.w;
    out_wn[0] = o.wn.x; out_wn[1] = o.wn.y; out_wn[2] = o.wn.z;
}

int orc_clip_near(const float clip_in[12], float near_clip, int numerics_fma, float out_clip[24]) {
    VOut v[3];
    std::memset(v, 0, sizeof(v));
    for (int i = 0; i < 3; ++i)
        v[i].clip = {clip_in[4 * i], clip_in[4 * i + 1], clip_in[4 * i + 2], clip_in[4 * i + 3]};
    VOut tris[2][3];
    int n = clip_near(v, near_clip, numerics_fma != 0, tris);
    for (int t = 0; t < n; ++t)
        for (int k = 0; k < 3; ++k) {
            float* o = out_clip + (t * 3 + k) * 4;
            o[0] = tris[t][k].clip.x; o[1] = tris[t][k].clip.y;
            o[2] = tris[t][k].clip.z; o[3] = tris[t][k].clip.w;
        }
    return n;
}

// ---------------------------------------------------------------------------
// Physics.Raycast (Physics.cs:19-179)
// ---------------------------------------------------------------------------
// Matrix4x4.Invert, .NET BCL scalar algorithm (cofactor expansion along the first column; fails when
// |det| < float.Epsilon).  Row-major M11..M44.
static bool invert4(const float* M, float* R) {
    float a = M[0], b = M[1], c = M[2], d = M[3], e = M[4], f = M[5], g = M[6], h = M[7];
    float i = M[8], j = M[9], k = M[10], l = M[11], m = M[12], n = M[13], o = M[14], p = M[15];
    float kp_lo = k * p - l * o, jp_ln = j * p - l * n, jo_kn = j * o - k * n;
    float ip_lm = i * p - l * m, io_km = i * o - k * m, in_jm = i * n - j * m;
    float a11 = +(f * kp_lo - g * jp_ln + h * jo_kn);
    float a12 = -(e * kp_lo - g * ip_lm + h * io_km);
    float a13 = +(e * jp_ln - f * ip_lm + h * in_jm);
    float a14 = -(e * jo_kn - f * io_km + g * in_jm);
    float det = a * a11 + b * a12 + c * a13 + d * a14;
    if (fabsf(det) < 1.401298464e-45f) return false;
    float invDet = 1.0f / det;
    R[0] = a11 * invDet; R[4] = a12 * invDet; R[8] = a13 * invDet; R[12] = a14 * invDet;
    R[1] = -(b * kp_lo - c * jp_ln + d * jo_kn) * invDet;
    R[5] = +(a * kp_lo - c * ip_lm + d * io_km) * invDet;
    R[9] = -(a * jp_ln - b * ip_lm + d * in_jm) * invDet;
    R[13] = +(a * jo_kn - b * io_km + c * in_jm) * invDet;
    float gp_ho = g * p - h * o, fp_hn = f * p - h * n, fo_gn = f * o - g * n;
    float ep_hm = e * p - h * m, eo_gm = e * o - g * m, en_fm = e * n - f * m;
    R[2] = +(b * gp_ho - c * fp_hn + d * fo_gn) * invDet;
    R[6] = -(a * gp_ho - c * ep_hm + d * eo_gm) * invDet;
    R[10] = +(a * fp_hn - b * ep_hm + d * en_fm) * invDet;
    R[14] = -(a * fo_gn - b * eo_gm + c * en_fm) * invDet;
    float gl_hk = g * l - h * k, fl_hj = f * l - h * j, fk_gj = f * k - g * j;
    float el_hi = e * l - h * i, ek_gi = e * k - g * i, ej_fi = e * j - f * i;
    R[3] = -(b * gl_hk - c * fl_hj + d * fk_gj) * invDet;
    R[7] = +(a * gl_hk - c * el_hi + d * ek_gi) * invDet;
    R[11] = -(a * fl_hj - b * el_hi + d * ej_fi) * invDet;
    R[15] = +(a * fk_gj - b * ek_gi + c * ej_fi) * invDet;
    return true;
}

static inline V3 cross3(V3 a, V3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// RayIntersectsTriangle (Physics.cs:137-178).
static bool ray_tri(V3 ro, V3 rd, V3 v0, V3 v1, V3 v2, int faceMask, float& distance, V3& bary) {
    const float Epsilon = 1e-8f;
    V3 edge1 = {v1.x - v0.x, v1.y - v0.y, v1.z - v0.z};
    V3 edge2 = {v2.x - v0.x, v2.y - v0.y, v2.z - v0.z};
    V3 pvec = cross3(rd, edge2);
    float det = dot3(edge1, pvec);
    if ((faceMask & 1) && det < Epsilon) return false;              // :156
    if ((faceMask & 2) && det > -Epsilon) return false;             // :157
    if (fabsf(det) < Epsilon) return false;                         // :158
    float invDet = 1.0f / det;
    V3 tvec = {ro.x - v0.x, ro.y - v0.y, ro.z - v0.z};
    float u = dot3(tvec, pvec) * invDet;
    if (u < 0 || u > 1) return false;
    V3 qvec = cross3(tvec, edge1);
    float v = dot3(rd, qvec) * invDet;
    if (v < 0 || u + v > 1) return false;
    distance = dot3(edge2, qvec) * invDet;
    if (distance < 0) return false;
    bary = {1.0f - u - v, u, v};
    return true;
}

int orc_raycast(const float origin[3], const float direction[3], const float* vertices, int n_vertices,
                const uint16_t* indices, int n_indices, const float model[16], int face_mask, int numerics_fma,
                float* out_distance, float out_point[3], float out_normal[3], int* out_tri) {
    bool fma = numerics_fma != 0;
    *out_distance = 3.402823466e+38f;                                // float.MaxValue (:31, :66)
    out_point[0] = out_point[1] = out_point[2] = 0.0f;
    out_normal[0] = out_normal[1] = out_normal[2] = 0.0f;
    if (out_tri) *out_tri = -1;
    float inv[16];
    if (!invert4(model, inv)) return 0;                              // :30-36
    float nmat[16];                                                  // Matrix4x4.Transpose(invModel) (:38)
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) nmat[4 * r + c] = inv[4 * c + r];
    std::vector<V3> wp(n_vertices), wn(n_vertices);
    for (int i = 0; i < n_vertices; ++i) {                           // :43-49
        const float* v = vertices + (size_t)i * 12;
        V4 p = transform4({v[0], v[1], v[2], 1.0f}, model, fma);     // Vector3.Transform(position, model)
        wp[i] = {p.x, p.y, p.z};
        V4 n = transform4({v[5], v[6], v[7], 0.0f}, nmat, fma);      // Vector4.Transform((normal,0), normalMatrix)
        wn[i] = normalize3({n.x, n.y, n.z});
    }
    V3 ro = {origin[0], origin[1], origin[2]};
    V3 rd = normalize3({direction[0], direction[1], direction[2]});  // :70
    float best = 3.402823466e+38f;
    int bestTri = -1;
    V3 bestBary = {0, 0, 0};
    for (int t = 0; t < n_indices / 3; ++t) {
        int i0 = indices[3 * t], i1 = indices[3 * t + 1], i2 = indices[3 * t + 2];
        float dist; V3 bary;
        if (!ray_tri(ro, rd, wp[i0], wp[i1], wp[i2], face_mask, dist, bary)) continue;
        if (dist < best) { best = dist; bestTri = t; bestBary = bary; }   // :104-107, strict: first (lowest index) wins ties
    }
    if (bestTri < 0) return 0;
    int i0 = indices[3 * bestTri], i1 = indices[3 * bestTri + 1], i2 = indices[3 * bestTri + 2];
    V3 n = {(wn[i0].x * bestBary.x + wn[i1].x * bestBary.y) + wn[i2].x * bestBary.z,       // :101
            (wn[i0].y * bestBary.x + wn[i1].y * bestBary.y) + wn[i2].y * bestBary.z,
            (wn[i0].z * bestBary.x + wn[i1].z * bestBary.y) + wn[i2].z * bestBary.z};
    n = normalize3(n);
    *out_distance = best;
    out_point[0] = ro.x + rd.x * best; out_point[1] = ro.y + rd.y * best; out_point[2] = ro.z + rd.z * best;   // :102
    out_normal[0] = n.x; out_normal[1] = n.y; out_normal[2] = n.z;
    if (out_tri) *out_tri = bestTri;
    return 1;
}

// FrustumCuller.IsSphereInFrustum (FrustumCuller.cs:153-224, 19-35).
int orc_sphere_in_frustum(const float sphere[4], const float M[16], const float V[16],
                          const float P[16], int numerics_fma) {
    bool fma = numerics_fma != 0;
    V4 wc4 = transform4({sphere[0], sphere[1], sphere[2], 1.0f}, M, fma);   // Vector3.Transform
    V3 c = {wc4.x, wc4.y, wc4.z};
    float s0 = sqrtf(M[0] * M[0] + M[1] * M[1] + M[2] * M[2]);
    float s1 = sqrtf(M[4] * M[4] + M[5] * M[5] + M[6] * M[6]);
    float s2 = sqrtf(M[8] * M[8] + M[9] * M[9] + M[10] * M[10]);
    float r = sphere[3] * net_max(net_max(s0, s1), s2);
    float VP[16];                                        // Matrix4x4.Multiply(view, proj)
    for (int i = 0; i < 4; ++i) {
        V4 row = transform4({V[4 * i], V[4 * i + 1], V[4 * i + 2], V[4 * i + 3]}, P, fma);
        VP[4 * i] = row.x; VP[4 * i + 1] = row.y; VP[4 * i + 2] = row.z; VP[4 * i + 3] = row.w;
    }
    // column k of VP (0-based): M1k..M4k = VP[k], VP[4+k], VP[8+k], VP[12+k]
    auto plane_ok = [&](int k, float sign) {
        float x = VP[3] + sign * VP[k], y = VP[7] + sign * VP[4 + k];
        float z = VP[11] + sign * VP[8 + k], d = VP[15] + sign * VP[12 + k];
        float mag = sqrtf(x * x + y * y + z * z);        // NormalizePlane
        V3 n = {x / mag, y / mag, z / mag};
        float dist = d / mag;
        n = normalize3(n);                               // Plane ctor renormalises (:26)
        return dot3(n, c) + dist > -r;                   // TestSphereAgainstPlane
    };
    // order: Left, Right, Top, Bottom, Near, Far (:211-216)
    return plane_ok(0, 1.0f) && plane_ok(0, -1.0f) && plane_ok(1, 1.0f) && plane_ok(1, -1.0f) &&
           plane_ok(2, 1.0f) && plane_ok(2, -1.0f);
}

}  // extern "C"
