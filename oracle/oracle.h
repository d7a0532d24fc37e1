/*
 * oracle.h -- C interface of the CPU oracle (TEST INFRASTRUCTURE, not product).
 *
 * The oracle is a C++ restatement of the per-frame rasterization path of
 * OCSYT/SoftwareRenderer (C#/.NET 9), written from the reference's semantics:
 *   Rasterizer.cs:58-65, 95-230, 342-539, 542-707; Shaders.cs:10-95;
 *   Renderer.cs:39-46, 830-860; Texture.cs:42-63; MainWindow.cs:378-436;
 *   FrustumCuller.cs:19-35, 153-224.
 *
 * PARITY UNPINNED: the reference ships no tests, golden images or fixtures and
 * cannot run here (no .NET runtime, no graphics stack), so this restatement
 * is pinned only by hand-derived known-answer tests (tests/test_oracle_kat.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product (libsrb200.so)
 * never links or calls it.
 */
#ifndef SRB_ORACLE_H
#define SRB_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Captured state of the stock shader pair (Renderer.cs:39-44) + switches. */
typedef struct OrcUniforms {
    float light_dir[4];      /* xyz = LightDirection (already normalised by the caller) */
    float light_color[4];
    float fog_color[4];
    float fog_start;
    float fog_end;
    float near_clip;         /* Rasterizer.NearClip */
    int32_t numerics_fma;    /* 1: System.Numerics Transform/Lerp fused (net9 on FMA hw) */
    int32_t tex_filter;      /* 0 nearest+repeat (reference); 1 bilinear+repeat (extension) */
    int32_t n_extra_lights;  /* extension: extra lights, 0..4 */
    int32_t debug_mode;      /* Rasterizer.RenderDebugMode: 0 None, 1 Wireframe (Rasterizer.cs:419-425) */
    int32_t _pad;
    float extra_dir[4][4];   /* w == 0: directional, xyz = direction; w != 0: point light, xyz = world position */
    float extra_color[4][4];
    float extra_atten[4][4]; /* point lights: AttenuationConstant, Linear, Quadratic (Light.cs:13-15) */
} OrcUniforms;

/* One RenderMesh call (Rasterizer.cs:163-174). */
typedef struct OrcDraw {
    const float*    vertices;   /* nV * 12 floats: pos3 uv2 normal3 color4 (Shaders.cs:10-24) */
    const uint16_t* indices;    /* nI ushort */
    const uint8_t*  tex_rgba8;  /* NULL = no texture (sample = 1,1,1,1; Renderer.cs:852) */
    int32_t n_vertices;
    int32_t n_indices;
    int32_t tex_w, tex_h;
    float model[16];            /* row-major M11..M44, row-vector convention */
    float view[16];
    float proj[16];
    int32_t cull;               /* 0 None 1 Back 2 Front  (Rasterizer.cs:45-50) */
    int32_t depth_test;         /* 0 Disabled 1 Less 2 LessEqual 3 Greater 4 GreaterEqual 5 Equal 6 NotEqual 7 Always */
    int32_t blend;              /* 0 None 1 Alpha 2 Additive 3 Multiply */
    int32_t _pad;
} OrcDraw;

typedef struct OrcWindow OrcWindow;

OrcWindow* orc_create(int width, int height);
void  orc_destroy(OrcWindow* w);
int   orc_width(const OrcWindow* w);
int   orc_height(const OrcWindow* w);
float* orc_color(OrcWindow* w);   /* W*H*4 floats */
float* orc_depth(OrcWindow* w);   /* W*H floats   */
void  orc_clear_color(OrcWindow* w, const float rgba[4]);  /* MainWindow.cs:400-407 */
void  orc_clear_depth(OrcWindow* w);                       /* MainWindow.cs:429-436 */
void  orc_clear_frame_mt(OrcWindow* w, const float rgba[4], int n_threads);   /* both, Parallel.For as in the reference */

/* Literal single-thread path, reference loop order (triangle -> tile row -> tile -> y -> x). */
void orc_render_mesh(OrcWindow* w, const OrcDraw* d, const OrcUniforms* u);

/* Tile-row-parallel variant: same arithmetic, draws/triangles in submission
 * order per tile; bit-identical to calling orc_render_mesh for each draw.
 * Restricts rasterisation to tile rows [tile_row0, tile_row1) (band support);
 * pass 0, INT_MAX for the full frame. */
void orc_render_frame_mt(OrcWindow* w, const OrcDraw* draws, int n_draws,
                         const OrcUniforms* u, int n_threads,
                         int tile_row0, int tile_row1);

/* FrustumCuller.IsSphereInFrustum (FrustumCuller.cs:201-218). sphere = cx,cy,cz,r */
int orc_sphere_in_frustum(const float sphere[4], const float model[16],
                          const float view[16], const float proj[16], int numerics_fma);

/* Texture.Sample (Texture.cs:42-63) / bilinear extension. */
void orc_texture_sample(const uint8_t* rgba8, int tw, int th, float u, float v,
                        int filter, float out[4]);

/* Stock vertex shader on one vertex (Renderer.cs:830-846): out = clip4, worldNormal3. */
void orc_vertex_shader(const float vtx[12], const float model[16], const float view[16],
                       const float proj[16], int numerics_fma, float out_clip[4], float out_wn[3]);

/* Near-plane clip of one triangle (Rasterizer.cs:95-160) on clip positions only:
 * returns number of output triangles (0..2), out_clip = n*3*4 floats. */
int orc_clip_near(const float clip_in[12], float near_clip, int numerics_fma, float out_clip[24]);

/* Physics.Raycast (Physics.cs:19-179): one ray against one mesh under `model`; closest hit by
 * Moller-Trumbore in world space.  face_mask: 1 IgnoreBackfaces, 2 IgnoreFrontfaces (Physics.cs:9-14).
 * The reference reduces per-thread partial results in completion order; ties in distance are broken here
 * by the lowest triangle index (the sequential order, one legal execution).
 * Matrix4x4.Invert / Transpose come from the .NET 9 BCL (not in the reference tree, unpinned): restated from
 * the BCL's published scalar cofactor algorithm; they only feed hit_normal.
 * Returns 1 on a hit.  out_tri receives the winning triangle index (not in the reference's signature). */
int orc_raycast(const float origin[3], const float direction[3], const float* vertices, int n_vertices,
                const uint16_t* indices, int n_indices, const float model[16], int face_mask, int numerics_fma,
                float* out_distance, float out_point[3], float out_normal[3], int* out_tri);

/* Debug statistics of the last orc_render_* call chain since orc_clear_depth. */
typedef struct OrcStats {
    int64_t triangles_in;       /* submitted */
    int64_t triangles_drawn;    /* reached RasterizeTriangle's pixel loops */
    int64_t fragments_tested;   /* centerInside */
    int64_t fragments_shaded;   /* depth test passed */
} OrcStats;
void orc_get_stats(const OrcWindow* w, OrcStats* out);

#ifdef __cplusplus
}
#endif
#endif
