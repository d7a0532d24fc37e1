/*
 * srb200.h -- C ABI of libsrb200.so, the B200 (sm_100a) rasterization backend
 * that sits under OCSYT/SoftwareRenderer's C# Renderer/Rasterizer/MainWindow/
 * Texture API.  POD types only: plain pointers, sizes and int status codes.
 *
 * The reference has no FFI of its own; each entry point below names the C#
 * member (file:line under /root/reference) whose BODY it replaces.  The C#
 * side binds these with [DllImport("srb200")] (csharp/Native.cs,
 * INTEGRATION.md).  There is no CPU fallback anywhere behind this header:
 * every call that needs the GPU fails with SRB_ERR_CUDA when there is none.
 *
 * Conventions:
 *   - Matrix4x4 = 16 floats row-major M11..M44, row-vector convention v' = v*M.
 *   - VertexInput = 12 floats: pos3, uv2, normal3, color4 (Shaders.cs:10-24).
 *   - colour buffer = float4 per pixel, depth buffer = float per pixel, both
 *     row-major, top-left origin, index = y*W + x (MainWindow.cs:30-31, 379).
 *   - stored depth is the reference's negated screen-space depth; the clear
 *     value is -FLT_MAX (float.MinValue, MainWindow.cs:429-436).
 *   - status: 0 = ok, <0 = error; srb_last_error() gives the message.
 *   - srb_draw* may be called from several host threads at once
 *     (Renderer.cs:444, 479 call RenderMesh from Parallel.ForEach); everything
 *     else is called from one thread.  Work is deferred to srb_end_frame.
 */
#ifndef SRB200_H
#define SRB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRB_OK             0
#define SRB_ERR_ARG       -1   /* ArgumentException analogue (Rasterizer.cs:71-74) */
#define SRB_ERR_CUDA      -2   /* CUDA runtime error or no device */
#define SRB_ERR_HANDLE    -3   /* unknown mesh/texture handle */
#define SRB_ERR_STATE     -4   /* call outside begin/end frame etc. */
#define SRB_ERR_UNSUPPORTED -5 /* non-stock shader delegates: no CPU fallback by design */

/* Rasterizer.CullMode / DepthTest / BlendMode (Rasterizer.cs:25-50), same values. */
enum { SRB_CULL_NONE = 0, SRB_CULL_BACK = 1, SRB_CULL_FRONT = 2 };
enum { SRB_DEPTH_DISABLED = 0, SRB_DEPTH_LESS = 1, SRB_DEPTH_LESS_EQUAL = 2, SRB_DEPTH_GREATER = 3,
       SRB_DEPTH_GREATER_EQUAL = 4, SRB_DEPTH_EQUAL = 5, SRB_DEPTH_NOT_EQUAL = 6, SRB_DEPTH_ALWAYS = 7 };
enum { SRB_BLEND_NONE = 0, SRB_BLEND_ALPHA = 1, SRB_BLEND_ADDITIVE = 2, SRB_BLEND_MULTIPLY = 3 };
/* Rasterizer.DebugMode (Rasterizer.cs:14-18), same values. */
enum { SRB_DEBUG_NONE = 0, SRB_DEBUG_WIREFRAME = 1 };

/* srb_begin_frame flags */
#define SRB_CLEAR_COLOR 1u     /* MainWindow.ClearColorBuffer (MainWindow.cs:400-407) */
#define SRB_CLEAR_DEPTH 2u     /* MainWindow.ClearDepthBuffer (MainWindow.cs:429-436) */

/* srb_draw flags */
#define SRB_DRAW_FRUSTUM_CULL 1u  /* run FrustumCuller.IsSphereInFrustum on the device for this draw */

/* pipeline selection (srb_set_pipeline) */
enum { SRB_PIPELINE_AUTO = 0,      /* visibility path when provably order-independent, else ordered */
       SRB_PIPELINE_ORDERED = 1,   /* always the submission-ordered path (any state) */
       SRB_PIPELINE_VISIBILITY = 2 /* force; SRB_ERR_STATE at end_frame if the frame is not eligible */ };

typedef struct SrbContext SrbContext;
typedef int32_t SrbMesh;      /* >0 valid */
typedef int32_t SrbTexture;   /* 0 = no texture (sample = 1,1,1,1; Renderer.cs:852), >0 valid */

/* Captured state of the stock shader pair (Renderer.cs:39-44, 848-860) plus
 * Rasterizer.NearClip (Rasterizer.cs:20) and the numerics switch. */
typedef struct SrbUniforms {
    float light_dir[4];       /* xyz = Renderer.LightDirection (normalised by the caller) */
    float light_color[4];     /* Renderer.LightColor */
    float fog_color[4];       /* Renderer.FogColor */
    float fog_start;          /* Renderer.FogStart */
    float fog_end;            /* Renderer.FogEnd */
    float near_clip;          /* Rasterizer.NearClip */
    int32_t numerics_fma;     /* 1 = System.Numerics Transform/Lerp fused (net9.0 on FMA hardware) */
    int32_t tex_filter;       /* 0 nearest+repeat (Texture.cs:42-63); 1 bilinear+repeat (extension) */
    int32_t n_extra_lights;   /* extension (BASELINE config 5): 0..4 extra lights */
    int32_t debug_mode;       /* Rasterizer.RenderDebugMode (Rasterizer.cs:14-22): 0 None, 1 Wireframe (every triangle is
                                 drawn as three DrawLine calls, Rasterizer.cs:419-425, 232-340; ordered pipeline only) */
    int32_t _pad;
    float extra_dir[4][4];    /* w == 0: directional light, xyz = Light.Direction; w != 0: point light, xyz = Light.Position */
    float extra_color[4][4];  /* Light.Color */
    float extra_atten[4][4];  /* point lights: AttenuationConstant, AttenuationLinear, AttenuationQuadratic (Light.cs:13-15) */
} SrbUniforms;

/* One Rasterizer.RenderMesh call (Rasterizer.cs:163-174) for batched submission. */
typedef struct SrbDrawCmd {
    SrbMesh    mesh;
    SrbTexture texture;
    float model[16];
    float view[16];
    float proj[16];
    float sphere[4];          /* Mesh.SphereBounds cx,cy,cz,r; read when SRB_DRAW_FRUSTUM_CULL */
    int32_t cull;
    int32_t depth_test;
    int32_t blend;
    uint32_t flags;
    int32_t submission_index; /* deterministic order key despite racing callers; <0 = call order */
    int32_t _pad[3];
} SrbDrawCmd;

typedef struct SrbStats {
    int64_t draws_submitted;
    int64_t draws_visible;        /* after the device frustum test */
    int64_t triangles_submitted;
    int64_t triangles_setup;      /* survived behind/NaN/area/cull/bbox(/band) and may cover a pixel (binned + direct) */
    int64_t tile_pairs;           /* (triangle, tile) bin entries */
    int64_t clipped_triangles;    /* went through the near-plane clipper */
    int32_t pipeline_used;        /* SRB_PIPELINE_ORDERED or SRB_PIPELINE_VISIBILITY */
    int32_t kernel_launches;      /* CUDA kernels launched by the last end_frame */
    float   gpu_ms;               /* device time of the last end_frame (CUDA events) */
    int32_t triangles_no_coverage;/* accepted by the reference's setup, but at most 4x4 px and proven to pass its
                                     coverage test nowhere: not binned (triangles_setup + this = the reference's count) */
    int32_t split_parts;          /* CTAs beyond one per tile that the visibility tile kernel spent on split bins */
    int32_t _pad;
} SrbStats;

/* ---- context / framebuffer: MainWindow buffers (MainWindow.cs:23-31, 298-361) ---- */
int srb_create(int device, int width, int height, SrbContext** out_ctx);
int srb_destroy(SrbContext* ctx);
int srb_resize(SrbContext* ctx, int width, int height);          /* MainWindow.HandleResize :298-361 */
int srb_get_size(SrbContext* ctx, int* width, int* height);      /* RenderWidth/RenderHeight :25-26 */
const char* srb_last_error(SrbContext* ctx);                     /* ctx may be NULL for create errors */
/* Multi-GPU: restrict this context to tile rows [row0,row1) (16-px rows). */
int srb_set_band(SrbContext* ctx, int tile_row0, int tile_row1);
int srb_set_pipeline(SrbContext* ctx, int pipeline);
/* Tuning (no reference counterpart; the image never depends on it): a tile whose bin holds at least min_entries
 * (triangle, tile) pairs is rasterized by several CTAs of part_entries (a multiple of 32) each, merged through 64-bit atomic max; one
 * threshold for contexts that render the whole frame, one for band contexts (multi-GPU).  0 = never split.
 * Defaults 0 / 256 / 128. */
int srb_set_bin_split(SrbContext* ctx, int min_entries_full_frame, int min_entries_band, int part_entries);
/* Tuning (the image never depends on it): how many frames may be in flight on the device, 2..4 buffer sets, or 0 =
 * chosen per frame (default).  With 2 the geometry stages of frame N+1 overlap the tile kernel of frame N; with
 * more, the geometry chains of consecutive frames also overlap each other (each set has its own stream) -- for
 * frames whose tile kernel is shorter than their chain of geometry launches: small frames, and the band contexts
 * of a multi-GPU frame. */
int srb_set_pipeline_depth(SrbContext* ctx, int buffer_sets);
/* Tuning: replay a frame's chain of geometry launches as an instantiated CUDA graph when the frame repeats the
 * shape (parameters, uniforms, grids, buffers) of the one before it on the same buffer set.  On by default; a
 * context whose capture fails falls back to plain launches by itself. */
int srb_set_graphs(SrbContext* ctx, int enabled);
/* Multi-GPU band gather fused into the tile kernel (no reference counterpart):
 * rank 0 exports its colour buffer (CUDA IPC handle, 64 bytes); a band rank passes that handle as
 * its colour target and from then on stores its band's final colour straight into rank 0's
 * buffer over NVLink (peer stores), so that no separate gather copy is needed -- only a barrier.
 * handle == NULL restores the local buffer.  Needs frames that clear colour. */
#define SRB_IPC_HANDLE_BYTES 64
int srb_export_color_handle(SrbContext* ctx, uint8_t handle_out[SRB_IPC_HANDLE_BYTES]);
int srb_set_color_target(SrbContext* ctx, const uint8_t handle[SRB_IPC_HANDLE_BYTES]);
/* The same gather at display precision (payload float4 or RGBA8): rank 0 exports an RGBA8 frame
 * (W*H*4 bytes, allocated by the call); from then on its own tile kernel stores clamp(c,0,1)*255 (round to
 * nearest) of its band there, beside the float4 ColorBuffer.  A band rank that passes the handle as its target
 * stores its band's bytes straight into rank 0's frame over NVLink (4 B per pixel instead of 16) and keeps its
 * float4 colour local.  handle == NULL turns it off.  srb_read_gathered_rgba8 copies the assembled frame to
 * the host (after the caller's frame barrier across ranks). */
int srb_gather_rgba8_export(SrbContext* ctx, uint8_t handle_out[SRB_IPC_HANDLE_BYTES]);
int srb_gather_rgba8_target(SrbContext* ctx, const uint8_t handle[SRB_IPC_HANDLE_BYTES]);
int srb_read_gathered_rgba8(SrbContext* ctx, uint8_t* rgba8_out);
void* srb_device_gathered_rgba8_ptr(SrbContext* ctx);
/* Queue the copy of the assembled frame to (pinned) host memory; same slots and rules as srb_present_rgba8_async. */
int srb_present_gathered_rgba8_async(SrbContext* ctx, uint8_t* rgba8_out);
/* Work per 16-px tile row of the last completed frame: (triangle, tile) pairs binned into the
 * row (out[0..n), n <= tile rows).  Input of cost-balanced band cuts. */
int srb_get_tile_row_pairs(SrbContext* ctx, uint32_t* out, int n);

/* ---- resources: Texture (Texture.cs:31-63), Mesh arrays (ModelLoader.cs:42-57) ---- */
int srb_texture_create(SrbContext* ctx, const uint8_t* rgba8, int width, int height, SrbTexture* out);
int srb_texture_destroy(SrbContext* ctx, SrbTexture tex);
/* Texture.Sample (Texture.cs:42-63) for n uv pairs, evaluated on the device; filter as SrbUniforms.tex_filter. */
int srb_texture_sample(SrbContext* ctx, SrbTexture tex, const float* uv /* n*2 */, int n, int filter,
                       float* rgba_out /* n*4 */);
int srb_mesh_create(SrbContext* ctx, const float* vertices, int n_vertices,
                    const uint16_t* indices, int n_indices, SrbMesh* out);
int srb_mesh_destroy(SrbContext* ctx, SrbMesh mesh);

/* ---- frame: Renderer.RenderScene call pattern (Renderer.cs:404-419) ---- */
int srb_set_uniforms(SrbContext* ctx, const SrbUniforms* u);
int srb_begin_frame(SrbContext* ctx, const float clear_rgba[4], uint32_t clear_flags);
/* Rasterizer.RenderMesh (Rasterizer.cs:163-230), stock shader pair only. */
int srb_draw(SrbContext* ctx, SrbMesh mesh, const float model[16], const float view[16],
             const float proj[16], SrbTexture tex, int cull, int depth_test, int blend,
             int submission_index);
int srb_submit_draws(SrbContext* ctx, const SrbDrawCmd* cmds, int n);
/* MainWindow.RenderFrame (MainWindow.cs:209-212): runs the whole pipeline, asynchronously. */
int srb_end_frame(SrbContext* ctx);
/* Re-run the pipeline on the draw list already resident on the device (benchmark loop). */
int srb_replay_frame(SrbContext* ctx);
int srb_sync(SrbContext* ctx);

/* ---- readback: MainWindow.ColorBuffer/DepthBuffer + accessors (MainWindow.cs:378-436) ---- */
int srb_read_color(SrbContext* ctx, float* rgba_out);            /* W*H*4 floats */
int srb_read_depth(SrbContext* ctx, float* depth_out);           /* W*H floats */
int srb_read_color_rgba8(SrbContext* ctx, uint8_t* rgba8_out);   /* 8-bit convenience copy (the parity metric's format) */
/* Present path: MainWindow.OnRender flattens ColorBuffer to packed Vector3 and uploads it with
 * TexSubImage2D(Rgb, Float) (MainWindow.cs:226-246).  srb_present_rgb delivers exactly that array
 * (W*H*3 floats) to the host.  The _async form returns once the copy is queued on a second stream, so
 * the next frame can be submitted and rendered while it runs; two copies may be outstanding (two
 * different, preferably pinned, host buffers); srb_present_wait returns when they have landed. */
int srb_present_rgb(SrbContext* ctx, float* rgb_out);
int srb_present_rgb_async(SrbContext* ctx, float* rgb_out);
/* Rows [y0,y1) only, written at their place in the full-frame array rgb_out points to.  Multi-GPU: every band
 * rank delivers its own rows to one shared (pinned) host frame over its own PCIe link, in parallel. */
int srb_present_rgb_rows_async(SrbContext* ctx, float* rgb_out, int y0, int y1);
/* The presented frame at display precision: the reference draws its float texture into
 * the window's 8-bit framebuffer (MainWindow.cs:251-253), so what reaches the screen is clamp(c,0,1)*255
 * rounded to nearest -- W*H*4 bytes R,G,B,A, a quarter of the float4 buffer.  Same queueing rules as the
 * float forms (they share the two present slots); the host side uploads it as (Rgba, UnsignedByte). */
int srb_present_rgba8(SrbContext* ctx, uint8_t* rgba8_out);
int srb_present_rgba8_async(SrbContext* ctx, uint8_t* rgba8_out);
int srb_present_rgba8_rows_async(SrbContext* ctx, uint8_t* rgba8_out, int y0, int y1);
/* Device-side fence: work queued on the pipeline stream (srb_stream) after this call starts only when the
 * present copies queued so far have landed.  A collective placed behind it is a "frame delivered" barrier
 * across band ranks that never blocks the host. */
int srb_present_fence(SrbContext* ctx);
/* Frame barrier between the band contexts of one frame, ONE CONTEXT PER GPU (no reference counterpart; SURVEY.md
 * 8(e)).  Every rank exports its row of flags, the rows are exchanged (any transport), every rank connects with all
 * of them (handles: world x 64 bytes, indexed by rank; its own entry is ignored).  srb_frame_barrier queues a
 * 32-thread kernel behind the frame on the context's stream: it posts "this rank has finished frame k" into every
 * rank's row over NVLink and waits until all ranks have posted frame k - slack (slack 0: every band of this frame
 * has landed, the kernel sits on the raster stream; slack >= 1: ranks only stay close -- the kernel runs on a side
 * stream and the raster stream waits for the previous call's barrier, i.e. ranks stay within slack + 1 frames).
 * The wait is bounded (4 s), a time-out turns into an error of the next call.  handles == NULL disconnects. */
int srb_barrier_export(SrbContext* ctx, uint8_t handle_out[SRB_IPC_HANDLE_BYTES]);
int srb_barrier_connect(SrbContext* ctx, int rank, int world, const uint8_t* handles);
int srb_frame_barrier(SrbContext* ctx, int slack);
int srb_present_wait(SrbContext* ctx);
int srb_read_color_rows(SrbContext* ctx, float* rgba_out, int y0, int y1);
int srb_get_pixel(SrbContext* ctx, int x, int y, float rgba_out[4]);   /* :390-398 */
int srb_set_pixel(SrbContext* ctx, int x, int y, const float rgba[4]); /* :382-388 */
int srb_get_depth(SrbContext* ctx, int x, int y, float* depth_out);    /* :419-427 */
int srb_set_depth(SrbContext* ctx, int x, int y, float depth);         /* :410-416 */
void* srb_device_color_ptr(SrbContext* ctx);   /* float4[W*H] on the device (headless present) */
void* srb_device_depth_ptr(SrbContext* ctx);   /* float[W*H] on the device */
void* srb_stream(SrbContext* ctx);             /* cudaStream_t the pipeline runs on */

/* ---- FrustumCuller.IsSphereInFrustum (FrustumCuller.cs:201-218), batched on the device ---- */
int srb_frustum_test(SrbContext* ctx, const float* spheres /* n*4 */, const float* models /* n*16 */,
                     const float view[16], const float proj[16], int n, int32_t* out_visible);

/* ---- Physics.Raycast (Physics.cs:19-179), batched on the device ----
 * n rays (6 floats each: origin xyz, direction xyz -- normalised by the call, Physics.cs:70) against one
 * resident mesh under `model`; closest hit per ray by the reference's Moller-Trumbore test (:137-178).
 * face_mask: RaycastFaceMask bits (Physics.cs:9-14).  A non-invertible model gives no hits (:30-36).
 * Where the reference's parallel reduction leaves ties in distance to thread timing, the lowest triangle
 * index wins here. */
#define SRB_RAY_IGNORE_BACKFACES  1
#define SRB_RAY_IGNORE_FRONTFACES 2
typedef struct SrbRayHit {
    int32_t hit;            /* return value of Physics.Raycast */
    float   distance;       /* hitDistance (float.MaxValue when no hit) */
    float   point[3];       /* hitPoint */
    float   normal[3];      /* hitNormal */
    int32_t triangle;       /* winning triangle index, -1 when no hit (not in the reference's signature) */
    int32_t _pad[3];
} SrbRayHit;
int srb_raycast(SrbContext* ctx, SrbMesh mesh, const float model[16], const float* rays /* n*6 */, int n,
                int face_mask, SrbRayHit* out /* n */);

/* srb_raycast for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same arguments,
 * same results bit for bit, but the call performs no cudaMalloc / cudaFree / allocating library routine: its scratch
 * (96 bytes per ray) is one fixed device array inside the library, 384 KiB, held by every GPU that loads the library
 * until the process exits.  n > SRB_RAYCAST_STATIC_MAX_RAYS is refused (SRB_ERR_ARG).  All calls on a device share
 * the array and it keeps the last call's contents; the call waits for its own result, but calls that overlap on the
 * GPU (different contexts of one device, from different threads) race -- serialise them. */
#define SRB_RAYCAST_STATIC_MAX_RAYS 4096
int srb_raycast_static(SrbContext* ctx, SrbMesh mesh, const float model[16], const float* rays /* n x 6 */, int n,
                       int face_mask, SrbRayHit* out /* n */);

int srb_get_stats(SrbContext* ctx, SrbStats* out);
/* Per-stage device times of the last completed frame (CUDA events on the pipeline stream).
 * Enable with srb_set_profiling(ctx, 1); stages: 0 frustum, 1 vertex, 2 setup+count, 3 scan,
 * 4 fill, 5 sort (ordered path only), 6 tile raster.  Returns ms per stage in out[0..n). */
#define SRB_N_STAGES 7
int srb_set_profiling(SrbContext* ctx, int enabled);
int srb_get_stage_ms(SrbContext* ctx, float* out, int n);
int srb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SRB200_H */
