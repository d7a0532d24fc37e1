// Native.cs -- P/Invoke binding of libsrb200.so (include/srb200.h) for OCSYT/SoftwareRenderer.
// Drop this file into the reference project (AllowUnsafeBlocks is already on, csproj:8) and apply
// the call-site edits listed in INTEGRATION.md.  It cannot be compiled in the build image (no .NET
// SDK); it is written against the C ABI and kept in sync by tests/test_abi.py, which checks every
// [DllImport] below against the header.
using System;
using System.Numerics;
using System.Runtime.InteropServices;

namespace SoftwareRenderer
{
    [StructLayout(LayoutKind.Sequential)]
    public unsafe struct SrbUniforms            // srb200.h: SrbUniforms (176 bytes)
    {
        public Vector4 LightDir;                // Renderer.LightDirection (xyz)
        public Vector4 LightColor;              // Renderer.LightColor
        public Vector4 FogColor;                // Renderer.FogColor
        public float FogStart, FogEnd;          // Renderer.FogStart / FogEnd
        public float NearClip;                  // Rasterizer.NearClip
        public int NumericsFma;                 // 1 on FMA hardware under net9.0
        public int TexFilter;                   // 0 = Texture.Sample (nearest, repeat)
        public int NExtraLights;
        public int DebugMode;                   // (int)Rasterizer.RenderDebugMode: 0 None, 1 Wireframe
        public int Pad;
        public fixed float ExtraDir[16];
        public fixed float ExtraColor[16];
        public fixed float ExtraAtten[16];
    }

    [StructLayout(LayoutKind.Sequential)]
    public unsafe struct SrbRayHit              // srb200.h: SrbRayHit (48 bytes)
    {
        public int Hit;                         // return value of Physics.Raycast
        public float Distance;                  // hitDistance
        public Vector3 Point;                   // hitPoint
        public Vector3 Normal;                  // hitNormal
        public int Triangle;
        public fixed int Pad[3];
    }

    [StructLayout(LayoutKind.Sequential)]
    public struct SrbStats
    {
        public long DrawsSubmitted, DrawsVisible, TrianglesSubmitted, TrianglesSetup, TilePairs, ClippedTriangles;
        public int PipelineUsed, KernelLaunches;
        public float GpuMs;
        public int TrianglesNoCoverage, SplitParts, Pad;
    }

    public static unsafe class Srb
    {
        const string Lib = "srb200";
        public const uint ClearColor = 1, ClearDepth = 2;

        [DllImport(Lib)] public static extern int srb_version();
        [DllImport(Lib)] public static extern int srb_create(int device, int width, int height, out IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_destroy(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_resize(IntPtr ctx, int width, int height);
        [DllImport(Lib)] public static extern int srb_get_size(IntPtr ctx, out int width, out int height);
        [DllImport(Lib)] public static extern IntPtr srb_last_error(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_set_band(IntPtr ctx, int tileRow0, int tileRow1);
        [DllImport(Lib)] public static extern int srb_barrier_export(IntPtr ctx, byte* handleOut);
        [DllImport(Lib)] public static extern int srb_barrier_connect(IntPtr ctx, int rank, int world, byte* handles);
        [DllImport(Lib)] public static extern int srb_frame_barrier(IntPtr ctx, int slack);
        [DllImport(Lib)] public static extern int srb_set_bin_split(IntPtr ctx, int minEntriesFullFrame, int minEntriesBand, int partEntries);
        [DllImport(Lib)] public static extern int srb_set_pipeline_depth(IntPtr ctx, int bufferSets);
        [DllImport(Lib)] public static extern int srb_set_graphs(IntPtr ctx, int enabled);
        [DllImport(Lib)] public static extern int srb_set_pipeline(IntPtr ctx, int pipeline);
        [DllImport(Lib)] public static extern int srb_export_color_handle(IntPtr ctx, byte* handleOut64);
        [DllImport(Lib)] public static extern int srb_set_color_target(IntPtr ctx, byte* handle64);
        [DllImport(Lib)] public static extern int srb_get_tile_row_pairs(IntPtr ctx, uint* rowsOut, int n);
        [DllImport(Lib)] public static extern int srb_texture_create(IntPtr ctx, byte* rgba8, int w, int h, out int tex);
        [DllImport(Lib)] public static extern int srb_texture_destroy(IntPtr ctx, int tex);
        [DllImport(Lib)] public static extern int srb_texture_sample(IntPtr ctx, int tex, float* uv, int n, int filter, float* rgbaOut);
        [DllImport(Lib)] public static extern int srb_mesh_create(IntPtr ctx, float* vertices, int nVertices, ushort* indices, int nIndices, out int mesh);
        [DllImport(Lib)] public static extern int srb_mesh_destroy(IntPtr ctx, int mesh);
        [DllImport(Lib)] public static extern int srb_set_uniforms(IntPtr ctx, in SrbUniforms u);
        [DllImport(Lib)] public static extern int srb_begin_frame(IntPtr ctx, float* clearRgba, uint clearFlags);
        [DllImport(Lib)] public static extern int srb_draw(IntPtr ctx, int mesh, float* model, float* view, float* proj, int tex, int cull, int depthTest, int blend, int submissionIndex);
        [DllImport(Lib)] public static extern int srb_submit_draws(IntPtr ctx, void* cmds, int n);
        [DllImport(Lib)] public static extern int srb_end_frame(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_replay_frame(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_sync(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_read_color(IntPtr ctx, float* rgbaOut);
        [DllImport(Lib)] public static extern int srb_read_depth(IntPtr ctx, float* depthOut);
        [DllImport(Lib)] public static extern int srb_read_color_rgba8(IntPtr ctx, byte* rgba8Out);
        [DllImport(Lib)] public static extern int srb_present_rgb(IntPtr ctx, float* rgbOut);
        [DllImport(Lib)] public static extern int srb_present_rgb_async(IntPtr ctx, float* rgbOut);
        [DllImport(Lib)] public static extern int srb_present_rgb_rows_async(IntPtr ctx, float* rgbOut, int y0, int y1);
        [DllImport(Lib)] public static extern int srb_present_rgba8(IntPtr ctx, byte* rgba8Out);
        [DllImport(Lib)] public static extern int srb_present_rgba8_async(IntPtr ctx, byte* rgba8Out);
        [DllImport(Lib)] public static extern int srb_present_rgba8_rows_async(IntPtr ctx, byte* rgba8Out, int y0, int y1);
        [DllImport(Lib)] public static extern int srb_raycast(IntPtr ctx, int mesh, float* model, float* rays, int n, int faceMask, SrbRayHit* hits);
        [DllImport(Lib)] public static extern int srb_raycast_static(IntPtr ctx, int mesh, float* model, float* rays, int n, int faceMask, SrbRayHit* hits);
        [DllImport(Lib)] public static extern int srb_gather_rgba8_export(IntPtr ctx, byte* handleOut);
        [DllImport(Lib)] public static extern int srb_gather_rgba8_target(IntPtr ctx, byte* handle);
        [DllImport(Lib)] public static extern int srb_read_gathered_rgba8(IntPtr ctx, byte* rgba8Out);
        [DllImport(Lib)] public static extern IntPtr srb_device_gathered_rgba8_ptr(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_present_gathered_rgba8_async(IntPtr ctx, byte* rgba8Out);
        [DllImport(Lib)] public static extern int srb_present_wait(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_present_fence(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_read_color_rows(IntPtr ctx, float* rgbaOut, int y0, int y1);
        [DllImport(Lib)] public static extern int srb_get_pixel(IntPtr ctx, int x, int y, float* rgbaOut);
        [DllImport(Lib)] public static extern int srb_set_pixel(IntPtr ctx, int x, int y, float* rgba);
        [DllImport(Lib)] public static extern int srb_get_depth(IntPtr ctx, int x, int y, out float depth);
        [DllImport(Lib)] public static extern int srb_set_depth(IntPtr ctx, int x, int y, float depth);
        [DllImport(Lib)] public static extern IntPtr srb_device_color_ptr(IntPtr ctx);
        [DllImport(Lib)] public static extern IntPtr srb_device_depth_ptr(IntPtr ctx);
        [DllImport(Lib)] public static extern IntPtr srb_stream(IntPtr ctx);
        [DllImport(Lib)] public static extern int srb_frustum_test(IntPtr ctx, float* spheres, float* models, float* view, float* proj, int n, int* outVisible);
        [DllImport(Lib)] public static extern int srb_get_stats(IntPtr ctx, out SrbStats stats);
        [DllImport(Lib)] public static extern int srb_set_profiling(IntPtr ctx, int enabled);
        [DllImport(Lib)] public static extern int srb_get_stage_ms(IntPtr ctx, float* outMs, int n);

        public static void Check(IntPtr ctx, int rc)
        {
            if (rc == 0) return;
            string msg = Marshal.PtrToStringUTF8(srb_last_error(ctx)) ?? "";
            if (rc == -1) throw new ArgumentException(msg);       // Rasterizer.cs:71-74 convention
            throw new InvalidOperationException($"srb200 error {rc}: {msg}");
        }
    }
}
