// The per-draw entry point under load: what the C# drop-in does every frame (Renderer.cs:444-459) -- RenderMesh, i.e.
// srb_draw, once per mesh from racing Parallel.ForEach threads, with explicit submission indices.  1,332 draws (the
// draw count of BASELINE config 3) from T host threads per frame, then srb_end_frame.  Prints the host cost per call
// and the frame rate of the whole loop (device included).  Build: make -C tools/micro   Run: tools/micro/per_draw_submit [threads]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#include "../../include/srb200.h"

int main(int argc, char** argv) {
    const int T = argc > 1 ? std::atoi(argv[1]) : 8, nDraws = 1332, frames = 200;
    SrbContext* ctx = nullptr;
    if (srb_create(0, 1920, 1080, &ctx) != SRB_OK) { std::fprintf(stderr, "srb_create: %s\n", srb_last_error(nullptr)); return 1; }
    // one small mesh (a 12 x 12 grid, 242 triangles), drawn at 1,332 places
    const int g = 12;
    std::vector<float> v;
    std::vector<uint16_t> idx;
    for (int y = 0; y < g; ++y)
        for (int x = 0; x < g; ++x) {
            const float p[12] = {x * 0.1f, y * 0.1f, 0.0f, x / float(g), y / float(g), 0, 0, 1, 0.8f, 0.7f, 0.6f, 1.0f};
            v.insert(v.end(), p, p + 12);
        }
    for (int y = 0; y + 1 < g; ++y)
        for (int x = 0; x + 1 < g; ++x) {
            const uint16_t q = (uint16_t)(y * g + x);
            const uint16_t t[6] = {q, (uint16_t)(q + 1), (uint16_t)(q + g + 1), q, (uint16_t)(q + g + 1), (uint16_t)(q + g)};
            idx.insert(idx.end(), t, t + 6);
        }
    SrbMesh mesh = 0;
    if (srb_mesh_create(ctx, v.data(), g * g, idx.data(), (int)idx.size(), &mesh) != SRB_OK) { std::fprintf(stderr, "%s\n", srb_last_error(ctx)); return 1; }
    std::vector<float> models((size_t)nDraws * 16, 0.0f);
    for (int i = 0; i < nDraws; ++i) {
        float* m = &models[(size_t)i * 16];
        m[0] = m[5] = m[10] = m[15] = 1.0f;
        m[12] = -18.0f + 1.0f * (i % 37); m[13] = -9.0f + 0.5f * (i / 37); m[14] = -20.0f;
    }
    const float view[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    const float f = 1.0f, a = 1920.0f / 1080.0f, zn = 0.1f, zf = 1000.0f;     // CreatePerspectiveFieldOfView(90 deg)
    const float proj[16] = {f / a, 0, 0, 0, 0, f, 0, 0, 0, 0, zf / (zn - zf), -1, 0, 0, zn * zf / (zn - zf), 0};
    const float clear[4] = {0.9137f, 0.7098f, 0.6588f, 1.0f};
    double callNs = 0.0;
    std::chrono::steady_clock::time_point t0;
    for (int fr = -5; fr < frames; ++fr) {
        if (fr == 0) { srb_sync(ctx); t0 = std::chrono::steady_clock::now(); callNs = 0.0; }
        srb_begin_frame(ctx, clear, SRB_CLEAR_COLOR | SRB_CLEAR_DEPTH);
        const auto c0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th;
        for (int t = 0; t < T; ++t)
            th.emplace_back([&, t] {
                for (int i = t; i < nDraws; i += T)
                    srb_draw(ctx, mesh, &models[(size_t)i * 16], view, proj, 0, SRB_CULL_BACK, SRB_DEPTH_LESS_EQUAL, SRB_BLEND_ALPHA, i);
            });
        for (auto& x : th) x.join();
        callNs += std::chrono::duration<double, std::nano>(std::chrono::steady_clock::now() - c0).count();
        if (srb_end_frame(ctx) != SRB_OK) { std::fprintf(stderr, "end_frame: %s\n", srb_last_error(ctx)); return 1; }
    }
    srb_sync(ctx);
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    SrbStats st;
    srb_get_stats(ctx, &st);
    std::printf("{\"per_draw_submit\": {\"threads\": %d, \"draws_per_frame\": %d, \"frames\": %d, \"frames_per_sec\": %.1f, "
                "\"submit_wall_us_per_frame\": %.1f, \"ns_per_srb_draw_call\": %.1f, \"triangles_setup\": %lld}}\n",
                T, nDraws, frames, frames / s, callNs / frames / 1e3, callNs / frames / nDraws, (long long)st.triangles_setup);
    srb_destroy(ctx);
    return 0;
}
