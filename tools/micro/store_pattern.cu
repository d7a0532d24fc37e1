// Micro-benchmark: how fast can a 3840x2160 float4 colour + float depth frame be written with streaming stores,
// depending on the shape of the region one CTA writes?  (Why the tile kernel's clear-only path takes 41 us.)
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o store_pattern store_pattern.cu && ./store_pattern
#include <cstdio>
#include <cuda_runtime.h>

template <int TW, int TH>
__global__ void __launch_bounds__(128) tile_store(float4* color, float* depth, int W, int H, int tilesX, const uint4* order) {
    int tile = order ? (int)__ldg(order + blockIdx.x).x : blockIdx.x;
    int tx0 = (tile % tilesX) * TW, ty0 = (tile / tilesX) * TH;
    const float4 c = make_float4(0.9f, 0.7f, 0.65f, 1.0f);
    for (int p = threadIdx.x; p < TW * TH; p += 128) {
        int x = tx0 + p % TW, y = ty0 + p / TW;
        if (x < W && y < H) {
            size_t i = (size_t)y * W + x;
            __stcs(color + i, c);
            __stcs(depth + i, -3.4e38f);
        }
    }
}
__global__ void linear_store(float4* color, float* depth, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { __stcs(color + i, make_float4(0.9f, 0.7f, 0.65f, 1.0f)); __stcs(depth + i, -3.4e38f); }
}

template <typename F> float timeit(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 5; ++i) f();
    cudaEventRecord(a);
    for (int i = 0; i < 50; ++i) f();
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms / 50 * 1000;
}

int main() {
    const int W = 3840, H = 2160;
    float4* color; float* depth; uint4* order;
    cudaMalloc(&color, (size_t)W * H * 16); cudaMalloc(&depth, (size_t)W * H * 4);
    const int n16 = (W / 16) * (H / 16);
    uint4* h = new uint4[n16];
    for (int i = 0; i < n16; ++i) h[i] = make_uint4(i, 0, 0, 0);
    cudaMalloc(&order, n16 * 16); cudaMemcpy(order, h, n16 * 16, cudaMemcpyHostToDevice);
    const double mb = (double)W * H * 20 / 1e6;
    auto rep = [&](const char* name, float us) { printf("%-44s %7.1f us  %6.2f TB/s\n", name, us, mb / us); };
    rep("linear, 256 threads", timeit([&] { linear_store<<<(W * H + 255) / 256, 256>>>(color, depth, (size_t)W * H); }));
    rep("16x16 per CTA (order load first)", timeit([&] { tile_store<16, 16><<<n16, 128>>>(color, depth, W, H, W / 16, order); }));
    rep("16x16 per CTA (no order load)", timeit([&] { tile_store<16, 16><<<n16, 128>>>(color, depth, W, H, W / 16, nullptr); }));
    rep("32x16 per CTA", timeit([&] { tile_store<32, 16><<<n16 / 2, 128>>>(color, depth, W, H, W / 32, nullptr); }));
    rep("64x16 per CTA", timeit([&] { tile_store<64, 16><<<n16 / 4, 128>>>(color, depth, W, H, W / 64, nullptr); }));
    rep("64x4 per CTA", timeit([&] { tile_store<64, 4><<<n16, 128>>>(color, depth, W, H, W / 64, nullptr); }));
    rep("256x1 per CTA", timeit([&] { tile_store<256, 1><<<n16, 128>>>(color, depth, W, H, W / 256, nullptr); }));
    rep("16x16 per CTA, 2 memsets instead", timeit([&] { cudaMemsetAsync(color, 0, (size_t)W * H * 16); cudaMemsetAsync(depth, 0, (size_t)W * H * 4); }));
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
