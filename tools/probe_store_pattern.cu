// Micro-benchmark: how fast can the 8 epilogue warps of one CTA per SM write a 128-pixel x 128-channel two-plane tile
// (64 KB) to global memory, as a function of the store pattern?
//   0: one pixel per lane, 32 B per store (what the conv epilogue does: 32 distinct 128-byte lines per instruction)
//   1: 8 lanes per pixel, 16 B per lane (full 128-byte lines: 4 lines per instruction)
//   2: 4 lanes per pixel, 16 B per lane (half lines, the 64-channel-tile case: 8 lines per instruction)
//   3: 4 lanes per pixel, 32 B per lane (full lines with 256-bit stores: 8 lines per instruction)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o probe_store_pattern probe_store_pattern.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void stg256(void* p, uint32_t a) {
    asm volatile("st.global.v8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" :: "l"(p), "r"(a) : "memory");
}
__device__ __forceinline__ void stg128(void* p, uint32_t a) {
    asm volatile("st.global.v4.b32 [%0], {%1,%1,%1,%1};" :: "l"(p), "r"(a) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(uint8_t* out, int tiles_per_cta, long long* cycles) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, hf = warp >> 2;              // lane quarter (32 pixels), column half (64 channels = 128 B)
    const long long t0 = clock64();
    for (int t = 0; t < tiles_per_cta; ++t) {
        const size_t tile = static_cast<size_t>(t) * gridDim.x + blockIdx.x;
        for (int plane = 0; plane < 2; ++plane) {
            uint8_t* base = out + (static_cast<size_t>(plane) * gridDim.x * tiles_per_cta + tile) * 128 * 256 + q * 32 * 256 + hf * 128;
            if (MODE == 0) {
#pragma unroll
                for (int c = 0; c < 4; ++c) stg256(base + lane * 256 + c * 32, t + c);
            } else if (MODE == 1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) stg128(base + (4 * j + (lane >> 3)) * 256 + (lane & 7) * 16, t + j);
            } else if (MODE == 2) {
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int j = 0; j < 4; ++j) stg128(base + (8 * j + (lane >> 2)) * 256 + h * 64 + (lane & 3) * 16, t + j);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) stg256(base + (8 * j + (lane >> 2)) * 256 + (lane & 3) * 32, t + j);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
    const int ctas = 148, tiles = 64;
    uint8_t* out; long long* cyc;
    cudaMalloc(&out, static_cast<size_t>(2) * ctas * tiles * 128 * 256);
    cudaMalloc(&cyc, ctas * sizeof(long long));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 4; ++mode) {
        float best = 1e9f; long long c0 = 0;
        for (int rep = 0; rep < 5; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<ctas, 256>>>(out, tiles, cyc);
            if (mode == 1) k<1><<<ctas, 256>>>(out, tiles, cyc);
            if (mode == 2) k<2><<<ctas, 256>>>(out, tiles, cyc);
            if (mode == 3) k<3><<<ctas, 256>>>(out, tiles, cyc);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) { best = ms; cudaMemcpy(&c0, cyc, sizeof(c0), cudaMemcpyDeviceToHost); }
        }
        const double bytes = 2.0 * ctas * tiles * 128 * 256;
        printf("mode %d: %.1f us, %.0f GB/s, %.0f cycles per 64 KB tile (CTA 0)\n", mode, best * 1e3, bytes / best * 1e-6, double(c0) / tiles);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
