// Hardware probe: can a K-major SWIZZLE_128B UMMA operand START at a row that is not a multiple of 8 rows (1024 B), with
// a stride between 8-row groups (SBO) that is not a multiple of 1024 B?  That is what a convolution needs to read all
// 9 taps of a 3x3 filter out of ONE halo box (18 rows x 10 pixels x 64 channels) written by a single TMA load:
// tile = 16 rows x 8 pixels, M row m = h*8 + w lives at box row (h + r)*10 + (w + s), i.e. start = (r*10 + s) rows,
// SBO = 10 rows = 1280 B.  TMA (and this probe's loader) swizzles by ABSOLUTE shared-memory address bits
// (16-byte chunk index ^= address bits [7,10)); the question is whether tcgen05.mma does the same on the read side.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/bin/probe_swizzle_offset tools/probe_swizzle_offset.cu
#include <cuda_bf16.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include "../styletransfer_b200/csrc/ptx.cuh"

constexpr int M = 128, N = 64, K = 64, BOX_W = 10, BOX_H = 18, BOX_ROWS = BOX_W * BOX_H;

__device__ __forceinline__ uint32_t idesc() {
    return (1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24);
}

__global__ void __launch_bounds__(128) probe(const uint16_t* box, const uint16_t* B, int r, int s, int base_offset, float* D) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                                  // BOX_ROWS rows x 128 B
    uint8_t* sB = smem + ((BOX_ROWS * 128 + 1023) & ~1023);
    uint64_t* bar = reinterpret_cast<uint64_t*>(sB + N * 128);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int t = threadIdx.x, warp = t >> 5;
    for (int i = t; i < BOX_ROWS * 8; i += 128) {
        const int row = i >> 3, c = i & 7;
        *reinterpret_cast<uint4*>(sA + row * 128 + ((c ^ (row & 7)) << 4)) = *reinterpret_cast<const uint4*>(box + row * K + c * 8);
    }
    for (int i = t; i < N * 8; i += 128) {
        const int row = i >> 3, c = i & 7;
        *reinterpret_cast<uint4*>(sB + row * 128 + ((c ^ (row & 7)) << 4)) = *reinterpret_cast<const uint4*>(B + row * K + c * 8);
    }
    if (t == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) { ptx::tmem_alloc(slot, 64); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *slot;
    if (t == 0) {
        const uint32_t a0 = ptx::umma_desc_lo(ptx::smem_u32(sA) + (r * BOX_W + s) * 128);
        const uint32_t b0 = ptx::umma_desc_lo(ptx::smem_u32(sB));
        // hi word: SBO >> 4 | base offset (bits 49..51 of the descriptor = 17..19 of this word) | version 1 | SWIZZLE_128B
        const uint32_t a_hi = uint32_t((BOX_W * 128) >> 4) | (uint32_t(base_offset & 7) << 17) | (1u << 14) | (2u << 29);
        for (int k = 0; k < K / 16; ++k) {
            // two operands need two "hi" words: issue with the A word for A and the standard one for B
            uint64_t da = (uint64_t(a_hi) << 32) | (a0 + ((k * 32) >> 4));
            uint64_t db = (uint64_t(ptx::umma_desc_hi_k_sw128()) << 32) | (b0 + ((k * 32) >> 4));
            uint32_t acc = k > 0;
            asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                         :: "r"(tmem), "l"(da), "l"(db), "r"(idesc()), "r"(acc) : "memory");
        }
        ptx::tc_commit(bar);
    }
    ptx::mbar_wait(bar, 0);
    ptx::tc_fence_after();
    for (int c = 0; c < N; c += 16) {
        uint32_t v[16];
        ptx::tmem_ld16(tmem + (uint32_t(warp * 32) << 16) + c, v);
        ptx::tmem_ld_wait();
        for (int i = 0; i < 16; ++i) D[t * N + c + i] = __uint_as_float(v[i]);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tmem, 64);
}

static uint16_t enc(float v, float* back) {
    __nv_bfloat16 b = __float2bfloat16_rn(v); *back = __bfloat162float(b); return *reinterpret_cast<uint16_t*>(&b);
}

int main() {
    uint16_t *hA = new uint16_t[BOX_ROWS * K], *hB = new uint16_t[N * K], *dA, *dB;
    float *fA = new float[BOX_ROWS * K], *fB = new float[N * K], *hD = new float[M * N], *dD;
    cudaMalloc(&dA, BOX_ROWS * K * 2); cudaMalloc(&dB, N * K * 2); cudaMalloc(&dD, M * N * 4);
    const int smem = 1024 + ((BOX_ROWS * 128 + 1023) & ~1023) + N * 128 + 64;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    srand(11);
    for (int i = 0; i < BOX_ROWS * K; ++i) hA[i] = enc((rand() % 2001 - 1000) / 517.0f, &fA[i]);
    for (int i = 0; i < N * K; ++i) hB[i] = enc((rand() % 2001 - 1000) / 733.0f, &fB[i]);
    cudaMemcpy(dA, hA, BOX_ROWS * K * 2, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, N * K * 2, cudaMemcpyHostToDevice);
    int ok_plain = 0, ok_bo = 0, total = 0;
    for (int mode = 0; mode < 2; ++mode)            // 0: base_offset field = 0, 1: base_offset = start row & 7
        for (int r = 0; r < 3; ++r)
            for (int s = 0; s < 3; ++s) {
                const int start = r * BOX_W + s;
                cudaMemset(dD, 0, M * N * 4);
                probe<<<1, 128, smem>>>(dA, dB, r, s, mode ? (start & 7) : 0, dD);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("tap r=%d s=%d mode=%d: CUDA error %s\n", r, s, mode, cudaGetErrorString(e)); return 2; }
                cudaMemcpy(hD, dD, M * N * 4, cudaMemcpyDeviceToHost);
                double worst = 0, ref_max = 0;
                for (int m = 0; m < M; ++m)
                    for (int n = 0; n < N; ++n) {
                        const int row = (m / 8 + r) * BOX_W + (m % 8) + s;
                        double acc = 0;
                        for (int k = 0; k < K; ++k) acc += double(fA[row * K + k]) * double(fB[n * K + k]);
                        worst = fmax(worst, fabs(acc - hD[m * N + n])); ref_max = fmax(ref_max, fabs(acc));
                    }
                const bool ok = worst < 1e-4 * ref_max;
                printf("tap r=%d s=%d start row %2d (phase %d) base_offset=%d: max abs err %.3e (max |ref| %.3f) -> %s\n",
                       r, s, start, start & 7, mode ? (start & 7) : 0, worst, ref_max, ok ? "OK" : "MISMATCH");
                if (mode == 0) ok_plain += ok; else ok_bo += ok;
                ++total;
            }
    printf("RESULT: base_offset=0: %d/9 taps exact; base_offset=phase: %d/9 taps exact\n", ok_plain, ok_bo);
    return 0;
}
