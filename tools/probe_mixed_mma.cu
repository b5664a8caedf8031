// Hardware probe: does tcgen05.mma kind::f16 accept DIFFERENT 16-bit formats for A and B (a_format != b_format in
// the instruction descriptor: fp16 x bf16)?  The f16x3 precision mode multiplies forward activations (fp16 planes)
// with gradients / per-image Gram-backward weights (bf16 planes) in the backward-weight and Gram-backward kernels.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o gpurun_out/probe_mixed_mma tools/probe_mixed_mma.cu && ./gpurun_out/probe_mixed_mma
// One CTA, M = 128, N = 64, K = 64, operands written to shared memory by hand in the K-major 128-byte-swizzle layout
// (what a TMA box with a 64-element inner dimension produces).  Prints the max error per format combination.
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include "../styletransfer_b200/csrc/ptx.cuh"

constexpr int M = 128, N = 64, K = 64;

__device__ __forceinline__ uint32_t idesc(int afmt, int bfmt) {     // 0 = f16, 1 = bf16
    return (1u << 4) | (uint32_t(afmt) << 7) | (uint32_t(bfmt) << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24);
}

__global__ void __launch_bounds__(128) probe(const uint16_t* A, const uint16_t* B, int afmt, int bfmt, float* D) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                 // 128 rows x 128 B
    uint8_t* sB = smem + M * 128;       // 64 rows x 128 B
    uint64_t* bar = reinterpret_cast<uint64_t*>(sB + N * 128);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int t = threadIdx.x, warp = t >> 5;
    for (int i = t; i < M * 8; i += 128) {           // 16-byte chunks
        const int r = i >> 3, c = i & 7;
        *reinterpret_cast<uint4*>(sA + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(A + r * K + c * 8);
    }
    for (int i = t; i < N * 8; i += 128) {
        const int r = i >> 3, c = i & 7;
        *reinterpret_cast<uint4*>(sB + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(B + r * K + c * 8);
    }
    if (t == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) { ptx::tmem_alloc(slot, 64); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *slot;
    if (t == 0) {
        const uint32_t a0 = ptx::umma_desc_lo(ptx::smem_u32(sA)), b0 = ptx::umma_desc_lo(ptx::smem_u32(sB));
        for (int k = 0; k < K / 16; ++k)
            ptx::mma_bf16_ss2(tmem, a0 + ((k * 32) >> 4), b0 + ((k * 32) >> 4), ptx::umma_desc_hi_k_sw128(), idesc(afmt, bfmt), k > 0);
        ptx::tc_commit(bar);
    }
    ptx::mbar_wait(bar, 0);
    ptx::tc_fence_after();
    for (int c = 0; c < N; c += 16) {
        uint32_t r[16];
        ptx::tmem_ld16(tmem + (uint32_t(warp * 32) << 16) + c, r);
        ptx::tmem_ld_wait();
        for (int i = 0; i < 16; ++i) D[t * N + c + i] = __uint_as_float(r[i]);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tmem, 64);
}

static uint16_t enc(float v, int fmt, float* back) {
    if (fmt == 0) { __half h = __float2half_rn(v); *back = __half2float(h); return *reinterpret_cast<uint16_t*>(&h); }
    __nv_bfloat16 b = __float2bfloat16_rn(v); *back = __bfloat162float(b); return *reinterpret_cast<uint16_t*>(&b);
}

int main() {
    uint16_t *hA = new uint16_t[M * K], *hB = new uint16_t[N * K], *dA, *dB;
    float *fA = new float[M * K], *fB = new float[N * K], *hD = new float[M * N], *dD;
    cudaMalloc(&dA, M * K * 2); cudaMalloc(&dB, N * K * 2); cudaMalloc(&dD, M * N * 4);
    const int smem = 1024 + M * 128 + N * 128 + 64;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int bad = 0;
    for (int combo = 0; combo < 4; ++combo) {
        const int af = combo >> 1, bf = combo & 1;
        srand(7);
        // values with > 8 significant bits so that a format mix-up (fp16 bits read as bf16) cannot go unnoticed
        for (int i = 0; i < M * K; ++i) hA[i] = enc((rand() % 2001 - 1000) / 517.0f, af, &fA[i]);
        for (int i = 0; i < N * K; ++i) hB[i] = enc((rand() % 2001 - 1000) / 733.0f, bf, &fB[i]);
        cudaMemcpy(dA, hA, M * K * 2, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, N * K * 2, cudaMemcpyHostToDevice);
        cudaMemset(dD, 0, M * N * 4);
        probe<<<1, 128, smem>>>(dA, dB, af, bf, dD);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("combo a=%s b=%s: CUDA error %s\n", af ? "bf16" : "f16", bf ? "bf16" : "f16", cudaGetErrorString(e)); return 2; }
        cudaMemcpy(hD, dD, M * N * 4, cudaMemcpyDeviceToHost);
        double worst = 0, ref_max = 0;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) {
                double s = 0;
                for (int k = 0; k < K; ++k) s += double(fA[m * K + k]) * double(fB[n * K + k]);
                worst = fmax(worst, fabs(s - hD[m * N + n])); ref_max = fmax(ref_max, fabs(s));
            }
        const bool ok = worst < 1e-4 * ref_max;
        printf("combo a=%s b=%s: max abs err %.3e (max |ref| %.3f) -> %s\n", af ? "bf16" : "f16", bf ? "bf16" : "f16", worst, ref_max, ok ? "OK" : "MISMATCH");
        bad += !ok;
    }
    printf(bad ? "RESULT: mixed formats NOT usable\n" : "RESULT: all four format combinations are exact\n");
    return bad ? 1 : 0;
}
