// Shared by the tcgen05 convolution kernels (conv_tc.cu, conv_tcx.cu): launch parameters and the
// fused epilogue of one accumulator row (= one output pixel, BN channels):
//   tcgen05.ld -> (+ correction accumulator) -> bias -> ReLU mask -> add -> split-bf16 store(s)
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace {

constexpr int TILE_H = 8, TILE_W = 16, BLOCK_M = 128;
constexpr int NUM_THREADS = 192;

struct ConvTcParams {
    int N, H, W, Cin, Cout;
    int tiles_h, tiles_w, n_blocks, work;
    int w_img_rows;                 // weight-row offset per image (0: shared weights)
    int w_rows_total;               // rows of one weight plane
    int out_planes;                 // bf16 planes of the output tensors (1..3)
    const float* bias;              // [Cout] or null
    const bf16* mask;               // [N,H,W,Cout] or null: acc *= (mask > 0)
    const bf16* add_hi;             // [N,H,W,Cout] or null: acc += add
    const bf16* add_lo;
    bf16* out_hi; bf16* out_lo;     // result (pre-activation)
    bf16* relu_hi; bf16* relu_lo;   // optional second output relu(result)
};

// taddr: TMEM address of this warp's lane quarter / accumulator stage; pix: output pixel index
template <int BN, int NACC>
__device__ __forceinline__ void tc_epilogue_row(const ConvTcParams& p, uint32_t taddr, int nb, size_t pix, bool valid) {
#pragma unroll 1
    for (int j = 0; j < BN / 16; ++j) {
        uint32_t r[16], rc[16];
        ptx::tmem_ld16(taddr + j * 16, r);
        if (NACC > 1) ptx::tmem_ld16(taddr + BN + j * 16, rc);
        ptx::tmem_ld_wait();
        if (valid) {
            const int ch = nb * BN + j * 16;
            const size_t off = pix * p.Cout + ch;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    v[i] = __uint_as_float(r[half * 8 + i]);
                    if (NACC > 1) v[i] += __uint_as_float(rc[half * 8 + i]);
                }
                const size_t o = off + half * 8;
                if (p.bias) {
                    const float4 b0 = *reinterpret_cast<const float4*>(p.bias + ch + half * 8);
                    const float4 b1 = *reinterpret_cast<const float4*>(p.bias + ch + half * 8 + 4);
                    v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
                    v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
                }
                if (p.mask) {
                    float mk[8];
                    load8(p.mask, nullptr, o, mk);
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = mk[i] > 0.f ? v[i] : 0.f;
                }
                if (p.add_hi) {
                    float ad[8];
                    load8p(p.add_hi, p.add_lo, p.add_lo ? p.out_planes : 1, o, ad);   // add carries the output's planes
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] += ad[i];
                }
                if (p.out_hi) store8p(p.out_hi, p.out_lo, p.out_planes, o, v);
                if (p.relu_hi) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
                    store8p(p.relu_hi, p.relu_lo, p.out_planes, o, v);
                }
            }
        }
    }
}

}  // namespace
