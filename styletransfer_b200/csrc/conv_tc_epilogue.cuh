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
    float* out_nchw; int nchw_c;    // optional fp32 NCHW output of the first nchw_c channels (unfused launches only)
};

// Epilogue of one accumulator row in groups of up to 64 channels.  The ReLU mask and the addend of a
// group do not depend on the accumulator: prefetch() issues those loads BEFORE the wait on the
// accumulator barrier, so their latency overlaps the MMAs of the tile instead of following them
// (the backward-data launches were bound by exactly that latency).  All global accesses are 256-bit:
// the 16 channels one tcgen05.ld delivers are one full 32-byte sector per plane.
template <int BN, int NACC>
struct TcEpilogue {
    static constexpr int GC = BN < 64 ? BN : 64;     // channels per group
    static constexpr int GV = GC / 16, NG = BN / GC;
    uint32_t mk[GV][8], ah[GV][8], al[GV][8];

    __device__ __forceinline__ void prefetch(const ConvTcParams& p, int nb, size_t pix, bool valid, int g) {
        if (!valid) return;
        const size_t off = pix * p.Cout + nb * BN + g * GC;
        if (p.mask) {
#pragma unroll
            for (int v = 0; v < GV; ++v) ldg256(p.mask + off + v * 16, mk[v]);
        }
        if (p.add_hi) {
#pragma unroll
            for (int v = 0; v < GV; ++v) ldg256(p.add_hi + off + v * 16, ah[v]);
            if (p.add_lo && p.out_planes >= 2) {
#pragma unroll
                for (int v = 0; v < GV; ++v) ldg256(p.add_lo + off + v * 16, al[v]);
            }
        }
    }

    // taddr: TMEM address of this warp's lane quarter / accumulator stage; pix: output pixel index
    __device__ __forceinline__ void process(const ConvTcParams& p, uint32_t taddr, int nb, size_t pix, bool valid, int g) {
#pragma unroll
        for (int j = 0; j < GV; ++j) {
            const int col = g * GC + j * 16;
            uint32_t r[16], rc[16], rd[16];
            ptx::tmem_ld16(taddr + col, r);
            if (NACC > 1) ptx::tmem_ld16(taddr + BN + col, rc);
            if (NACC > 2) ptx::tmem_ld16(taddr + 2 * BN + col, rd);
            ptx::tmem_ld_wait();
            if (valid) {
                const int ch = nb * BN + col;
                const size_t o = pix * p.Cout + ch;
                float v[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    v[i] = __uint_as_float(r[i]);
                    if (NACC == 2) v[i] += __uint_as_float(rc[i]);
                    if (NACC == 3) v[i] += __uint_as_float(rc[i]) + __uint_as_float(rd[i]);   // small terms first
                }
                if (p.bias) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 b = *reinterpret_cast<const float4*>(p.bias + ch + i * 4);
                        v[4 * i] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
                    }
                }
                if (p.mask) {
                    float mv[16];
                    unpack16(mk[j], mv);
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = mv[i] > 0.f ? v[i] : 0.f;
                }
                if (p.add_hi) {
                    float ad[16], t[16];
                    unpack16(ah[j], ad);
                    if (p.add_lo && p.out_planes >= 2) {
                        unpack16(al[j], t);
#pragma unroll
                        for (int i = 0; i < 16; ++i) ad[i] += t[i];
                        const bf16* p2 = third_plane(p.add_hi, p.add_lo, p.out_planes);   // add carries the output's planes
                        if (p2) {
                            uint32_t w[8];
                            ldg256(p2 + o, w);
                            unpack16(w, t);
#pragma unroll
                            for (int i = 0; i < 16; ++i) ad[i] += t[i];
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] += ad[i];
                }
                if (p.out_hi) store16p(p.out_hi, p.out_lo, p.out_planes, o, v);
                if (p.relu_hi) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
                    store16p(p.relu_hi, p.relu_lo, p.out_planes, o, v);
                }
            }
        }
    }
};

// Rolled epilogue for launches without mask / addend (forward convs, Gram backward).
template <int BN, int NACC>
__device__ __forceinline__ void tc_epilogue_row(const ConvTcParams& p, uint32_t taddr, int nb, size_t pix, bool valid) {
#pragma unroll 1
    for (int j = 0; j < BN / 16; ++j) {
        uint32_t r[16], rc[16], rd[16];
        ptx::tmem_ld16(taddr + j * 16, r);
        if (NACC > 1) ptx::tmem_ld16(taddr + BN + j * 16, rc);
        if (NACC > 2) ptx::tmem_ld16(taddr + 2 * BN + j * 16, rd);
        ptx::tmem_ld_wait();
        if (valid) {
            const int ch = nb * BN + j * 16;
            const size_t o = pix * p.Cout + ch;
            float v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                v[i] = __uint_as_float(r[i]);
                if (NACC == 2) v[i] += __uint_as_float(rc[i]);
                if (NACC == 3) v[i] += __uint_as_float(rc[i]) + __uint_as_float(rd[i]);   // small terms first
            }
            if (p.bias) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 b = *reinterpret_cast<const float4*>(p.bias + ch + i * 4);
                    v[4 * i] += b.x; v[4 * i + 1] += b.y; v[4 * i + 2] += b.z; v[4 * i + 3] += b.w;
                }
            }
            if (p.out_hi) store16p(p.out_hi, p.out_lo, p.out_planes, o, v);
            if (p.out_nchw && j == 0) {
                // module-boundary output (image gradient): lanes are consecutive pixels of a row
                const size_t HW = static_cast<size_t>(p.H) * p.W;
                const size_t n = pix / HW, rem = pix - n * HW;
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (c < p.nchw_c) p.out_nchw[(n * p.nchw_c + c) * HW + rem] = v[c];
            }
            if (p.relu_hi) {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
                store16p(p.relu_hi, p.relu_lo, p.out_planes, o, v);
            }
        }
    }
}

}  // namespace
