// Shared by the tcgen05 convolution kernels (conv_tc.cu, conv_tcx.cu): launch parameters and the
// fused epilogue of one accumulator row (= one output pixel, BN channels):
//   tcgen05.ld -> (+ correction accumulator) -> bias -> ReLU mask -> add -> split-bf16 store(s)
#pragma once
#include "common.cuh"
#include "ptx.cuh"

#ifndef TC_TRACE
#define TC_TRACE(role, tile, ev) do { } while (0)
#endif

namespace {

constexpr int TILE_H = 8, TILE_W = 16, BLOCK_M = 128;
// three warpgroups: warp 0 TMA producer, warp 1 MMA issuer (warps 2, 3 idle), warps 4..11 epilogue.  After set-up the
// first warpgroup hands most of its registers to the two epilogue warpgroups (setmaxnreg): the epilogue keeps a
// 32-column group of up to three accumulators plus the prefetched mask / addend of the next tile in registers.
constexpr int NUM_THREADS = 384;
constexpr int EPI_WARPS = 8, EPI_FIRST_WARP = 4;
constexpr int REGS_CONTROL = 72, REGS_EPILOGUE = 216;      // 128 * 72 + 256 * 216 = 64512 <= 65536

struct ConvTcParams {
    int N, H, W, Cin, Cout;
    int tiles_h, tiles_w, n_blocks, work;
    int w_img_rows;                 // weight-row offset per image (0: shared weights)
    int w_rows_total;               // rows of one weight plane
    int w_resident;                 // conv_tc: all weight tiles stay in shared memory (loaded once per CTA)
    uint32_t fmt_bits;              // operand-format bits of the instruction descriptor (ptx::umma_fmt_bits)
    int out_planes;                 // plane CODE (common.cuh) of the output tensors
    int add_code;                   // plane code of the addend
    float acc_mul;                  // acc *= acc_mul * (acc_mul_slot ? acc_mul_slot[1] : 1) before the bias: undoes the
    const float* acc_mul_slot;      //   fixed factor of fp16 weights and the power-of-two scale slot of a scaled operand
    const float* add_mul_slot;      // the addend is multiplied by add_mul_slot[1] (a scaled fp16 gradient)
    const float* bias;              // [Cout] or null
    const bf16* mask;               // [N,H,W,Cout] or null: acc *= (mask > 0); hi plane, bf16 or fp16 (sign test only)
    const bf16* add_hi;             // [N,H,W,Cout] or null: acc += add
    const bf16* add_lo;
    bf16* out_hi; bf16* out_lo;     // result (pre-activation)
    bf16* relu_hi; bf16* relu_lo;   // optional second output relu(result)
    float* out_nchw; int nchw_c;    // optional fp32 NCHW output of the first nchw_c channels (unfused launches only)
    float* out_f32;                 // optional fp32 [pixel][Cout] output (unfused launches only)
    int dbg;                        // experiments (tools/bench_conv.py --dbg): 1 no global stores, 2 no MMAs, 4 no activation loads
};

// ---------------------------------------------------------------------------------------------
// Epilogue: 8 warps, two per TMEM lane quarter, each owning half of the tile's BN columns.
//   unfused launches (no mask / addend): rolled loop over 16-column chunks;
//   fused launches: the mask / addend of a 32-channel group live in a register buffer that is
//   filled one tile AHEAD (two buffers), so their global-load latency never sits between the
//   accumulator becoming ready and its store (the backward-data launches were bound by that).
// All global accesses are 256-bit: the 16 channels one tcgen05.ld delivers are one 32-byte sector
// per plane.
// ---------------------------------------------------------------------------------------------
struct TcPix { int nb; size_t pix; bool valid; };
struct TcPre { uint32_t mk[2][8], ah[2][8], al[2][8]; };

template <int GCH>
__device__ __forceinline__ void tc_prefetch(const ConvTcParams& p, TcPre& b, const TcPix& px, int ch0) {
    if (!px.valid) return;
    const size_t off = px.pix * p.Cout + ch0;
    if (p.mask) {
#pragma unroll
        for (int v = 0; v < GCH / 16; ++v) ldg256(p.mask + off + v * 16, b.mk[v]);
    }
    if (p.add_hi) {
#pragma unroll
        for (int v = 0; v < GCH / 16; ++v) ldg256(p.add_hi + off + v * 16, b.ah[v]);
        if (p.add_lo && pc_planes(p.add_code) >= 2) {
#pragma unroll
            for (int v = 0; v < GCH / 16; ++v) ldg256(p.add_lo + off + v * 16, b.al[v]);
        }
    }
}

// one 16-channel chunk: accumulators -> bias -> mask -> addend -> stores
struct TcMul { float acc, add; };      // accumulator / addend multipliers of this launch (1 unless fp16 operands are involved)
__device__ __forceinline__ TcMul tc_muls(const ConvTcParams& p) {
    TcMul m;
    m.acc = p.acc_mul * slot_inv(p.acc_mul_slot);
    m.add = slot_inv(p.add_mul_slot);
    return m;
}

// accumulator columns [col, col + 16) of up to three accumulators -> registers (no wait)
template <int BN, int NACC>
__device__ __forceinline__ void tc_load_chunk(uint32_t taddr, int col, uint32_t (&r)[16], uint32_t (&rc)[16], uint32_t (&rd)[16]) {
    ptx::tmem_ld16(taddr + col, r);
    if (NACC > 1) ptx::tmem_ld16(taddr + BN + col, rc);
    if (NACC > 2) ptx::tmem_ld16(taddr + 2 * BN + col, rd);
}

// one 16-channel chunk, accumulators already in registers: bias -> mask -> addend -> stores.
// bias: [Cout] in shared memory (conv_tc stages it once per CTA) or global, or null.
template <int BN, int NACC, bool FUSED>
__device__ __forceinline__ void tc_chunk_compute(const ConvTcParams& p, const TcMul& mul, const float* bias, const TcPix& px, int col,
                                                 const uint32_t (&r)[16], const uint32_t (&rc)[16], const uint32_t (&rd)[16],
                                                 const uint32_t* mk, const uint32_t* ah, const uint32_t* al) {
    const int ch = px.nb * BN + col;
    const size_t o = px.pix * p.Cout + ch;
    float v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        v[i] = __uint_as_float(r[i]);
        if (NACC == 2) v[i] += __uint_as_float(rc[i]);
        if (NACC == 3) v[i] += __uint_as_float(rc[i]) + __uint_as_float(rd[i]);   // small terms first
    }
    if (bias) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 b = *reinterpret_cast<const float4*>(bias + ch + i * 4);
            v[4 * i] = fmaf(v[4 * i], mul.acc, b.x); v[4 * i + 1] = fmaf(v[4 * i + 1], mul.acc, b.y);
            v[4 * i + 2] = fmaf(v[4 * i + 2], mul.acc, b.z); v[4 * i + 3] = fmaf(v[4 * i + 3], mul.acc, b.w);
        }
    } else if (mul.acc != 1.f) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] *= mul.acc;
    }
    if (FUSED) {
        if (p.mask) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                v[2 * i] = pos16_lo(mk[i]) ? v[2 * i] : 0.f;
                v[2 * i + 1] = pos16_hi(mk[i]) ? v[2 * i + 1] : 0.f;
            }
        }
        if (p.add_hi) {
            const bool lo = p.add_lo && pc_planes(p.add_code) >= 2, f16 = pc_f16(p.add_code);
            float a[16];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float2 t = unpack16x2(ah[i], f16);
                if (lo) { const float2 u = unpack16x2(al[i], f16); t.x += u.x; t.y += u.y; }
                a[2 * i] = t.x; a[2 * i + 1] = t.y;
            }
            const bf16* p2 = lo ? third_plane(p.add_hi, p.add_lo, p.add_code) : nullptr;
            if (p2) {
                uint32_t w[8];
                ldg256(p2 + o, w);
#pragma unroll
                for (int i = 0; i < 8; ++i) { a[2 * i] += bf16lo_to_f(w[i]); a[2 * i + 1] += bf16hi_to_f(w[i]); }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = fmaf(a[i], mul.add, v[i]);
        }
    }
    if (p.out_hi) store16p(p.out_hi, p.out_lo, p.out_planes, o, v);
    if (!FUSED && p.out_f32) {
        uint32_t w0[8], w1[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { w0[i] = __float_as_uint(v[i]); w1[i] = __float_as_uint(v[8 + i]); }
        stg256(p.out_f32 + o, w0); stg256(p.out_f32 + o + 8, w1);
    }
    if (!FUSED && p.out_nchw && ch == 0) {
        // module-boundary output (image gradient): lanes are consecutive pixels of a row
        const size_t HW = static_cast<size_t>(p.H) * p.W;
        const size_t n = px.pix / HW, rem = px.pix - n * HW;
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

// A group of GCH (16 or 32) columns: ALL its accumulator loads are issued before the single wait, and when the group is
// the last one of the tile the TMEM stage is handed back to the MMA warp right after that wait -- the arithmetic and
// the global stores of a tile run while the tensor core already fills the stage again.  (The epilogue is latency
// bound: 8 warps, long dependent chains; per-chunk load -> wait -> store sequences held a stage for the whole epilogue
// and the tensor pipe idled 40 % of the time, profiles/r02o_conv_tc_ncu_full_summary.csv.)
template <int BN, int NACC, int GCH, bool FUSED, class EndTile>
__device__ __forceinline__ void tc_group(const ConvTcParams& p, const TcMul& mul, const float* bias, const TcPre* pre, uint32_t taddr,
                                         const TcPix& px, int col0, bool last, const EndTile& end_tile) {
    constexpr int NCH = GCH / 16;
    uint32_t r[NCH][16], rc[NCH][16], rd[NCH][16];
#pragma unroll
    for (int j = 0; j < NCH; ++j) tc_load_chunk<BN, NACC>(taddr, col0 + j * 16, r[j], rc[j], rd[j]);
    ptx::tmem_ld_wait();
    if (last) end_tile();
    if (!px.valid || (p.dbg & 1)) return;
#pragma unroll
    for (int j = 0; j < NCH; ++j)
        tc_chunk_compute<BN, NACC, FUSED>(p, mul, bias, px, col0 + j * 16, r[j], rc[j], rd[j],
                                          FUSED ? pre->mk[j] : nullptr, FUSED ? pre->ah[j] : nullptr, FUSED ? pre->al[j] : nullptr);
}

// dec(wk, m) -> TcPix of accumulator row m of work item wk.  bias: see tc_chunk_compute.
template <int BN, int NACC, int FMODE = -1, class Decode>
__device__ __forceinline__ void tc_epilogue_warps(const ConvTcParams& p, const float* bias, uint32_t tmem_base, int stage_cols,
                                                  uint64_t* t_full, uint64_t* t_empty, int warp, int lane, const Decode& dec) {
    constexpr int CW = BN >= 32 ? BN / 2 : BN;            // columns per warp
    constexpr int GCH = CW < 32 ? CW : 32;                // channels per register buffer
    constexpr int NGRP = CW / GCH;
    const int q = warp & 3, hf = (warp - EPI_FIRST_WARP) >> 2;
    const int m = q * 32 + lane;
    const bool active = BN >= 32 || hf == 0;
    const int col_base = BN >= 32 ? hf * CW : 0;
    // FMODE >= 0: the launch picked a kernel that holds only one of the two epilogues (half the code: short launches
    // spent a quarter of their samples on instruction-cache misses, profiles/r02al)
    const bool fused = FMODE < 0 ? (p.mask || p.add_hi) : (FMODE == 1);                // warp-uniform
    const TcMul mul = tc_muls(p);
    const int step = gridDim.x;
    auto taddr_of = [&](int it) { return tmem_base + (static_cast<uint32_t>(q * 32) << 16) + (it & 1) * stage_cols; };
    auto begin_tile = [&](int it) { ptx::mbar_wait(&t_full[it & 1], (it >> 1) & 1); ptx::tc_fence_after(); };

    if (!fused || !active) {
        int it = 0;
        for (int wk = blockIdx.x; wk < p.work; wk += step, ++it) {
            const TcPix px = dec(wk, m);
            begin_tile(it);
            if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 0);
            auto end_tile = [&]() { ptx::tc_fence_before(); __syncwarp(); if (lane == 0) ptx::mbar_arrive(&t_empty[it & 1]);
                                    if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 1); };
            if (active) {
#pragma unroll
                for (int g = 0; g < NGRP; ++g)
                    tc_group<BN, NACC, GCH, false>(p, mul, bias, nullptr, taddr_of(it), px, col_base + g * GCH, g == NGRP - 1, end_tile);
            } else {
                end_tile();
            }
            if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 2);
        }
        return;
    }
    TcPre A, B;
    if (NGRP == 2) {
        // buffer A = first 32 channels, B = second 32 channels; each is refilled for the next tile right after use
        int wk = blockIdx.x;
        if (wk < p.work) { const TcPix px = dec(wk, m); tc_prefetch<GCH>(p, A, px, px.nb * BN + col_base); tc_prefetch<GCH>(p, B, px, px.nb * BN + col_base + GCH); }
        for (int it = 0; wk < p.work; wk += step, ++it) {
            const TcPix px = dec(wk, m);
            const bool more = wk + step < p.work;
            TcPix nx = px;
            if (more) nx = dec(wk + step, m);
            begin_tile(it);
            if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 0);
            auto end_tile = [&]() { ptx::tc_fence_before(); __syncwarp(); if (lane == 0) ptx::mbar_arrive(&t_empty[it & 1]);
                                    if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 1); };
            tc_group<BN, NACC, GCH, true>(p, mul, bias, &A, taddr_of(it), px, col_base, false, end_tile);
            if (more) tc_prefetch<GCH>(p, A, nx, nx.nb * BN + col_base);
            tc_group<BN, NACC, GCH, true>(p, mul, bias, &B, taddr_of(it), px, col_base + GCH, true, end_tile);
            if (more) tc_prefetch<GCH>(p, B, nx, nx.nb * BN + col_base + GCH);
            if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 2);
        }
    } else {
        // buffers alternate between tiles: A holds tile k, B tile k + 1; tile k + 2 is fetched into A after tile k
        int wk0 = blockIdx.x, wk1 = wk0 + step;
        if (wk0 < p.work) { const TcPix px = dec(wk0, m); tc_prefetch<GCH>(p, A, px, px.nb * BN + col_base); }
        if (wk1 < p.work) { const TcPix px = dec(wk1, m); tc_prefetch<GCH>(p, B, px, px.nb * BN + col_base); }
        int it = 0;
        auto end_tile = [&]() { ptx::tc_fence_before(); __syncwarp(); if (lane == 0) ptx::mbar_arrive(&t_empty[it & 1]); };
        while (wk0 < p.work) {
            {
                const TcPix px = dec(wk0, m);
                begin_tile(it);
                tc_group<BN, NACC, GCH, true>(p, mul, bias, &A, taddr_of(it), px, col_base, true, end_tile);
                ++it;
            }
            const int wk2 = wk0 + 2 * step;
            if (wk2 < p.work) { const TcPix px = dec(wk2, m); tc_prefetch<GCH>(p, A, px, px.nb * BN + col_base); }
            if (wk1 >= p.work) break;
            {
                const TcPix px = dec(wk1, m);
                begin_tile(it);
                tc_group<BN, NACC, GCH, true>(p, mul, bias, &B, taddr_of(it), px, col_base, true, end_tile);
                ++it;
            }
            const int wk3 = wk1 + 2 * step;
            if (wk3 < p.work) { const TcPix px = dec(wk3, m); tc_prefetch<GCH>(p, B, px, px.nb * BN + col_base); }
            wk0 = wk2; wk1 = wk3;
        }
    }
}

}  // namespace
