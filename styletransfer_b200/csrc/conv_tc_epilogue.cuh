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

// n / d for 0 <= n < 2^31 and a run-time d >= 1 as a multiply-high, an add and a shift: the work-item decode of every tile
// sat on the epilogue's critical path with three hardware-less 32-bit divisions (each a dependent chain of I2F / MUFU.RCP /
// F2I / R2UR, a few hundred cycles of latency that the compiler also rematerialised inside the channel chunks).
struct FastDiv {
    uint32_t mul, sh, d;
    void init(int div) {
        d = static_cast<uint32_t>(div);
        sh = 0; while ((1ull << sh) < d) ++sh;
        mul = static_cast<uint32_t>((((1ull << sh) - d) << 32) / d + 1);
    }
    __device__ __forceinline__ uint32_t div(uint32_t n) const { return (__umulhi(n, mul) + n) >> sh; }
    __device__ __forceinline__ void divmod(uint32_t n, uint32_t& q, uint32_t& r) const { q = div(n); r = n - q * d; }
};

struct ConvTcParams {
    FastDiv fd_nb, fd_tw, fd_th, fd_W, fd_H;   // n_blocks, tiles_w, tiles_h, W, H
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
    // add_route != null: the addend is the gradient of a 2x2 max-pool's OUTPUT, [N,H/2,W/2,Cout], routed on the fly:
    // pixel (h, w) takes channel c of pooled pixel (h/2, w/2) where the routing byte equals (h&1)*2 + (w&1), else 0
    // (relu_maxpool_bwd_idx_kernel folded into this epilogue; bytes of 4 mark windows whose maximum was <= 0)
    const unsigned char* add_route;
    int route_Ho, route_Wo;
    // pool_idx != null (unfused 3x3 launches): the epilogue also writes relu(maxpool2x2(result)) [N][H/2][W/2][Cout] in the
    // planes of `out` and the routing byte of every pooled element (relu_maxpool_fwd_idx_kernel folded into the conv that
    // produces its input); pool_lane_h = lane distance of the pixel one row down in the accumulator tile (its width)
    bf16* pool_hi; bf16* pool_lo; unsigned char* pool_idx;
    int pool_lane_h, pool_Ho, pool_Wo;
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
struct TcPre { uint32_t mk[2][8], ah[2][8], al[2][8], rt[2][4], pos; };

template <int GCH>
__device__ __forceinline__ void tc_prefetch(const ConvTcParams& p, TcPre& b, const TcPix& px, int ch0) {
    if (!px.valid) return;
#ifdef ST_CONV_TRACE
    if (p.dbg & 64) return;      // experiment: no mask / addend loads (the buffers keep whatever they held)
#endif
    const size_t off = px.pix * p.Cout + ch0;
    if (p.mask) {
#pragma unroll
        for (int v = 0; v < GCH / 16; ++v) ldg256(p.mask + off + v * 16, b.mk[v]);
    }
    if (p.add_hi) {
        size_t aoff = off;
        if (p.add_route) {
            // 32-bit arithmetic (the launcher checks N*H*W < 2^31)
            uint32_t t, w, n, h;
            p.fd_W.divmod(static_cast<uint32_t>(px.pix), t, w);
            p.fd_H.divmod(t, n, h);
            const int oh = static_cast<int>(h >> 1), ow = static_cast<int>(w >> 1);
            const bool in = oh < p.route_Ho && ow < p.route_Wo;         // odd sizes: the last row / column is not pooled
            // unpooled border: position 5 matches no routing byte (0..4) -- decided here without touching the loaded bytes,
            // so nothing in this function waits for a load
            b.pos = in ? (((h & 1) << 1) | (w & 1)) : 5u;
            aoff = ((static_cast<size_t>(n) * p.route_Ho + (in ? oh : 0)) * p.route_Wo + (in ? ow : 0)) * p.Cout + ch0;
#pragma unroll
            for (int v = 0; v < GCH / 16; ++v) {
                const uint4 r = *reinterpret_cast<const uint4*>(p.add_route + aoff + v * 16);
                b.rt[v][0] = r.x; b.rt[v][1] = r.y; b.rt[v][2] = r.z; b.rt[v][3] = r.w;
            }
        }
#pragma unroll
        for (int v = 0; v < GCH / 16; ++v) ldg256(p.add_hi + aoff + v * 16, b.ah[v]);
        if (p.add_lo && pc_planes(p.add_code) >= 2) {
#pragma unroll
            for (int v = 0; v < GCH / 16; ++v) ldg256(p.add_lo + aoff + v * 16, b.al[v]);
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

// eight 16-bit-pair words per plane -> 16 floats (hi [+ lo])
template <bool F16, bool LO>
__device__ __forceinline__ void tc_unpack_add(const uint32_t (&h)[8], const uint32_t (&l)[8], float (&a)[16]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float2 t = F16 ? unpack_f16x2(h[i]) : make_float2(bf16lo_to_f(h[i]), bf16hi_to_f(h[i]));
        if (LO) {
            const float2 u = F16 ? unpack_f16x2(l[i]) : make_float2(bf16lo_to_f(l[i]), bf16hi_to_f(l[i]));
            t.x += u.x; t.y += u.y;
        }
        a[2 * i] = t.x; a[2 * i + 1] = t.y;
    }
}

// one 16-channel chunk, accumulators already in registers: bias -> mask -> addend -> stores.
// bias: [Cout] in shared memory (conv_tc stages it once per CTA) or global, or null.
template <int BN, int NACC, bool FUSED>
__device__ __forceinline__ void tc_chunk_values(const ConvTcParams& p, const TcMul& mul, const float* bias, const TcPix& px, int col,
                                                const uint32_t (&r)[16], const uint32_t (&rc)[16], const uint32_t (&rd)[16],
                                                const uint32_t* mk, const uint32_t* ah, const uint32_t* al,
                                                const uint32_t* rt, uint32_t pos, float (&v)[16]) {
    const int ch = px.nb * BN + col;
    const size_t o = px.pix * p.Cout + ch;
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
#ifdef ST_CONV_TRACE
        const bool dbg_no_add = p.dbg & 256, dbg_no_route = p.dbg & 128;
#else
        constexpr bool dbg_no_add = false, dbg_no_route = false;
#endif
        if (p.add_hi && !dbg_no_add) {
            const bool lo = p.add_lo && pc_planes(p.add_code) >= 2, f16 = pc_f16(p.add_code);
            uint32_t h[8], l[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) { h[i] = ah[i]; l[i] = lo ? al[i] : 0u; }
            if (p.add_route && !dbg_no_route) {
                // routing bytes are 0..4, pos 0..3 (5: no gradient): a byte of (rt ^ pos4) is 0 exactly where the gradient goes.  Byte-wise
                // zero test without carries (bytes <= 7), then PRMT replicates each byte's top bit over a 16-bit lane:
                // the words of both planes are masked BEFORE they are unpacked (5 + 4 instructions per 4 channels)
                const uint32_t pos4 = pos * 0x01010101u;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t e = ~((rt[j] ^ pos4) + 0x7f7f7f7fu) & 0x80808080u;
                    const uint32_t m0 = ptx::prmt(e, 0u, 0x9988u), m1 = ptx::prmt(e, 0u, 0xbbaau);    // selector bit 3: replicate the byte's sign
                    h[2 * j] &= m0; h[2 * j + 1] &= m1; l[2 * j] &= m0; l[2 * j + 1] &= m1;
                }
            }
            float a[16];
            // one warp-uniform branch per chunk instead of a predicated copy of both conversions per element
            if (f16) {
                if (lo) tc_unpack_add<true, true>(h, l, a); else tc_unpack_add<true, false>(h, l, a);
            } else {
                if (lo) tc_unpack_add<false, true>(h, l, a); else tc_unpack_add<false, false>(h, l, a);
            }
            const bf16* p2 = (lo && !p.add_route) ? third_plane(p.add_hi, p.add_lo, p.add_code) : nullptr;
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
}

template <int BN, bool FUSED>
__device__ __forceinline__ void tc_chunk_store(const ConvTcParams& p, const TcPix& px, int col, float (&v)[16]);

template <int BN, int NACC, bool FUSED>
__device__ __forceinline__ void tc_chunk_compute(const ConvTcParams& p, const TcMul& mul, const float* bias, const TcPix& px, int col,
                                                 const uint32_t (&r)[16], const uint32_t (&rc)[16], const uint32_t (&rd)[16],
                                                 const uint32_t* mk, const uint32_t* ah, const uint32_t* al,
                                                 const uint32_t* rt = nullptr, uint32_t pos = 0) {
    const int ch = px.nb * BN + col;
    const size_t o = px.pix * p.Cout + ch;
    float v[16];
    tc_chunk_values<BN, NACC, FUSED>(p, mul, bias, px, col, r, rc, rd, mk, ah, al, rt, pos, v);
    tc_chunk_store<BN, FUSED>(p, px, col, v);
}

// the stores of one 16-channel chunk of a valid pixel
template <int BN, bool FUSED>
__device__ __forceinline__ void tc_chunk_store(const ConvTcParams& p, const TcPix& px, int col, float (&v)[16]) {
    const int ch = px.nb * BN + col;
    const size_t o = px.pix * p.Cout + ch;
#ifdef ST_CONV_TRACE
    if (p.dbg & 32) {      // experiment: all the arithmetic of the split, (practically) no store
        uint32_t hw[8], mw[8], x = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { split2h(v[2 * i], v[2 * i + 1], hw[i], mw[i]); x ^= hw[i] ^ mw[i]; }
        if (x == 0x9e3779b9u) stg256(p.out_hi + o, hw);
        return;
    }
#endif
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

// relu + 2x2 max-pool of one 16-channel chunk inside the accumulator tile: the four pixels of a window are lanes l, l^1,
// l^W, l^(W|1) of one warp (tile rows are W = pool_lane_h pixels wide, a warp holds whole row pairs); ALL lanes call this.
// First maximum in PyTorch's scan order wins, byte 4 = the window maximum is <= 0 (relu' = 0 everywhere).
template <int BN>
__device__ __forceinline__ void tc_pool_chunk(const ConvTcParams& p, const TcPix& px, int col, const float (&v)[16]) {
    const int lane = threadIdx.x & 31, dh = p.pool_lane_h;
    float best[16];
    uint32_t idx[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float a = v[i];
        const float b = __shfl_xor_sync(0xffffffffu, a, 1), c = __shfl_xor_sync(0xffffffffu, a, dh), d = __shfl_xor_sync(0xffffffffu, a, dh | 1);
        float m = a; uint32_t k = 0u;
        if (b > m) { m = b; k = 1u; }
        if (c > m) { m = c; k = 2u; }
        if (d > m) { m = d; k = 3u; }
        if (!(m > 0.f)) { m = 0.f; k = 4u; }
        best[i] = m;
        idx[i >> 2] |= k << (8 * (i & 3));
    }
    if (!(lane & 1) && !(lane & dh) && px.valid) {
        uint32_t t, w, n, h;
        p.fd_W.divmod(static_cast<uint32_t>(px.pix), t, w);
        p.fd_H.divmod(t, n, h);
        const int oh = static_cast<int>(h >> 1), ow = static_cast<int>(w >> 1);
        if (oh < p.pool_Ho && ow < p.pool_Wo) {
            const size_t o = ((static_cast<size_t>(n) * p.pool_Ho + oh) * p.pool_Wo + ow) * p.Cout + px.nb * BN + col;
            store16p(p.pool_hi, p.pool_lo, p.out_planes, o, best);
            *reinterpret_cast<uint4*>(p.pool_idx + o) = make_uint4(idx[0], idx[1], idx[2], idx[3]);
        }
    }
}

// A group of GCH (16 or 32) columns: ALL its accumulator loads are issued before the single wait, and when the group is
// the last one of the tile the TMEM stage is handed back to the MMA warp right after that wait -- the arithmetic and
// the global stores of a tile run while the tensor core already fills the stage again.  (The epilogue is latency
// bound: 8 warps, long dependent chains; per-chunk load -> wait -> store sequences held a stage for the whole epilogue
// and the tensor pipe idled 40 % of the time, profiles/r02o_conv_tc_ncu_full_summary.csv.)
// TS (transposed stores): the group's 32 channels x 2 planes of the warp's 32 pixels go through a 4 KB shared-memory tile
// of the warp and leave as 16-byte stores whose lanes run along the CHANNELS of a pixel (4 lanes = 64 contiguous bytes,
// 8 pixels per instruction) instead of one 32-byte sector per lane in 32 different 128-byte lines: the LSU spends about
// two cycles per line an instruction touches (tools/probe_store_pattern.cu: 4.6 TB/s for the one-pixel-per-lane pattern
// with nothing else going on), which made the 1x1 launches store-issue bound.  Rows of the tile are 64 bytes, the
// 16-byte pieces of a row are XOR-swizzled with bits 1..2 of the row so that writes (8 rows, one piece) and reads
// (2 rows, 4 pieces) both cover all 32 banks.
constexpr int TS_WARP_BYTES = 4096;
template <int BN, int NACC, int GCH, bool FUSED, bool TS, class EndTile>
__device__ __forceinline__ void tc_group(const ConvTcParams& p, const TcMul& mul, const float* bias, const TcPre* pre, uint32_t taddr,
                                         const TcPix& px, int col0, bool last, const EndTile& end_tile, uint8_t* stage = nullptr) {
    constexpr int NCH = GCH / 16;
    uint32_t r[NCH][16], rc[NCH][16], rd[NCH][16];
#pragma unroll
    for (int j = 0; j < NCH; ++j) tc_load_chunk<BN, NACC>(taddr, col0 + j * 16, r[j], rc[j], rd[j]);
    ptx::tmem_ld_wait();
    if (last) end_tile();
    if constexpr (!TS) {
        if (!FUSED && p.pool_idx) {
            // every lane computes its values (the pool exchanges them between lanes), valid pixels store
#pragma unroll
            for (int j = 0; j < NCH; ++j) {
                float v[16];
                tc_chunk_values<BN, NACC, false>(p, mul, bias, px, col0 + j * 16, r[j], rc[j], rd[j], nullptr, nullptr, nullptr, nullptr, 0u, v);
                tc_pool_chunk<BN>(p, px, col0 + j * 16, v);
                if (px.valid && (p.out_hi || p.relu_hi)) tc_chunk_store<BN, false>(p, px, col0 + j * 16, v);
            }
            return;
        }
        if (!px.valid || (p.dbg & 1)) return;
#pragma unroll
        for (int j = 0; j < NCH; ++j)
            tc_chunk_compute<BN, NACC, FUSED>(p, mul, bias, px, col0 + j * 16, r[j], rc[j], rd[j],
                                              FUSED ? pre->mk[j] : nullptr, FUSED ? pre->ah[j] : nullptr, FUSED ? pre->al[j] : nullptr,
                                              FUSED ? pre->rt[j] : nullptr, FUSED ? pre->pos : 0u);
    } else {
        static_assert(GCH == 32, "transposed stores: 32-channel groups");
        const int lane = threadIdx.x & 31;
        const bool f16 = pc_f16(p.out_planes);
        float v[NCH][16];
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[j][i] = 0.f;
            if (px.valid)
                tc_chunk_values<BN, NACC, FUSED>(p, mul, bias, px, col0 + j * 16, r[j], rc[j], rd[j],
                                                 FUSED ? pre->mk[j] : nullptr, FUSED ? pre->ah[j] : nullptr, FUSED ? pre->al[j] : nullptr,
                                                 FUSED ? pre->rt[j] : nullptr, FUSED ? pre->pos : 0u, v[j]);
        }
        const unsigned vm = (p.dbg & 1) ? 0u : __ballot_sync(0xffffffffu, px.valid);
        const unsigned long long obase = static_cast<unsigned long long>(px.pix) * p.Cout + px.nb * BN + col0;
        const int wsw = (lane >> 1) & 3, c = lane & 3;
        // split the group's values into two 16-bit planes, stage them, store them with lanes along the channels
        auto flush = [&](bf16* dst_hi, bf16* dst_lo, bool relu) {
            const bool two = dst_lo != nullptr;
#pragma unroll
            for (int j = 0; j < NCH; ++j) {
                uint32_t hw[8], mw[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float a = relu ? fmaxf(v[j][2 * i], 0.f) : v[j][2 * i], b = relu ? fmaxf(v[j][2 * i + 1], 0.f) : v[j][2 * i + 1];
                    if (f16) split2h(a, b, hw[i], mw[i]); else split2(a, b, hw[i], mw[i]);
                }
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint8_t* dst = stage + lane * 64 + (((2 * j + h) ^ wsw) << 4);
                    *reinterpret_cast<uint4*>(dst) = make_uint4(hw[4 * h], hw[4 * h + 1], hw[4 * h + 2], hw[4 * h + 3]);
                    if (two) *reinterpret_cast<uint4*>(dst + 2048) = make_uint4(mw[4 * h], mw[4 * h + 1], mw[4 * h + 2], mw[4 * h + 3]);
                }
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int prow = 8 * j + (lane >> 2);
                const unsigned long long o = __shfl_sync(0xffffffffu, obase, prow) + c * 8;
                const uint8_t* src = stage + prow * 64 + ((c ^ ((prow >> 1) & 3)) << 4);
                if ((vm >> prow) & 1u) {
                    *reinterpret_cast<uint4*>(dst_hi + o) = *reinterpret_cast<const uint4*>(src);
                    if (two) *reinterpret_cast<uint4*>(dst_lo + o) = *reinterpret_cast<const uint4*>(src + 2048);
                }
            }
            __syncwarp();
        };
        if (p.out_hi) flush(p.out_hi, p.out_lo, false);
        if (p.relu_hi) flush(p.relu_hi, p.relu_lo, true);
    }
}

// dec(wk, m) -> TcPix of accumulator row m of work item wk.  bias: see tc_chunk_compute.
// TS: transposed stores (tc_group); ts_stage: 8 x TS_WARP_BYTES of shared memory, 16-byte aligned
template <int BN, int NACC, int FMODE = -1, bool TS = false, class Decode>
__device__ __forceinline__ void tc_epilogue_warps(const ConvTcParams& p, const float* bias, uint32_t tmem_base, int stage_cols,
                                                  uint64_t* t_full, uint64_t* t_empty, int warp, int lane, const Decode& dec,
                                                  uint8_t* ts_stage = nullptr) {
    uint8_t* stage = TS ? ts_stage + (warp - EPI_FIRST_WARP) * TS_WARP_BYTES : nullptr;
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
#pragma unroll 1
                for (int g = 0; g < NGRP; ++g)
                    tc_group<BN, NACC, GCH, false, TS>(p, mul, bias, nullptr, taddr_of(it), px, col_base + g * GCH, g == NGRP - 1, end_tile, stage);
            } else {
                end_tile();
            }
            if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 2);
        }
        return;
    }
    // One rolled loop over (tile, 32-channel group) steps.  B receives the NEXT step's mask / addend at the start of a step
    // (before this step's arithmetic and stores enter the LSU queue) and is copied into A when that step begins.  The
    // epilogue runs each instruction once per tile, so its speed is set by instruction fetch: unrolled pairs of tiles /
    // groups with alternating buffers doubled the straight-line code of the loop and cost more than the copy.
    TcPre A, B;
    int it = 0, g = 0, wk = blockIdx.x;
    auto end_tile = [&]() { ptx::tc_fence_before(); __syncwarp(); if (lane == 0) ptx::mbar_arrive(&t_empty[it & 1]);
                            if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 1); };
    TcPix px;
    if (wk < p.work) { px = dec(wk, m); tc_prefetch<GCH>(p, B, px, px.nb * BN + col_base); }
#pragma unroll 1
    while (wk < p.work) {
        A = B;
        int ng = g + 1, nwk = wk;
        if (ng == NGRP) { ng = 0; nwk = wk + step; }
        TcPix nx = px;
        if (nwk < p.work) {
            if (ng == 0) nx = dec(nwk, m);
            tc_prefetch<GCH>(p, B, nx, nx.nb * BN + col_base + ng * GCH);
        }
        if (g == 0) {
            begin_tile(it);
            if (warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 0);
        }
        tc_group<BN, NACC, GCH, true, TS>(p, mul, bias, &A, taddr_of(it), px, col_base + g * GCH, g == NGRP - 1, end_tile, stage);
        if (g == NGRP - 1 && warp == EPI_FIRST_WARP && lane == 0) TC_TRACE(2, it, 2);
        g = ng; px = nx;
        if (ng == 0) { wk = nwk; ++it; }
    }
}

}  // namespace
