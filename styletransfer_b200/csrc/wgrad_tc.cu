// tcgen05 / TMEM / TMA backward-weight of a stride-1 "same" convolution (transform net:
// autograd of the nn.Conv2d call sites network.py:468-481, :582-599).
//
//   dW[co][ci][r][s] = sum over pixels p of  dy[p][co] * x[p + (r - pad, s - pad)][ci]
//
// GEMM view (both operands are pixel-major in memory, NHWC, so both are "MN-major" UMMA
// operands: the contraction index K = pixel is the slow one):
//   K = 128 pixels (an 8 x 16 patch), N = Cdy gradient channels (32 / 64 / 128),
//   M = 128 = TWO vertical taps (r, r+1) x 64 input channels of one horizontal tap s:
//       one TMA box of (8 + KH - 1) x 16 pixels x 64 channels holds every vertical tap, tap r
//       starts r * 2048 bytes into it, so the M-group stride (LBO) is simply 2048.
//   An odd last tap pairs with 128 rows of don't-care data (rows are independent; they are
//   never written out).
// Work item = (horizontal tap s, 64-channel chunk of Cin, pixel split); the CTA walks its pixel
// tiles accumulating ceil(KH/2) accumulators of 128 x Cdy in TMEM, then writes fp32 partials; a
// finish kernel sums the splits in a fixed order (deterministic, no atomics) into the PyTorch
// weight-gradient layout.  Gradients carry 2 bf16 planes: 3 MMAs per product.
//
// Warp roles (192 threads): warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 epilogue.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.cuh"

static int g_gram_max_splits = 0;     // 0: automatic (see plan_gram); > 0: tuning override (st_set_gram_max_splits)
extern "C" int st_set_gram_max_splits(int v) { const int old = g_gram_max_splits; g_gram_max_splits = v; return old; }

namespace {

constexpr int TILE_H = 8, TILE_W = 16, TILE_PIX = 128;
constexpr int NUM_THREADS = 192;
constexpr int PL = 2;                       // planes read from each operand

struct WgradParams {
    int N, H, W, Cin, Cdy;
    int tiles_h, tiles_w, tiles;            // pixel tiles
    int chunks;                             // Cin / 64
    int splits, tiles_per_split;
    int pad_w;                              // horizontal padding (KW / 2)
    uint32_t fmt_bits;                      // operand format bits of the instruction descriptor (both operands alike)
    float* partial;                         // [split][s][chunk][acc][128][Cdy]
};

template <int KH, int CDY> struct WCfg {
    static constexpr int A_ROWS = (TILE_H + KH - 1) * TILE_W;
    static constexpr int A_PLANE = A_ROWS * 128;                 // 64 channels per row
    static constexpr int NBOX = CDY > 64 ? CDY / 64 : 1;            // dy boxes per plane
    static constexpr int DROW = CDY >= 64 ? 128 : CDY * 2;          // bytes per dy row (per box)
    static constexpr int B_BOX = TILE_PIX * DROW;
    static constexpr int B_PLANE = NBOX * B_BOX;
    static constexpr int STAGE = PL * (A_PLANE + B_PLANE);
    static constexpr int NACC = (KH + 1) / 2;
    static constexpr int TMEM_NEED = NACC * CDY;
    static constexpr int TMEM_COLS = TMEM_NEED <= 32 ? 32 : TMEM_NEED <= 64 ? 64 : TMEM_NEED <= 128 ? 128 : TMEM_NEED <= 256 ? 256 : 512;
    static_assert(TMEM_NEED <= 512, "accumulators exceed TMEM");
    static constexpr int SMEM_MAX = 232448;
    static constexpr int SLACK = 4096;                               // the odd tap's partner reads past the last box
    static constexpr int NS_FIT = (SMEM_MAX - 1024 - 256 - SLACK) / STAGE;
    static constexpr int NS = NS_FIT > 4 ? 4 : NS_FIT;
    static_assert(NS >= 2, "need two pipeline stages");
    static constexpr int SMEM = 1024 + NS * STAGE + SLACK + 256;
};

// MN-major shared-memory descriptor halves (SWIZZLE_128B: layout 2, rows of 128 B, 8-row groups
// 1024 B apart; SWIZZLE_64B: layout 4, rows of 64 B, groups 512 B apart)
__device__ __forceinline__ uint32_t mn_desc_lo(uint32_t addr, uint32_t lbo_bytes) {
    return ((addr & 0x3ffff) >> 4) | ((lbo_bytes >> 4) << 16);
}
__host__ __device__ constexpr uint32_t mn_desc_hi(int row_bytes) {
    return static_cast<uint32_t>((8 * row_bytes) >> 4) | (1u << 14) | ((row_bytes == 128 ? 2u : 4u) << 29);
}
// kind::f16 instruction descriptor, fp32 accumulate, BOTH operands MN-major (bits 15, 16); the operand format
// bits (ptx::umma_fmt_bits: bf16 or fp16, the same for both operands) are OR-ed in at run time
__host__ __device__ constexpr uint32_t idesc_mn(int M, int N) {
    return (1u << 4) | (1u << 15) | (1u << 16) |
           (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                       uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(d_tmem),
        "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

template <int KH, int CDY>
__global__ void __launch_bounds__(NUM_THREADS, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmD, const WgradParams p) {
    pdl_prologue();
    using C = WCfg<KH, CDY>;
    constexpr int NS = C::NS;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NS * C::STAGE + C::SLACK);
    uint64_t* full = bars;              // [NS]
    uint64_t* empty = full + NS;        // [NS]
    uint64_t* done = empty + NS;        // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int pad = KH / 2;
    // work item
    const int split = blockIdx.x % p.splits;
    int item = blockIdx.x / p.splits;
    const int chunk = item % p.chunks;
    const int s = item / p.chunks;                    // horizontal tap (KW = KH or 1: the host sizes the grid)
    const int t0 = split * p.tiles_per_split;
    const int t1 = min(p.tiles, t0 + p.tiles_per_split);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmX); ptx::prefetch_tmap(&tmD);
        for (int i = 0; i < NS; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
        ptx::mbar_init(done, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) { ptx::tmem_alloc(tmem_slot, C::TMEM_COLS); ptx::tmem_relinquish(); }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int st = 0; uint32_t ph = 0;
            for (int t = t0; t < t1; ++t) {
                int q = t; const int tw = q % p.tiles_w; q /= p.tiles_w;
                const int th = q % p.tiles_h; const int n = q / p.tiles_h;
                const int h0 = th * TILE_H, w0 = tw * TILE_W;
                ptx::mbar_wait(&empty[st], ph ^ 1);
                ptx::mbar_expect_tx(&full[st], C::STAGE);
                uint8_t* dst = smem + st * C::STAGE;
#pragma unroll
                for (int pl = 0; pl < PL; ++pl) {
                    // planes of one tensor are equally spaced: plane index rides on the batch coordinate
                    ptx::tma_load_4d(dst + pl * C::A_PLANE, &tmX, &full[st], chunk * 64, w0 + s - p.pad_w, h0 - pad, n + pl * p.N);
                }
                uint8_t* db = dst + PL * C::A_PLANE;
#pragma unroll
                for (int pl = 0; pl < PL; ++pl)
#pragma unroll
                    for (int b = 0; b < C::NBOX; ++b)
                        ptx::tma_load_4d(db + pl * C::B_PLANE + b * C::B_BOX, &tmD, &full[st], b * 64, w0, h0, n + pl * p.N);
                if (++st == NS) { st = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const bool leader = ptx::elect_one();
        const uint32_t idesc = idesc_mn(128, CDY) | p.fmt_bits;
        constexpr uint32_t a_hi = mn_desc_hi(128);
        constexpr uint32_t b_hi = mn_desc_hi(C::DROW);
        int st = 0; uint32_t ph = 0;
        uint32_t accumulate = 0;
        for (int t = t0; t < t1; ++t) {
            ptx::mbar_wait(&full[st], ph);
            ptx::tc_fence_after();
            const uint32_t a0 = ptx::smem_u32(smem + st * C::STAGE);
            const uint32_t b0 = a0 + PL * C::A_PLANE;
#pragma unroll
            for (int g = 0; g < C::NACC; ++g) {
#pragma unroll
                for (int k = 0; k < TILE_PIX / 16; ++k) {
                    // a*b = a0*b0 + a0*b1 + a1*b0  (planes)
#pragma unroll
                    for (int pa = 0; pa < PL; ++pa) {
#pragma unroll
                        for (int pb = 0; pb < PL; ++pb) {
                            if (pa + pb > 1) continue;
                            const uint32_t a_lo = mn_desc_lo(a0 + pa * C::A_PLANE + g * 2 * TILE_W * 128 + k * 16 * 128, TILE_W * 128);
                            const uint32_t b_lo = mn_desc_lo(b0 + pb * C::B_PLANE + k * 16 * C::DROW, C::B_BOX);
                            if (leader) mma_ss(tmem_base + g * CDY, a_lo, a_hi, b_lo, b_hi, idesc, (k | pa | pb) ? 1u : accumulate);
                        }
                    }
                }
            }
            accumulate = 1;
            if (leader) ptx::tc_commit(&empty[st]);
            __syncwarp();
            if (++st == NS) { st = 0; ph ^= 1; }
        }
        if (leader) ptx::tc_commit(done);
        __syncwarp();
    } else {
        // ===================== epilogue (warps 2..5): TMEM -> fp32 partials =====================
        const int q = warp & 3;
        const int m = q * 32 + lane;                  // accumulator row: (tap within the pair) * 64 + channel
        ptx::mbar_wait(done, 0);
        ptx::tc_fence_after();
        const bool any = t1 > t0;
        float* out = p.partial + (static_cast<size_t>(blockIdx.x) * C::NACC) * 128 * CDY;
#pragma unroll 1
        for (int g = 0; g < C::NACC; ++g) {
#pragma unroll 1
            for (int j = 0; j < CDY / 16; ++j) {
                uint32_t r[16];
                ptx::tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + g * CDY + j * 16, r);
                ptx::tmem_ld_wait();
                float4* dst = reinterpret_cast<float4*>(out + (static_cast<size_t>(g) * 128 + m) * CDY + j * 16);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    dst[i] = any ? make_float4(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]),
                                               __uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        ptx::tc_fence_before();
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
}

// dw[co][ci][r][s] = sum over splits of partial[split][s][chunk][acc = r / 2][(r % 2) * 64 + ci % 64][co]
// block = 32 output channels x 8 consecutive j = (ci, r, s): the partials are read with co fastest
// (coalesced), the tile is transposed through shared memory and written with j fastest.
__global__ void __launch_bounds__(256)
wgrad_tc_finish_kernel(const float* __restrict__ partial, int splits, int KW, int chunks, int nacc, int CDY,
                       int Cout, int Cin, int KH, const float* __restrict__ dy_slot, float* __restrict__ dw) {
    pdl_prologue();
    __shared__ float tile[8][33];
    const int J = Cin * KH * KW;
    const int co0 = blockIdx.y * 32, j0 = blockIdx.x * 8;
    const int cl = threadIdx.x & 31, jl = threadIdx.x >> 5;
    const size_t per_cta = static_cast<size_t>(nacc) * 128 * CDY;
    {
        const int j = j0 + jl, co = co0 + cl;
        float a0 = 0.f, a1 = 0.f;
        if (j < J && co < Cout) {
            int q = j; const int s = q % KW; q /= KW; const int r = q % KH; const int ci = q / KH;
            const int chunk = ci / 64;
            const size_t within = (static_cast<size_t>(r / 2) * 128 + (r % 2) * 64 + (ci % 64)) * CDY + co;
            const float* base = partial + static_cast<size_t>(s * chunks + chunk) * splits * per_cta + within;
            float a2 = 0.f, a3 = 0.f;
            int sp = 0;
            for (; sp + 3 < splits; sp += 4) {
                a0 += base[static_cast<size_t>(sp) * per_cta];     a1 += base[static_cast<size_t>(sp + 1) * per_cta];
                a2 += base[static_cast<size_t>(sp + 2) * per_cta]; a3 += base[static_cast<size_t>(sp + 3) * per_cta];
            }
            for (; sp < splits; ++sp) a0 += base[static_cast<size_t>(sp) * per_cta];
            a0 += a2; a1 += a3;
        }
        tile[jl][cl] = a0 + a1;
    }
    __syncthreads();
    {
        const int co = co0 + (threadIdx.x >> 3), j = j0 + (threadIdx.x & 7);
        if (co < Cout && j < J) dw[static_cast<size_t>(co) * J + j] = tile[threadIdx.x & 7][threadIdx.x >> 3] * slot_inv(dy_slot);
    }
}

// ---------------------------------------------------------------------------------------------
// Gram matrix  G[b] = F_b^T F_b  (StyleLoss.gram_matrix, network.py:92-108) on the same MN-major
// scheme: F is [B][HW][C] pixel-major, a tile is TP consecutive pixels x C channels (C/64 TMA boxes
// per plane), M = 128 = two 64-channel groups (LBO = one box), N = C, and the SAME shared-memory
// tile is both operands.  One CTA = (image, pixel split); fp32 partials, deterministic finish.
// ---------------------------------------------------------------------------------------------
struct GramParams {
    int B, HW, tiles, splits, tiles_per_split;
    uint32_t fmt_bits;
    float* partial;                         // [b][split][acc][128][C]
};

template <int C_> struct GCfg {
    static constexpr int TP = C_ > 128 ? 64 : 128;                // pixels per tile
    static constexpr int NBOX = C_ / 64;
    static constexpr int BOX = TP * 128;
    static constexpr int PLANE = NBOX * BOX;
    static constexpr int STAGE = PL * PLANE;
    static constexpr int NACC = C_ > 128 ? C_ / 128 : 1;
    // C <= 128: ONE wide MMA per k-step, hi^T x [hi | lo] (the planes of a stage are adjacent 64-channel boxes, so N = 2C
    // is one operand with the box stride as leading dimension): accumulators [Hh | X] with X = hi^T lo.  The third product
    // lo^T hi is X^T and is added by the epilogue -- 8 KB (12 KB at C = 128) of shared-memory operand reads per k-step
    // instead of 18 (24) KB for three N = C MMAs; the kernel was bound by that, not by HBM.
    static constexpr bool WIDE = C_ <= 128;
    static constexpr int TMEM_NEED = WIDE ? 2 * C_ : C_;          // C = 256: one 128-row half per CTA
    static constexpr int TMEM_COLS = TMEM_NEED <= 64 ? 64 : TMEM_NEED <= 128 ? 128 : TMEM_NEED <= 256 ? 256 : 512;
    static constexpr int SLACK = C_ == 64 ? BOX : 0;              // C = 64: the second M group is don't-care data
    static constexpr int NS_FIT = (232448 - 1024 - 256 - SLACK) / STAGE;
    static constexpr int NS = NS_FIT > 4 ? 4 : NS_FIT;
    static_assert(NS >= 2, "need two pipeline stages");
    static constexpr int SMEM = 1024 + NS * STAGE + SLACK + 256;
};

template <int C_>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gram_tc_kernel(const __grid_constant__ CUtensorMap tmF, const GramParams p) {
    pdl_prologue();
    using G = GCfg<C_>;
    constexpr int NS = G::NS;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NS * G::STAGE + G::SLACK);
    uint64_t* full = bars; uint64_t* empty = full + NS; uint64_t* done = empty + NS;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // C > 128: blockIdx = ((b * splits) + split) * NACC + g, this CTA accumulates rows [g*128, g*128 + 128) of G only
    const int gsel = blockIdx.x % G::NACC, cta = blockIdx.x / G::NACC;
    const int split = cta % p.splits, b = cta / p.splits;
    const int t0 = split * p.tiles_per_split, t1 = min(p.tiles, t0 + p.tiles_per_split);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmF);
        for (int i = 0; i < NS; ++i) { ptx::mbar_init(&full[i], 1); ptx::mbar_init(&empty[i], 1); }
        ptx::mbar_init(done, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) { ptx::tmem_alloc(tmem_slot, G::TMEM_COLS); ptx::tmem_relinquish(); }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int st = 0; uint32_t ph = 0;
            for (int t = t0; t < t1; ++t) {
                ptx::mbar_wait(&empty[st], ph ^ 1);
                ptx::mbar_expect_tx(&full[st], G::STAGE);
                uint8_t* dst = smem + st * G::STAGE;
#pragma unroll
                for (int pl = 0; pl < PL; ++pl)
#pragma unroll
                    for (int bx = 0; bx < G::NBOX; ++bx)
                        ptx::tma_load_3d(dst + pl * G::PLANE + bx * G::BOX, &tmF, &full[st], bx * 64, t * G::TP, b + pl * p.B);
                if (++st == NS) { st = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        const bool leader = ptx::elect_one();
        const uint32_t idesc = idesc_mn(128, C_) | p.fmt_bits;
        constexpr uint32_t hi = mn_desc_hi(128);
        int st = 0; uint32_t ph = 0;
        uint32_t accumulate = 0;
        for (int t = t0; t < t1; ++t) {
            ptx::mbar_wait(&full[st], ph);
            ptx::tc_fence_after();
            const uint32_t base = ptx::smem_u32(smem + st * G::STAGE);
            if constexpr (G::WIDE) {
                constexpr uint32_t idesc_w = idesc_mn(128, 2 * C_);
#pragma unroll
                for (int k = 0; k < G::TP / 16; ++k) {
                    const uint32_t a_lo = mn_desc_lo(base + k * 16 * 128, G::BOX);
                    if (leader) mma_ss(tmem_base, a_lo, hi, a_lo, hi, idesc_w | p.fmt_bits, k ? 1u : accumulate);
                }
            } else {
                const int g = gsel;
#pragma unroll
                for (int k = 0; k < G::TP / 16; ++k) {
#pragma unroll
                    for (int pa = 0; pa < PL; ++pa) {
#pragma unroll
                        for (int pb = 0; pb < PL; ++pb) {
                            if (pa + pb > 1) continue;
                            const uint32_t a_lo = mn_desc_lo(base + pa * G::PLANE + g * 2 * G::BOX + k * 16 * 128, G::BOX);
                            const uint32_t b_lo = mn_desc_lo(base + pb * G::PLANE + k * 16 * 128, G::BOX);
                            if (leader) mma_ss(tmem_base, a_lo, hi, b_lo, hi, idesc, (k | pa | pb) ? 1u : accumulate);
                        }
                    }
                }
            }
            accumulate = 1;
            if (leader) ptx::tc_commit(&empty[st]);
            __syncwarp();
            if (++st == NS) { st = 0; ph ^= 1; }
        }
        if (leader) ptx::tc_commit(done);
        __syncwarp();
    } else {
        const int q = warp & 3;
        const int m = q * 32 + lane;
        ptx::mbar_wait(done, 0);
        ptx::tc_fence_after();
        const bool any = t1 > t0;
        // partial layout [b][split][half][128][C]: blockIdx.x enumerates (b, split, half) in exactly that order
        float* out = p.partial + static_cast<size_t>(blockIdx.x) * 128 * C_;
        if constexpr (G::WIDE) {
            // X -> shared memory (the pipeline stages are free now), row stride C + 1 floats; then out = Hh + X + X^T
            float* xs = reinterpret_cast<float*>(smem);
            constexpr int LD = C_ + 1;
            static_assert(LD * C_ * 4 <= NS * G::STAGE, "transpose buffer does not fit the stages");
            if (m < C_ && any) {
#pragma unroll 1
                for (int j = 0; j < C_ / 16; ++j) {
                    uint32_t r[16];
                    ptx::tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + C_ + j * 16, r);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) xs[m * LD + j * 16 + i] = __uint_as_float(r[i]);
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (m < C_) {
#pragma unroll 1
                for (int j = 0; j < C_ / 16; ++j) {
                    uint32_t r[16], x[16];
                    ptx::tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + j * 16, r);
                    ptx::tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + C_ + j * 16, x);
                    ptx::tmem_ld_wait();
                    float4* dst = reinterpret_cast<float4*>(out + static_cast<size_t>(m) * C_ + j * 16);
                    float v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = any ? __uint_as_float(r[i]) + (__uint_as_float(x[i]) + xs[(j * 16 + i) * LD + m]) : 0.f;
#pragma unroll
                    for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                }
            }
        } else if (m < C_) {                                         // C = 256 (half per CTA)
#pragma unroll 1
            for (int j = 0; j < C_ / 16; ++j) {
                uint32_t r[16];
                ptx::tmem_ld16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + j * 16, r);
                ptx::tmem_ld_wait();
                float4* dst = reinterpret_cast<float4*>(out + static_cast<size_t>(m) * C_ + j * 16);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    dst[i] = any ? make_float4(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]),
                                               __uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        ptx::tc_fence_before();
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, G::TMEM_COLS);
}

// G[b][i][j] = scale * sum over splits of partial[b][split][i / 128][i % 128][j]
__global__ void __launch_bounds__(256)
gram_tc_finish_kernel(const float* __restrict__ partial, int splits, int nacc, int C, float scale, float* __restrict__ Gm, int total) {
    pdl_prologue();
    __shared__ float sm[8][33];
    const int ol = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int idx = blockIdx.x * 32 + ol;
    float s = 0.f;
    if (idx < total) {
        int q = idx; const int j = q % C; q /= C; const int i = q % C; const int b = q / C;
        const size_t per = static_cast<size_t>(nacc) * 128 * C;
        const float* base = partial + (static_cast<size_t>(b) * splits) * per + (static_cast<size_t>(i / 128) * 128 + (i % 128)) * C + j;
        // four independent accumulators: the ~18 dependent L2 round trips of a 148-way split become ~5
        float s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int sp = sl;
        for (; sp + 24 < splits; sp += 32) {
            s += base[static_cast<size_t>(sp) * per];       s1 += base[static_cast<size_t>(sp + 8) * per];
            s2 += base[static_cast<size_t>(sp + 16) * per]; s3 += base[static_cast<size_t>(sp + 24) * per];
        }
        for (; sp < splits; sp += 8) s += base[static_cast<size_t>(sp) * per];
        s = (s + s1) + (s2 + s3);
    }
    sm[sl][ol] = s;
    __syncthreads();
    if (sl == 0 && idx < total) {
        float t = 0.f;
#pragma unroll
        for (int l = 0; l < 8; ++l) t += sm[l][ol];
        Gm[idx] = t * scale;
    }
}

// the same sum for a SMALL split count (batched training: 148 / B splits per image): one thread per 4 adjacent
// outputs walks the splits itself -- the 8-lane layout above would leave most lanes of its 32768 blocks idle
__global__ void __launch_bounds__(256)
gram_tc_finish_small_kernel(const float* __restrict__ partial, int splits, int nacc, int C, float scale, float* __restrict__ Gm, int total4) {
    pdl_prologue();
    const int v = blockIdx.x * 256 + threadIdx.x;
    if (v >= total4) return;
    int q = v * 4; const int j = q % C; q /= C; const int i = q % C; const int b = q / C;
    const size_t per = static_cast<size_t>(nacc) * 128 * C;
    const float* base = partial + (static_cast<size_t>(b) * splits) * per + (static_cast<size_t>(i / 128) * 128 + (i % 128)) * C + j;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    int sp = 0;
    for (; sp + 1 < splits; sp += 2) {
        const float4 x0 = *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * per);
        const float4 x1 = *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp + 1) * per);
        a0.x += x0.x; a0.y += x0.y; a0.z += x0.z; a0.w += x0.w;
        a1.x += x1.x; a1.y += x1.y; a1.z += x1.z; a1.w += x1.w;
    }
    if (sp < splits) {
        const float4 x0 = *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * per);
        a0.x += x0.x; a0.y += x0.y; a0.z += x0.z; a0.w += x0.w;
    }
    *reinterpret_cast<float4*>(Gm + static_cast<size_t>(v) * 4) =
        make_float4((a0.x + a1.x) * scale, (a0.y + a1.y) * scale, (a0.z + a1.z) * scale, (a0.w + a1.w) * scale);
}

// ---- Gram finish fused with StyleLoss.forward (network.py:110-123) -------------------------------------------------------
// diff = G - T (T [C][C] broadcast over images), loss = mean(diff^2), amax = max |diff| -- in the launch that reduces
// the split partials, instead of three more launches per style node (finish -> diff -> sum).  Every block leaves its
// sum of squares and its maximum in `part`; the block that draws the last ticket adds them up in a fixed order
// (deterministic) and hands the ticket counter back at zero.
struct StyleTail {
    const float* T; float* diff; float* Gm;          // Gm optional
    float* part; int* ticket; float* loss; float* amax; float inv_total;
};
__device__ __forceinline__ void style_tail_finish(const StyleTail& t, float s, float m, float* sm) {
    s = block_sum(s, sm);
    m = block_max(m, sm);
    __shared__ int last;
    if (threadIdx.x == 0) {
        t.part[blockIdx.x] = s; t.part[gridDim.x + blockIdx.x] = m;
        __threadfence();
        last = atomicAdd(t.ticket, 1) == static_cast<int>(gridDim.x) - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // eight independent load pairs per round (a rolled loop is one L2 round trip per 256 partials: 8 us at 2048 blocks)
    float ss = 0.f, mm = 0.f;
    const int nb = static_cast<int>(gridDim.x), step = static_cast<int>(blockDim.x);
    for (int i0 = threadIdx.x; i0 < nb; i0 += 8 * step) {
        float a[8], b[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int i = i0 + j * step;
            a[j] = i < nb ? __ldcg(t.part + i) : 0.f;
            b[j] = i < nb ? __ldcg(t.part + nb + i) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { ss += a[j]; mm = fmaxf(mm, b[j]); }
    }
    ss = block_sum(ss, sm);
    mm = block_max(mm, sm);
    if (threadIdx.x == 0) { t.loss[0] = ss * t.inv_total; if (t.amax) t.amax[0] = mm; *t.ticket = 0; }
}

__global__ void __launch_bounds__(256)
gram_tc_finish_style_kernel(const float* __restrict__ partial, int splits, int nacc, int C, float scale, int total, StyleTail t) {
    pdl_prologue();
    __shared__ float sm[8][33];
    __shared__ float red[32];
    const int ol = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int idx = blockIdx.x * 32 + ol;
    float s = 0.f;
    if (idx < total) {
        int q = idx; const int j = q % C; q /= C; const int i = q % C; const int b = q / C;
        const size_t per = static_cast<size_t>(nacc) * 128 * C;
        const float* base = partial + (static_cast<size_t>(b) * splits) * per + (static_cast<size_t>(i / 128) * 128 + (i % 128)) * C + j;
        float s1 = 0.f, s2 = 0.f, s3 = 0.f;
        if (splits <= 160) {
            // every partial of this lane in flight at once (<= 20 loads), summed in the order of the loop below: the loop is
            // one L2 round trip per four partials, five in a row for the 147-way split of a single image
            float v[20];
#pragma unroll
            for (int i = 0; i < 20; ++i) {
                const int sp = sl + 8 * i;
                v[i] = sp < splits ? __ldcg(base + static_cast<size_t>(sp) * per) : 0.f;
            }
#pragma unroll
            for (int i = 0; i < 20; i += 4) {
                // the loop adds four per round only while sp + 24 < splits, the remainder goes to s alone
                const int sp = sl + 8 * i;
                if (sp + 24 < splits) { s += v[i]; s1 += v[i + 1]; s2 += v[i + 2]; s3 += v[i + 3]; }
                else { s += v[i]; s += v[i + 1]; s += v[i + 2]; s += v[i + 3]; }
            }
        } else {
        int sp = sl;
        for (; sp + 24 < splits; sp += 32) {
            s += base[static_cast<size_t>(sp) * per];       s1 += base[static_cast<size_t>(sp + 8) * per];
            s2 += base[static_cast<size_t>(sp + 16) * per]; s3 += base[static_cast<size_t>(sp + 24) * per];
        }
        for (; sp < splits; sp += 8) s += base[static_cast<size_t>(sp) * per];
        }
        s = (s + s1) + (s2 + s3);
    }
    sm[sl][ol] = s;
    __syncthreads();
    float d2 = 0.f, am = 0.f;
    if (sl == 0 && idx < total) {
        float g = 0.f;
#pragma unroll
        for (int l = 0; l < 8; ++l) g += sm[l][ol];
        g *= scale;
        if (t.Gm) t.Gm[idx] = g;
        const float d = g - t.T[idx % (C * C)];
        t.diff[idx] = d;
        d2 = d * d; am = fabsf(d);
    }
    style_tail_finish(t, d2, am, red);
}

// Same reduction with FOUR consecutive outputs per thread (one 16-byte load per split): a quarter of the blocks for the
// same bytes -- at C = 256 the scalar kernel is 2048 blocks whose per-block tail (two block reductions, a ticket) and
// 2048-entry last-block reduction cost more than the partials they read.  Used when an image has more than 16 splits
// and C >= 128; summation order per output as in gram_tc_finish_style_kernel.
__global__ void __launch_bounds__(256)
gram_tc_finish_style4_kernel(const float* __restrict__ partial, int splits, int nacc, int C, float scale, int total4, StyleTail t) {
    pdl_prologue();
    __shared__ float4 sm[8][33];
    __shared__ float red[32];
    const int ol = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int v4 = blockIdx.x * 32 + ol;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    auto add4 = [](float4& a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; };
    if (v4 < total4) {
        int q = v4 * 4; const int j = q % C; q /= C; const int i = q % C; const int b = q / C;
        const size_t per = static_cast<size_t>(nacc) * 128 * C;
        const float* base = partial + (static_cast<size_t>(b) * splits) * per + (static_cast<size_t>(i / 128) * 128 + (i % 128)) * C + j;
        float4 s1 = s, s2 = s, s3 = s;
        if (splits <= 160) {
            float4 v[20];
#pragma unroll
            for (int k = 0; k < 20; ++k) {
                const int sp = sl + 8 * k;
                v[k] = sp < splits ? __ldcg(reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * per)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int k = 0; k < 20; k += 4) {
                const int sp = sl + 8 * k;
                if (sp + 24 < splits) { add4(s, v[k]); add4(s1, v[k + 1]); add4(s2, v[k + 2]); add4(s3, v[k + 3]); }
                else { add4(s, v[k]); add4(s, v[k + 1]); add4(s, v[k + 2]); add4(s, v[k + 3]); }
            }
        } else {
            int sp = sl;
            for (; sp + 24 < splits; sp += 32) {
                add4(s, *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * per));
                add4(s1, *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp + 8) * per));
                add4(s2, *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp + 16) * per));
                add4(s3, *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp + 24) * per));
            }
            for (; sp < splits; sp += 8) add4(s, *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * per));
        }
        s = make_float4((s.x + s1.x) + (s2.x + s3.x), (s.y + s1.y) + (s2.y + s3.y), (s.z + s1.z) + (s2.z + s3.z), (s.w + s1.w) + (s2.w + s3.w));
    }
    sm[sl][ol] = s;
    __syncthreads();
    float d2 = 0.f, am = 0.f;
    if (sl == 0 && v4 < total4) {
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int l = 0; l < 8; ++l) add4(g, sm[l][ol]);
        g = make_float4(g.x * scale, g.y * scale, g.z * scale, g.w * scale);
        const size_t idx = static_cast<size_t>(v4) * 4;
        if (t.Gm) *reinterpret_cast<float4*>(t.Gm + idx) = g;
        const float4 tt = *reinterpret_cast<const float4*>(t.T + idx % (static_cast<size_t>(C) * C));
        const float4 d = make_float4(g.x - tt.x, g.y - tt.y, g.z - tt.z, g.w - tt.w);
        *reinterpret_cast<float4*>(t.diff + idx) = d;
        d2 = (d.x * d.x + d.y * d.y) + (d.z * d.z + d.w * d.w);
        am = fmaxf(fmaxf(fabsf(d.x), fabsf(d.y)), fmaxf(fabsf(d.z), fabsf(d.w)));
    }
    style_tail_finish(t, d2, am, red);
}

__global__ void __launch_bounds__(256)
gram_tc_finish_small_style_kernel(const float* __restrict__ partial, int splits, int nacc, int C, float scale, int total4, StyleTail t) {
    pdl_prologue();
    __shared__ float red[32];
    const int v = blockIdx.x * 256 + threadIdx.x;
    float d2 = 0.f, am = 0.f;
    if (v < total4) {
        int q = v * 4; const int j = q % C; q /= C; const int i = q % C; const int b = q / C;
        const size_t per = static_cast<size_t>(nacc) * 128 * C;
        const float* base = partial + (static_cast<size_t>(b) * splits) * per + (static_cast<size_t>(i / 128) * 128 + (i % 128)) * C + j;
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
        int sp = 0;
        for (; sp + 1 < splits; sp += 2) {
            const float4 x0 = *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * per);
            const float4 x1 = *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp + 1) * per);
            a0.x += x0.x; a0.y += x0.y; a0.z += x0.z; a0.w += x0.w;
            a1.x += x1.x; a1.y += x1.y; a1.z += x1.z; a1.w += x1.w;
        }
        if (sp < splits) {
            const float4 x0 = *reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * per);
            a0.x += x0.x; a0.y += x0.y; a0.z += x0.z; a0.w += x0.w;
        }
        const float4 g = make_float4((a0.x + a1.x) * scale, (a0.y + a1.y) * scale, (a0.z + a1.z) * scale, (a0.w + a1.w) * scale);
        if (t.Gm) *reinterpret_cast<float4*>(t.Gm + static_cast<size_t>(v) * 4) = g;
        const float4 tt = *reinterpret_cast<const float4*>(t.T + (static_cast<size_t>(v) * 4) % (static_cast<size_t>(C) * C));
        const float4 d = make_float4(g.x - tt.x, g.y - tt.y, g.z - tt.z, g.w - tt.w);
        *reinterpret_cast<float4*>(t.diff + static_cast<size_t>(v) * 4) = d;
        d2 = (d.x * d.x + d.y * d.y) + (d.z * d.z + d.w * d.w);
        am = fmaxf(fmaxf(fabsf(d.x), fabsf(d.y)), fmaxf(fabsf(d.z), fabsf(d.w)));
    }
    style_tail_finish(t, d2, am, red);
}

struct GramPlan { int tiles, splits, tiles_per_split, grid, nacc; };
GramPlan plan_gram(int B, int HW, int C) {
    GramPlan q;
    const int tp = C > 128 ? 64 : 128;
    q.tiles = (HW + tp - 1) / tp;
    // C = 256: a CTA owns ONE 128-row half of G (grid = B x splits x 2), so a 148-SM GPU runs 74 pixel splits x 2 halves:
    // the fp32 partials (and what the finish launch reads back) halve -- 19 instead of 39 MB for a 128 x 128 tap
    const int halves = C > 128 ? C / 128 : 1;
    int splits = st_num_sms() / (B * halves); if (splits < 1) splits = 1; if (splits > q.tiles) splits = q.tiles;
    // Tuning override only.  Measured (tools/bench_gram.py, profiles/r01n_gram_split_sweep.txt): capping the split
    // to cut the C x C partial traffic (39 MB at C = 256) makes every tap SLOWER -- the kernel is bound by the
    // per-CTA tile loop (MMA issue + TMEM epilogue), not by the partials -- so the automatic plan uses every SM.
    if (g_gram_max_splits > 0 && splits > g_gram_max_splits) splits = g_gram_max_splits;
    q.tiles_per_split = (q.tiles + splits - 1) / splits;
    q.splits = (q.tiles + q.tiles_per_split - 1) / q.tiles_per_split;
    q.grid = B * q.splits * halves; q.nacc = halves;
    return q;
}

template <int C_>
int launch_gram(const CUtensorMap& tf, const GramParams& p, int grid, cudaStream_t stream) {
    using G = GCfg<C_>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gram_tc_kernel<C_>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM);
        if (e != cudaSuccess) { st_set_error("gram_tc smem attribute: %s", cudaGetErrorString(e)); return ST_ERR_CUDA; }
        attr_set = true;
    }
    st_launch(grid, NUM_THREADS, G::SMEM, stream)(gram_tc_kernel<C_>, tf, p);
    ST_CHECK_LAUNCH("gram_tc_kernel");
    return ST_OK;
}

template <int KH, int CDY>
int launch_wgrad(const CUtensorMap& tx, const CUtensorMap& td, const WgradParams& p, int grid, cudaStream_t stream) {
    using C = WCfg<KH, CDY>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(wgrad_tc_kernel<KH, CDY>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { st_set_error("wgrad_tc smem attribute: %s", cudaGetErrorString(e)); return ST_ERR_CUDA; }
        attr_set = true;
    }
    st_launch(grid, NUM_THREADS, C::SMEM, stream)(wgrad_tc_kernel<KH, CDY>, tx, td, p);
    ST_CHECK_LAUNCH("wgrad_tc_kernel");
    return ST_OK;
}

struct WgradPlan { int tiles, chunks, kw, items, splits, tiles_per_split, grid, nacc; };
WgradPlan plan_wgrad(int N, int H, int W, int CinPad, int kh, int kw) {
    WgradPlan q;
    q.tiles = N * ((H + TILE_H - 1) / TILE_H) * ((W + TILE_W - 1) / TILE_W);
    q.chunks = CinPad / 64; q.kw = kw; q.items = kw * q.chunks;
    int splits = st_num_sms() / q.items; if (splits < 1) splits = 1; if (splits > q.tiles) splits = q.tiles;
    q.tiles_per_split = (q.tiles + splits - 1) / splits;
    q.splits = (q.tiles + q.tiles_per_split - 1) / q.tiles_per_split;
    q.grid = q.items * q.splits; q.nacc = (kh + 1) / 2;
    return q;
}

}  // namespace

bool gram_tc_eligible(int C, bool two_planes) { return (C == 64 || C == 128 || C == 256) && two_planes; }
size_t gram_tc_workspace_bytes(int B, int HW, int C) {
    const GramPlan q = plan_gram(B, HW, C);
    return static_cast<size_t>(q.grid) * 128 * C * sizeof(float);          // one 128-row block of partials per CTA
}
// blocks of the finish launch (= entries of each half of the style partials)
int gram_tc_finish_blocks(int B, int HW, int C) {
    const GramPlan q = plan_gram(B, HW, C);
    const int total = B * C * C;
    return q.splits <= 16 ? (total / 4 + 255) / 256 : (total + 31) / 32;
}
// f [planes][B][HW][C] (plane stride B*HW*C elements; f16: the planes are fp16) -> Gm [B][C][C] * scale.
// With T != null the finish launch also evaluates StyleLoss.forward: diff = G - T, loss = mean(diff^2), amax = max|diff|
// (Gm then optional; style_part: 2 * gram_tc_finish_blocks floats; ticket: int, zero before the first call).
int gram_tc_run(const bf16* f_hi, bool f16, int B, int HW, int C, float scale, float* Gm, void* workspace, size_t workspace_bytes,
                const float* T, float* diff, float* loss, float* amax, float* style_part, int* ticket, cudaStream_t stream) {
    const GramPlan q = plan_gram(B, HW, C);
    const size_t need = static_cast<size_t>(q.grid) * 128 * C * sizeof(float);
    ST_CHECK_ARG(workspace && workspace_bytes >= need, "gram_tc: workspace too small (%zu < %zu)", workspace_bytes, need);
    GramParams p;
    p.B = B; p.HW = HW; p.tiles = q.tiles; p.splits = q.splits; p.tiles_per_split = q.tiles_per_split;
    p.fmt_bits = ptx::umma_fmt_bits(f16);
    p.partial = static_cast<float*>(workspace);
    CUtensorMap tf;
    uint64_t fd[3] = {(uint64_t)C, (uint64_t)HW, (uint64_t)B * PL};
    uint32_t fb[3] = {64, (uint32_t)(C > 128 ? 64 : 128), 1};
    int rc;
    if ((rc = st_make_tmap(&tf, f_hi, 3, fd, fb, 128))) return rc;
    if (C == 64) rc = launch_gram<64>(tf, p, q.grid, stream);
    else if (C == 128) rc = launch_gram<128>(tf, p, q.grid, stream);
    else rc = launch_gram<256>(tf, p, q.grid, stream);
    if (rc) return rc;
    const int total = B * C * C;
    if (T) {
        StyleTail t;
        t.T = T; t.diff = diff; t.Gm = Gm; t.part = style_part; t.ticket = ticket; t.loss = loss; t.amax = amax;
        t.inv_total = 1.0f / static_cast<float>(total);
        if (q.splits <= 16) {
            st_launch((total / 4 + 255) / 256, 256, 0, stream)(gram_tc_finish_small_style_kernel, p.partial, q.splits, q.nacc, C, scale, total / 4, t);
            ST_CHECK_LAUNCH("gram_tc_finish_small_style_kernel");
        } else {
            if (C >= 128) {
                st_launch((total / 4 + 31) / 32, 256, 0, stream)(gram_tc_finish_style4_kernel, p.partial, q.splits, q.nacc, C, scale, total / 4, t);
                ST_CHECK_LAUNCH("gram_tc_finish_style4_kernel");
            } else {
                st_launch((total + 31) / 32, 256, 0, stream)(gram_tc_finish_style_kernel, p.partial, q.splits, q.nacc, C, scale, total, t);
                ST_CHECK_LAUNCH("gram_tc_finish_style_kernel");
            }
        }
        return ST_OK;
    }
    if (q.splits <= 16) {
        st_launch((total / 4 + 255) / 256, 256, 0, stream)(gram_tc_finish_small_kernel, p.partial, q.splits, q.nacc, C, scale, Gm, total / 4);
        ST_CHECK_LAUNCH("gram_tc_finish_small_kernel");
    } else {
        st_launch((total + 31) / 32, 256, 0, stream)(gram_tc_finish_kernel, p.partial, q.splits, q.nacc, C, scale, Gm, total);
        ST_CHECK_LAUNCH("gram_tc_finish_kernel");
    }
    return ST_OK;
}

bool wgrad_tc_eligible(int ks, int stride, int pad, int up, int CinPad, int Cdy, bool two_planes) {
    return ks == 3 && stride == 1 && pad == 1 && !up && CinPad % 64 == 0 && (Cdy == 32 || Cdy == 64 || Cdy == 128) && two_planes;
}
size_t wgrad_tc_workspace_bytes(int N, int H, int W, int CinPad, int Cdy, int ks) {
    const WgradPlan q = plan_wgrad(N, H, W, CinPad, ks, ks);
    return static_cast<size_t>(q.grid) * q.nacc * 128 * Cdy * sizeof(float);
}

// dy [planes][N][H][W][Cdy], x [planes][N][H][W][CinPad] (plane stride = N*H*W*C elements), dw [Cout][Cin][ks][ks].
// f16: both tensors hold fp16 planes; dy_slot (optional): the power-of-two scale slot of dy, undone by the finish kernel
int wgrad_tc_run(const bf16* dy_hi, const bf16* x_hi, bool f16, const float* dy_slot, int N, int H, int W, int CinPad, int Cdy, int ks,
                 int Cout, int Cin, float* dw, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    const WgradPlan q = plan_wgrad(N, H, W, CinPad, ks, ks);
    const size_t need = static_cast<size_t>(q.grid) * q.nacc * 128 * Cdy * sizeof(float);
    ST_CHECK_ARG(workspace && workspace_bytes >= need, "wgrad_tc: workspace too small (%zu < %zu)", workspace_bytes, need);
    WgradParams p;
    p.N = N; p.H = H; p.W = W; p.Cin = CinPad; p.Cdy = Cdy;
    p.tiles_h = (H + TILE_H - 1) / TILE_H; p.tiles_w = (W + TILE_W - 1) / TILE_W; p.tiles = q.tiles;
    p.chunks = q.chunks; p.splits = q.splits; p.tiles_per_split = q.tiles_per_split; p.pad_w = ks / 2;
    p.fmt_bits = ptx::umma_fmt_bits(f16);
    p.partial = static_cast<float*>(workspace);
    CUtensorMap tx, td;
    uint64_t xd[4] = {(uint64_t)CinPad, (uint64_t)W, (uint64_t)H, (uint64_t)N * PL};
    uint32_t xb[4] = {64, TILE_W, (uint32_t)(TILE_H + ks - 1), 1};
    uint64_t dd[4] = {(uint64_t)Cdy, (uint64_t)W, (uint64_t)H, (uint64_t)N * PL};
    uint32_t db[4] = {(uint32_t)(Cdy < 64 ? Cdy : 64), TILE_W, TILE_H, 1};
    int rc;
    if ((rc = st_make_tmap(&tx, x_hi, 4, xd, xb, 128))) return rc;
    if ((rc = st_make_tmap(&td, dy_hi, 4, dd, db, Cdy < 64 ? 64 : 128))) return rc;
    if (Cdy == 128) rc = launch_wgrad<3, 128>(tx, td, p, q.grid, stream);
    else if (Cdy == 64) rc = launch_wgrad<3, 64>(tx, td, p, q.grid, stream);
    else rc = launch_wgrad<3, 32>(tx, td, p, q.grid, stream);
    if (rc) return rc;
    const int J = Cin * ks * ks;
    st_launch(dim3((J + 7) / 8, (Cout + 31) / 32), 256, 0, stream)(wgrad_tc_finish_kernel, p.partial, q.splits, ks, q.chunks, q.nacc, Cdy, Cout, Cin, ks, dy_slot, dw);
    ST_CHECK_LAUNCH("wgrad_tc_finish_kernel");
    return ST_OK;
}
