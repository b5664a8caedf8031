// Backward-weight of the transform net's 9x9 layers (network.py:525-530 / :850-855 and :604-609)
// on the tensor cores:
//
//   O[b][c][r][s] = sum over pixels p of  B[p][b] * S[c][p + (r - 4, s - 4)]       b < 32, c < 3
//
// as a GEMM  D[m][b] = sum_p A[m][p] * B[p][b]  with m = (c, r, s) (243 rows, two M = 128 blocks),
// N = 32 channels of the big tensor, K = pixels (blocks of 64 consecutive pixels of an image row).
//   B operand: the split-bf16 NHWC tensor itself, "MN-major" (channels contiguous), one TMA box of
//              64 pixels x 32 channels per plane, 64-byte swizzle.
//   A operand: does not exist in memory.  Row (c, r, s) of a K block is the image row
//              S[c][h + r - 4][w0 + s - 4 ...] -- 9 shifted copies per (c, r).  216 builder threads
//              (one per (c, r, 16-byte chunk)) read 16 fp32 pixels, split them into bf16 hi / lo and
//              store the 9 shifted windows K-major with the 128-byte swizzle a TMA load would
//              produce; fence.proxy.async + an mbarrier hand the tile to the MMA warp.
// Products: (a0 + a1)(b0 + b1) without a1 b1 as two MMAs per K step: a0 x [b0 | b1] (N = 64, the two
// planes are adjacent N groups) and a1 x b0 into the second half of the accumulator.
// Each CTA keeps its accumulators in TMEM over all its K blocks and writes one fp32 partial;
// conv9_wgrad_finish_kernel adds the partials in a fixed order.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.cuh"

namespace {

constexpr int T9 = 9, TPAD = 4, TCB = 32;
constexpr int KB = 64;                         // pixels per K block
constexpr int MROWS = 243;                     // (c, r, s) rows
constexpr int A_MBLOCK = 128 * 128;            // bytes of one M block of one plane (128 rows x 128 B)
constexpr int A_PLANE = 2 * A_MBLOCK;
constexpr int A_STAGE = 2 * A_PLANE;           // hi, lo
constexpr int B_PLANE = KB * 64;               // 64 pixels x 32 channels bf16
constexpr int B_STAGE = 2 * B_PLANE;
constexpr int STAGE9 = A_STAGE + B_STAGE;
constexpr int NS9 = 2;
constexpr int BUILD_WARPS = 7, BUILD_TASKS = 3 * T9 * 8;        // 216 = (c, r, chunk)
constexpr int BUILD_GROUPS = 2;                 // builder groups alternate K blocks; group g owns pipeline stage g
constexpr int THREADS9 = (2 + BUILD_GROUPS * BUILD_WARPS) * 32;
constexpr int SMEM9 = 1024 + NS9 * STAGE9 + 256;
constexpr int TMEM9 = 128;                      // 2 M blocks x (32 main + 32 correction) columns

struct C9Params {
    const float* s; int sc;                     // small tensor: channel c of sample n at s + (n*sc + c)*H*W
    int N, H, W, wblocks, work;
    float* partial; int part_stride;            // one partial per CTA: [32][cs_total][81]
    int cs_total, c_off;
    int f16;                                    // the big tensor holds fp16 planes: the builders make fp16 rows too
    const float* small_slot;                    // optional {s, 1/s}: the small tensor is multiplied by s before the fp16 split
};

// (operand format bits, ptx::umma_fmt_bits, are OR-ed in at run time)
__host__ __device__ constexpr uint32_t idesc_k_mn(int M, int N) {     // A K-major, B MN-major
    return (1u << 4) | (1u << 16) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
// 16 floats -> 16-bit patterns of the hi / lo planes (bf16, or fp16 of mul * v)
__device__ __forceinline__ void split16(const float (&v)[16], bool f16, float mul, uint32_t (&hi)[16], uint32_t (&lo)[16]) {
    if (f16) {
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
            uint32_t h2, l2; split2h(v[i] * mul, v[i + 1] * mul, h2, l2);
            hi[i] = h2 & 0xffffu; hi[i + 1] = h2 >> 16; lo[i] = l2 & 0xffffu; lo[i + 1] = l2 >> 16;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const bf16 h = __float2bfloat16_rn(v[i]);
            const bf16 l = __float2bfloat16_rn(v[i] - __bfloat162float(h));
            hi[i] = __bfloat16_as_ushort(h); lo[i] = __bfloat16_as_ushort(l);
        }
    }
}
__device__ __forceinline__ uint32_t mn_lo(uint32_t addr, uint32_t lbo_bytes) { return ((addr & 0x3ffff) >> 4) | ((lbo_bytes >> 4) << 16); }
__device__ __forceinline__ void mma2(uint32_t d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(acc)
        : "memory");
}

__global__ void __launch_bounds__(THREADS9, 1)
conv9_correlate_tc_kernel(const __grid_constant__ CUtensorMap tmB, const C9Params p) {
    pdl_prologue();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NS9 * STAGE9);
    uint64_t* a_full = bars; uint64_t* b_full = a_full + NS9; uint64_t* empty = b_full + NS9; uint64_t* done = empty + NS9;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int NH = p.N * p.H;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmB);
        for (int i = 0; i < NS9; ++i) { ptx::mbar_init(&a_full[i], BUILD_WARPS); ptx::mbar_init(&b_full[i], 1); ptx::mbar_init(&empty[i], 1); }
        ptx::mbar_init(done, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) { ptx::tmem_alloc(tmem_slot, TMEM9); ptx::tmem_relinquish(); }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer: the big tensor's K block, both planes =====================
        if (lane == 0) {
            int st = 0; uint32_t ph = 0;
            for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x) {
                const int wb = wk % p.wblocks; const int row = wk / p.wblocks;      // row = n * H + h
                ptx::mbar_wait(&empty[st], ph ^ 1);
                ptx::mbar_expect_tx(&b_full[st], B_STAGE);
                uint8_t* dst = smem + st * STAGE9 + A_STAGE;
                ptx::tma_load_3d(dst, &tmB, &b_full[st], 0, wb * KB, row);
                ptx::tma_load_3d(dst + B_PLANE, &tmB, &b_full[st], 0, wb * KB, NH + row);
                if (++st == NS9) { st = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const bool leader = ptx::elect_one();
        const uint32_t fmt = ptx::umma_fmt_bits(p.f16 != 0);
        constexpr uint32_t a_hi = ptx::umma_desc_hi_k_sw128();
        constexpr uint32_t b_hi = static_cast<uint32_t>((8 * 64) >> 4) | (1u << 14) | (4u << 29);      // MN-major, 64-byte swizzle
        int st = 0; uint32_t ph = 0;
        uint32_t first = 0;
        for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x) {
            ptx::mbar_wait(&b_full[st], ph);
            ptx::mbar_wait(&a_full[st], ph);
            ptx::tc_fence_after();
            const uint32_t a0 = ptx::smem_u32(smem + st * STAGE9), b0 = a0 + A_STAGE;
#pragma unroll
            for (int mb = 0; mb < 2; ++mb) {
#pragma unroll
                for (int k = 0; k < KB / 16; ++k) {
                    const uint32_t ak = a0 + mb * A_MBLOCK + k * 32;
                    const uint32_t bk = mn_lo(b0 + k * 16 * 64, B_PLANE);
                    if (leader) {
                        mma2(tmem_base + mb * 64, ptx::umma_desc_lo(ak), a_hi, bk, b_hi, idesc_k_mn(128, 64) | fmt, k ? 1u : first);
                        mma2(tmem_base + mb * 64 + 32, ptx::umma_desc_lo(ak + A_PLANE), a_hi, bk, b_hi, idesc_k_mn(128, 32) | fmt, 1u);
                    }
                }
            }
            first = 1;
            if (leader) ptx::tc_commit(&empty[st]);
            __syncwarp();
            if (++st == NS9) { st = 0; ph ^= 1; }
        }
        if (leader) ptx::tc_commit(done);
        __syncwarp();
    } else {
        // ===================== A-tile builders (two groups of 7 warps), then epilogue (warps 2..5) =====================
        static_assert(NS9 == BUILD_GROUPS, "one pipeline stage per builder group");
        const int grp = (warp - 2) / BUILD_WARPS;
        const int task = threadIdx.x - 64 - grp * BUILD_WARPS * 32;
        const bool active = task < BUILD_TASKS;
        const int j = task & 7, cr = task >> 3;
        const int r = cr % T9, c = active ? cr / T9 : 0;
        const size_t HW = static_cast<size_t>(p.H) * p.W;
        float cur[16], nxt[16];
        auto fetch = [&](int wk, float (&v)[16]) {
            const int wb = wk % p.wblocks; const int row = wk / p.wblocks;
            const int n = row / p.H, h = row - n * p.H;
            const int ih = h + r - TPAD, iw0 = wb * KB + j * 8 - TPAD;
            const float* src = p.s + (static_cast<size_t>(n) * p.sc + c) * HW + static_cast<size_t>(ih) * p.W;
            const bool rok = active && ih >= 0 && ih < p.H;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int iw = iw0 + i;
                v[i] = (rok && iw >= 0 && iw < p.W) ? __ldg(src + iw) : 0.f;
            }
        };
        const int wk_first = blockIdx.x + grp * gridDim.x, wk_step = BUILD_GROUPS * gridDim.x;
        if (wk_first < p.work) fetch(wk_first, cur);
        const int st = grp; uint32_t ph = 0;
        for (int wk = wk_first; wk < p.work; wk += wk_step) {
            const int nwk = wk + wk_step;
            if (nwk < p.work) fetch(nwk, nxt);
            // split into hi / lo planes (16-bit patterns) in the format of the big tensor
            uint32_t hi[16], lo[16];
            split16(cur, p.f16 != 0, p.small_slot ? p.small_slot[0] : 1.f, hi, lo);
            ptx::mbar_wait(&empty[st], ph ^ 1);
            if (active) {
                uint8_t* A = smem + st * STAGE9;
#pragma unroll
                for (int s = 0; s < T9; ++s) {
                    const int m = cr * T9 + s;
                    const int row = m & 127;
                    uint8_t* dst = A + (m >> 7) * A_MBLOCK + row * 128 + ((j ^ (row & 7)) << 4);
                    uint4 wh, wl;
                    wh.x = hi[s] | (hi[s + 1] << 16); wh.y = hi[s + 2] | (hi[s + 3] << 16);
                    wh.z = hi[s + 4] | (hi[s + 5] << 16); wh.w = hi[s + 6] | (hi[s + 7] << 16);
                    wl.x = lo[s] | (lo[s + 1] << 16); wl.y = lo[s + 2] | (lo[s + 3] << 16);
                    wl.z = lo[s + 4] | (lo[s + 5] << 16); wl.w = lo[s + 6] | (lo[s + 7] << 16);
                    *reinterpret_cast<uint4*>(dst) = wh;
                    *reinterpret_cast<uint4*>(dst + A_PLANE) = wl;
                }
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&a_full[st]);
            ph ^= 1;
#pragma unroll
            for (int i = 0; i < 16; ++i) cur[i] = nxt[i];
        }
        if (warp < 6) {
            const int q = warp & 3;
            ptx::mbar_wait(done, 0);
            ptx::tc_fence_after();
            float* out = p.partial + static_cast<size_t>(blockIdx.x) * p.part_stride + p.c_off * 81;
#pragma unroll 1
            for (int mb = 0; mb < 2; ++mb) {
                const int m = mb * 128 + q * 32 + lane;
#pragma unroll 1
                for (int hb = 0; hb < 2; ++hb) {
                    uint32_t d0[16], d1[16];
                    const uint32_t ta = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + mb * 64 + hb * 16;
                    ptx::tmem_ld16(ta, d0);
                    ptx::tmem_ld16(ta + 32, d1);
                    ptx::tmem_ld_wait();
                    if (m < MROWS) {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            out[static_cast<size_t>(hb * 16 + i) * p.cs_total * 81 + m] = __uint_as_float(d0[i]) + __uint_as_float(d1[i]);
                    }
                }
            }
            ptx::tc_fence_before();
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, TMEM9);
}

// ---------------------------------------------------------------------------------------------
// expand: out[p][32] = sum_{c,r,s} S[c][p + (r,s) - 4] * Wt[(c,r,s)][32] (+ bias)      (layer-0 forward;
// layer-22 backward-data with flipped taps).  GEMM with M = 128 consecutive pixels of an image
// row, N = 32, K = (c, r, s).  The SAME shifted-window rows as above are built, but now they are the
// MN-major A operand (a row = one k, 64 pixels contiguous; M = 128 = two 64-pixel groups at LBO).
// K is walked in "quarters" of 7 (c, r) pairs = 63 rows (+1 zero row): one pipeline stage = one
// quarter of one tile.  The weights are split into bf16 planes once per CTA and stay in shared
// memory, K-major, planes adjacent so that a0 x [w0|w1|w2], a1 x [w0|w1], a2 x w0 are three MMAs
// into adjacent accumulators D0 | D1 | D2.
// ---------------------------------------------------------------------------------------------
constexpr int XQ_PAIRS = 7, XQ_ROWS = 64;              // (c, r) pairs and k rows per quarter
constexpr int XA_GROUP = XQ_ROWS * 128;                // one 64-pixel group of one plane: 8 KB
constexpr int XB_PLANE = TCB * 128;                    // 32 weight rows x 64 k: 4 KB
constexpr int X_BUILD_WARPS = 4;                       // 112 tasks = (group, pair in quarter, chunk)
constexpr int X_GROUPS = 2;                            // builder groups alternate pipeline items
constexpr int X_THREADS = (2 + X_GROUPS * X_BUILD_WARPS + 4) * 32; // + 4 epilogue warps

struct X9Params {
    const float* s0; const float* s1; int sc;
    const float* w; const float* bias; int flip;
    bf16* out_hi; bf16* out_lo; int planes;
    int N, H, W, wtiles, tiles;
};

template <int CS, int PL> struct X9Cfg {
    static constexpr int PAIRS = 9 * CS;
    static constexpr int NQ = (PAIRS + XQ_PAIRS - 1) / XQ_PAIRS;
    static constexpr int A_STAGE = PL * 2 * XA_GROUP;
    static constexpr int B_BYTES = NQ * PL * XB_PLANE;
    static constexpr int NSTG = (232448 - 2048 - B_BYTES) / A_STAGE >= 3 ? 3 : 2;
    static constexpr int SMEM = 1024 + NSTG * A_STAGE + B_BYTES + 256;
    static constexpr int STAGE_COLS = PL * TCB;                    // D0 | D1 (| D2)
    static constexpr int TMEM_COLS = 2 * STAGE_COLS <= 128 ? 128 : 256;
};

__host__ __device__ constexpr uint32_t idesc_mn_k(int M, int N) {     // A MN-major, B K-major
    return (1u << 4) | (1u << 15) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}

template <int CS, int PL>
__global__ void __launch_bounds__(X_THREADS, 1)
conv9_expand_tc_kernel(const X9Params p) {
    pdl_prologue();
    using C = X9Cfg<CS, PL>;
    constexpr int NQ = C::NQ, NSTG = C::NSTG;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + NSTG * C::A_STAGE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + C::B_BYTES);
    uint64_t* a_full = bars; uint64_t* a_empty = a_full + NSTG; uint64_t* t_full = a_empty + NSTG; uint64_t* t_empty = t_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool f16 = PL == 2 && pc_f16(p.planes);          // operand format = format of the output planes

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < NSTG; ++i) { ptx::mbar_init(&a_full[i], X_BUILD_WARPS); ptx::mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&t_full[i], 1); ptx::mbar_init(&t_empty[i], 4); }
        ptx::fence_barrier_init();
    }
    if (warp == 1) { ptx::tmem_alloc(tmem_slot, C::TMEM_COLS); ptx::tmem_relinquish(); }
    // weights -> bf16 planes, K-major with the 128-byte swizzle: [quarter][plane][co][64 k]; zero the
    // A rows no builder writes (row 63 of every quarter, pairs past the last one)
    for (int i = threadIdx.x; i < NQ * TCB * XQ_ROWS; i += X_THREADS) {
        const int kk = i % XQ_ROWS; int t = i / XQ_ROWS; const int co = t % TCB; const int q = t / TCB;
        const int pair = q * XQ_PAIRS + kk / 9, sft = kk % 9;
        float v = 0.f;
        if (kk < 63 && pair < C::PAIRS) {
            const int c = pair / 9, r = pair % 9;
            v = p.flip ? p.w[((c * TCB + co) * 9 + (8 - r)) * 9 + (8 - sft)] : p.w[((co * CS + c) * 9 + r) * 9 + sft];
        }
        uint8_t* dst = sB + q * PL * XB_PLANE + co * 128 + (((kk >> 3) ^ (co & 7)) << 4) + (kk & 7) * 2;
        if (f16) {
            // fp16 operands (PL = 2): planes of ST_W_F16_SCALE * w; the epilogue multiplies by the inverse
            uint32_t h2, l2; split2h(v * ST_W_F16_SCALE, 0.f, h2, l2);
            *reinterpret_cast<uint16_t*>(dst) = static_cast<uint16_t>(h2 & 0xffffu);
            *reinterpret_cast<uint16_t*>(dst + XB_PLANE) = static_cast<uint16_t>(l2 & 0xffffu);
            continue;
        }
        float rem = v;
#pragma unroll
        for (int pl = 0; pl < PL; ++pl) {
            const bf16 h = __float2bfloat16_rn(rem);
            *reinterpret_cast<bf16*>(dst + pl * XB_PLANE) = h;
            rem -= __bfloat162float(h);
        }
    }
    for (int i = threadIdx.x; i < NSTG * C::A_STAGE / 16; i += X_THREADS) reinterpret_cast<uint4*>(sA)[i] = make_uint4(0, 0, 0, 0);
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 1) {
        // ===================== MMA issuer =====================
        const bool leader = ptx::elect_one();
        const uint32_t fmt = ptx::umma_fmt_bits(f16);
        constexpr uint32_t a_hi = static_cast<uint32_t>(1024 >> 4) | (1u << 14) | (2u << 29);      // MN-major SW128: 8 k rows = 1024 B
        constexpr uint32_t b_hi = ptx::umma_desc_hi_k_sw128();
        int st = 0; uint32_t ph = 0; int it = 0;
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
            const int acc = it & 1;
            ptx::mbar_wait(&t_empty[acc], ((it >> 1) & 1) ^ 1);
            ptx::tc_fence_after();
            const uint32_t d0 = tmem_base + acc * C::STAGE_COLS;
            uint32_t accum = 0;
#pragma unroll 1
            for (int q = 0; q < NQ; ++q) {
                ptx::mbar_wait(&a_full[st], ph);
                ptx::tc_fence_after();
                const uint32_t a0 = ptx::smem_u32(sA + st * C::A_STAGE), b0 = ptx::smem_u32(sB + q * PL * XB_PLANE);
#pragma unroll
                for (int k = 0; k < XQ_ROWS / 16; ++k) {
                    const uint32_t bk = ptx::umma_desc_lo(b0 + k * 32);
                    if (leader) {
#pragma unroll
                        for (int pa = 0; pa < PL; ++pa) {
                            const uint32_t ak = mn_lo(a0 + pa * 2 * XA_GROUP + k * 16 * 128, XA_GROUP);
                            mma2(d0 + pa * TCB, ak, a_hi, bk, b_hi, idesc_mn_k(128, (PL - pa) * TCB) | fmt, pa ? 1u : accum);
                        }
                    }
                    accum = 1;
                }
                if (leader) ptx::tc_commit(&a_empty[st]);
                __syncwarp();
                if (++st == NSTG) { st = 0; ph ^= 1; }
            }
            if (leader) ptx::tc_commit(&t_full[acc]);
            __syncwarp();
        }
    } else if (warp >= 2 && warp < 2 + X_GROUPS * X_BUILD_WARPS) {
        // ===================== A builders: task = (pixel group, pair within the quarter, 16-byte chunk);
        // the two builder groups take alternate items (item = quarter of a tile) =====================
        const int grp = (warp - 2) / X_BUILD_WARPS;
        const int task = threadIdx.x - 64 - grp * X_BUILD_WARPS * 32;
        const bool active = task < 2 * XQ_PAIRS * 8;
        const int j = task & 7, pq = (task >> 3) % XQ_PAIRS, g = active ? (task >> 3) / XQ_PAIRS : 0;
        const size_t HW = static_cast<size_t>(p.H) * p.W;
        float cur[16], nxt[16];
        auto fetch = [&](int tile, int q, float (&v)[16]) {
            const int wt = tile % p.wtiles; const int row = tile / p.wtiles;
            const int n = row / p.H, h = row - n * p.H;
            const int pair = q * XQ_PAIRS + pq;
            const int c = pair / 9, r = pair % 9;
            const int ih = h + r - TPAD, iw0 = wt * 128 + g * 64 + j * 8 - TPAD;
            const bool rok = active && pair < C::PAIRS && ih >= 0 && ih < p.H;
            const float* base = (CS == 6 && c >= 3) ? p.s1 : p.s0;
            const float* src = base + (static_cast<size_t>(n) * p.sc + (c >= 3 ? c - 3 : c)) * HW + static_cast<size_t>(ih) * p.W;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int iw = iw0 + i;
                v[i] = (rok && iw >= 0 && iw < p.W) ? __ldg(src + iw) : 0.f;
            }
        };
        // item i of this CTA = (tile blockIdx.x + (i / NQ) * gridDim.x, quarter i % NQ), pipeline stage i % NSTG
        const int my_tiles = p.tiles > static_cast<int>(blockIdx.x) ? (p.tiles - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
        const int items = my_tiles * NQ;
        auto fetch_item = [&](int i, float (&v)[16]) { fetch(blockIdx.x + (i / NQ) * gridDim.x, i % NQ, v); };
        if (grp < items) fetch_item(grp, cur);
        for (int i = grp; i < items; i += X_GROUPS) {
            if (i + X_GROUPS < items) fetch_item(i + X_GROUPS, nxt);
            const int st = i % NSTG; const uint32_t ph = (i / NSTG) & 1;
            uint32_t pw[PL][16];
            if (f16) {
#pragma unroll
                for (int k = 0; k < 16; k += 2) {
                    uint32_t h2, l2; split2h(cur[k], cur[k + 1], h2, l2);
                    pw[0][k] = h2 & 0xffffu; pw[0][k + 1] = h2 >> 16;
                    pw[PL - 1][k] = l2 & 0xffffu; pw[PL - 1][k + 1] = l2 >> 16;
                }
            } else {
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    float rem = cur[k];
#pragma unroll
                    for (int pl = 0; pl < PL; ++pl) {
                        const bf16 h = __float2bfloat16_rn(rem);
                        pw[pl][k] = __bfloat16_as_ushort(h);
                        rem -= __bfloat162float(h);
                    }
                }
            }
            ptx::mbar_wait(&a_empty[st], ph ^ 1);
            if (active) {
                uint8_t* A = sA + st * C::A_STAGE + g * XA_GROUP;
#pragma unroll
                for (int sft = 0; sft < T9; ++sft) {
                    const int row = pq * 9 + sft;
                    uint8_t* dst = A + row * 128 + ((j ^ (row & 7)) << 4);
#pragma unroll
                    for (int pl = 0; pl < PL; ++pl) {
                        uint4 wv;
                        wv.x = pw[pl][sft] | (pw[pl][sft + 1] << 16); wv.y = pw[pl][sft + 2] | (pw[pl][sft + 3] << 16);
                        wv.z = pw[pl][sft + 4] | (pw[pl][sft + 5] << 16); wv.w = pw[pl][sft + 6] | (pw[pl][sft + 7] << 16);
                        *reinterpret_cast<uint4*>(dst + pl * 2 * XA_GROUP) = wv;
                    }
                }
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&a_full[st]);
#pragma unroll
            for (int k = 0; k < 16; ++k) cur[k] = nxt[k];
        }
    } else if (warp >= 2 + X_GROUPS * X_BUILD_WARPS) {
        // ===================== epilogue: accumulators -> bias -> split-bf16 NHWC =====================
        const int q = warp & 3;
        const int m = q * 32 + lane;
        int it = 0;
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
            const int acc = it & 1;
            const int wt = tile % p.wtiles; const int row = tile / p.wtiles;
            const int wq = wt * 128 + m;
            const bool valid = wq < p.W;
            const size_t o = (static_cast<size_t>(row) * p.W + wq) * TCB;
            ptx::mbar_wait(&t_full[acc], (it >> 1) & 1);
            ptx::tc_fence_after();
            const uint32_t ta = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * C::STAGE_COLS;
#pragma unroll
            for (int hb = 0; hb < 2; ++hb) {
                uint32_t d[PL][16];
#pragma unroll
                for (int pl = 0; pl < PL; ++pl) ptx::tmem_ld16(ta + pl * TCB + hb * 16, d[pl]);
                ptx::tmem_ld_wait();
                if (valid) {
                    float v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        float t = __uint_as_float(d[PL - 1][i]);                  // smallest terms first
#pragma unroll
                        for (int pl = PL - 2; pl >= 0; --pl) t += __uint_as_float(d[pl][i]);
                        if (f16) t *= ST_W_F16_INV_SCALE;
                        v[i] = t + (p.bias ? p.bias[hb * 16 + i] : 0.f);
                    }
                    store16p(p.out_hi, p.out_lo, p.planes, o + hb * 16, v);
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&t_empty[acc]);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
}

template <int CS, int PL>
int launch_expand(const X9Params& p, cudaStream_t stream) {
    using C = X9Cfg<CS, PL>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv9_expand_tc_kernel<CS, PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { st_set_error("conv9_expand_tc smem attribute: %s", cudaGetErrorString(e)); return ST_ERR_CUDA; }
        attr_set = true;
    }
    const int grid = p.tiles < st_num_sms() ? p.tiles : st_num_sms();
    st_launch(grid, X_THREADS, C::SMEM, stream)(conv9_expand_tc_kernel<CS, PL>, p);
    ST_CHECK_LAUNCH("conv9_expand_tc_kernel");
    return ST_OK;
}

}  // namespace

// planes: 2 or 3 output planes (= operand planes)
int conv9_expand_tc_run(const float* s0, const float* s1, int sc, int CS, const float* w, const float* bias, int flip,
                        bf16* out_hi, bf16* out_lo, int planes, int N, int H, int W, cudaStream_t stream) {
    X9Params p;
    p.s0 = s0; p.s1 = s1; p.sc = sc; p.w = w; p.bias = bias; p.flip = flip;
    p.out_hi = out_hi; p.out_lo = out_lo; p.planes = planes;
    p.N = N; p.H = H; p.W = W; p.wtiles = (W + 127) / 128; p.tiles = N * H * p.wtiles;
    const bool three = pc_planes(planes) >= 3 && !pc_f16(planes);
    if (CS == 3) return three ? launch_expand<3, 3>(p, stream) : launch_expand<3, 2>(p, stream);
    return three ? launch_expand<6, 3>(p, stream) : launch_expand<6, 2>(p, stream);
}

// One 3-channel slice of the small tensor.  Returns the number of partials written (= grid).
int conv9_correlate_tc_run(const bf16* b_hi, bool f16, const float* small_slot, const float* s, int sc, int cs_total, int c_off,
                           float* partial, int N, int H, int W, int* nparts, cudaStream_t stream) {
    C9Params p;
    p.f16 = f16 ? 1 : 0; p.small_slot = small_slot;
    p.s = s; p.sc = sc; p.N = N; p.H = H; p.W = W;
    p.wblocks = (W + KB - 1) / KB; p.work = N * H * p.wblocks;
    p.partial = partial; p.part_stride = TCB * cs_total * 81; p.cs_total = cs_total; p.c_off = c_off;
    CUtensorMap tb;
    uint64_t bd[3] = {(uint64_t)TCB, (uint64_t)W, (uint64_t)N * H * 2};
    uint32_t bb[3] = {(uint32_t)TCB, (uint32_t)KB, 1};
    int rc;
    if ((rc = st_make_tmap(&tb, b_hi, 3, bd, bb, 64))) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv9_correlate_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM9);
        if (e != cudaSuccess) { st_set_error("conv9_correlate_tc smem attribute: %s", cudaGetErrorString(e)); return ST_ERR_CUDA; }
        attr_set = true;
    }
    const int grid = p.work < st_num_sms() ? p.work : st_num_sms();
    st_launch(grid, THREADS9, SMEM9, stream)(conv9_correlate_tc_kernel, tb, p);
    ST_CHECK_LAUNCH("conv9_correlate_tc_kernel");
    *nparts = grid;
    return ST_OK;
}
