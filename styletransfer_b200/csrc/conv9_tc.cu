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
constexpr int THREADS9 = (2 + BUILD_WARPS) * 32;
constexpr int SMEM9 = 1024 + NS9 * STAGE9 + 256;
constexpr int TMEM9 = 128;                      // 2 M blocks x (32 main + 32 correction) columns

struct C9Params {
    const float* s; int sc;                     // small tensor: channel c of sample n at s + (n*sc + c)*H*W
    int N, H, W, wblocks, work;
    float* partial; int part_stride;            // one partial per CTA: [32][cs_total][81]
    int cs_total, c_off;
};

__host__ __device__ constexpr uint32_t idesc_k_mn(int M, int N) {     // A K-major, B MN-major
    return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
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
                        mma2(tmem_base + mb * 64, ptx::umma_desc_lo(ak), a_hi, bk, b_hi, idesc_k_mn(128, 64), k ? 1u : first);
                        mma2(tmem_base + mb * 64 + 32, ptx::umma_desc_lo(ak + A_PLANE), a_hi, bk, b_hi, idesc_k_mn(128, 32), 1u);
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
        // ===================== A-tile builders (warps 2..8), then epilogue (warps 2..5) =====================
        const int task = threadIdx.x - 64;
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
        if (static_cast<int>(blockIdx.x) < p.work) fetch(blockIdx.x, cur);
        int st = 0; uint32_t ph = 0;
        for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x) {
            const int nwk = wk + gridDim.x;
            if (nwk < p.work) fetch(nwk, nxt);
            // split into bf16 hi / lo (16-bit patterns)
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const bf16 h = __float2bfloat16_rn(cur[i]);
                const bf16 l = __float2bfloat16_rn(cur[i] - __bfloat162float(h));
                hi[i] = __bfloat16_as_ushort(h); lo[i] = __bfloat16_as_ushort(l);
            }
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
            if (++st == NS9) { st = 0; ph ^= 1; }
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

}  // namespace

// One 3-channel slice of the small tensor.  Returns the number of partials written (= grid).
int conv9_correlate_tc_run(const bf16* b_hi, const float* s, int sc, int cs_total, int c_off, float* partial,
                           int N, int H, int W, int* nparts, cudaStream_t stream) {
    C9Params p;
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
    conv9_correlate_tc_kernel<<<grid, THREADS9, SMEM9, stream>>>(tb, p);
    ST_CHECK_LAUNCH("conv9_correlate_tc_kernel");
    *nparts = grid;
    return ST_OK;
}
