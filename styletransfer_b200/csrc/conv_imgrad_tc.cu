// Backward-data of VGG conv1_1 to the image (autograd's backward to `input_image`, reference network.py:451) on the
// tensor cores in "project, then gather" form:
//     P[q][(c, t)] = sum_co g[q][co] * Wf[(c, t)][co]          one GEMM, M = pixels q, K = 64, N = 27 (-> 32)
//     gx[c][h][w]  = sum_t  P[(h + r - 1, w + s - 1)][(c, t)]   t = 3 r + s, nine adds per output on the CUDA cores
// Wf is the packed backward weight the implicit-GEMM kernel already uses (flipped taps, ops.PackedWeight mode 2,
// layout [c][tap][co]), so (c, tap) is simply the row index of its first 27 rows.  The implicit-GEMM form of this layer
// (conv_tc_kernel<3, 16, 2, RES>) issues 72 MMAs of N = 32 / 16 per 128 output pixels, each re-reading a 4 KB activation
// operand from shared memory; this form issues 16 (two M blocks cover the 18 x 10 halo box of a 16 x 8 pixel tile:
// 1.4 x redundant rows, 4.5 x fewer MMAs) and was 40 us of a 512 x 512 Gatys closure.
//
// One persistent CTA per SM, 320 threads: warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 epilogue (two per TMEM lane
// quarter: warps 2..5 own M block 0 = box pixels 0..127, warps 6..9 M block 1 = box pixels 128..255, of which 128..179
// exist).  Operand split as everywhere else: hi x [w_hi | w_lo] as ONE N = 64 MMA (main | correction accumulators), then
// lo x w_hi (N = 32) into the correction accumulator.  The epilogue writes (main + correction) * acc_mul of its box pixel
// to a shared-memory table P[256][27] (two tables, alternating between tiles: one barrier per tile), then the 256 epilogue
// threads gather the nine taps of their output pixel and channel(s).
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.cuh"
#include <stdlib.h>

namespace {

constexpr int TH = 16, TW = 8, BOX_H = TH + 2, BOX_W = TW + 2, BOX_PIX = BOX_H * BOX_W;      // 180 box pixels
constexpr int A_BYTES = BOX_PIX * 128;                       // one plane of a box: 64 channels x 2 B per pixel row
constexpr int A_PLANE = (A_BYTES + 1023) & ~1023;            // 23552
constexpr int A_SLOT = 2 * A_PLANE, NA = 3;
constexpr int A_TAIL = 256 * 128 - A_PLANE;                  // M block 1 of the lo plane of the last slot reads this far past it
constexpr int W_ROWS = 27, W_PLANE = 32 * 128, W_BYTES = 2 * W_PLANE;     // [w_hi rows 0..31 | w_lo rows 32..63]
constexpr int P_LD = 27, P_TABLE = 256 * P_LD, P_BYTES = 2 * P_TABLE * 4;      // two tables: tiles alternate
constexpr int BAR_BYTES = 256;
constexpr int SMEM = 1024 + NA * A_SLOT + A_TAIL + W_BYTES + P_BYTES + BAR_BYTES;
constexpr int THREADS = 320, EPI_FIRST = 2;
constexpr int STAGE_COLS = 128;                              // per stage: block 0 [main 32 | corr 32], block 1 [main 32 | corr 32]
static_assert(SMEM <= 232448, "shared memory");

struct ImgradParams {
    int N, H, W, tiles_h, tiles_w, work, nchw_c;
    uint32_t fmt_bits;
    float acc_mul; const float* acc_mul_slot;
    float* out;                                              // [N][nchw_c][H][W] fp32
};

__global__ void __launch_bounds__(THREADS, 1)
conv_imgrad_tc_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                      const __grid_constant__ CUtensorMap tmW_hi, const __grid_constant__ CUtensorMap tmW_lo, const ImgradParams p) {
    pdl_trigger();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sW = smem + NA * A_SLOT + A_TAIL;               // 1024-byte aligned (A_SLOT and A_TAIL are multiples of 1024)
    float* sP = reinterpret_cast<float*>(sW + W_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sP) + P_BYTES);
    uint64_t* a_full = bars;            // [NA]
    uint64_t* a_empty = a_full + NA;    // [NA]
    uint64_t* w_full = a_empty + NA;    // [1]
    uint64_t* t_full = w_full + 1;      // [2]
    uint64_t* t_empty = t_full + 2;     // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA_hi); ptx::prefetch_tmap(&tmA_lo); ptx::prefetch_tmap(&tmW_hi); ptx::prefetch_tmap(&tmW_lo);
        for (int i = 0; i < NA; ++i) { ptx::mbar_init(&a_full[i], 1); ptx::mbar_init(&a_empty[i], 1); }
        ptx::mbar_init(w_full, 1);
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&t_full[i], 1); ptx::mbar_init(&t_empty[i], 8); }
        ptx::fence_barrier_init();
    }
    if (warp == 1) { ptx::tmem_alloc(tmem_slot, 2 * STAGE_COLS); ptx::tmem_relinquish(); }
    // weight rows 27..31 of both planes are never written by TMA: zero the tile first (the MMA reads all 32 / 64 rows), and
    // the bytes behind the last activation slot that M block 1 reads as rows 184..255 (finite garbage would do; zeros are free)
    for (int i = threadIdx.x; i < W_BYTES / 16; i += THREADS) reinterpret_cast<uint4*>(sW)[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int i = threadIdx.x; i < A_TAIL / 16; i += THREADS) reinterpret_cast<uint4*>(sA + NA * A_SLOT)[i] = make_uint4(0u, 0u, 0u, 0u);
    ptx::fence_proxy_async();
    pdl_wait();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int tiles_per_img = p.tiles_h * p.tiles_w;

    if (warp == 0) {
        if (lane == 0) {
            ptx::mbar_expect_tx(w_full, 2 * W_ROWS * 128);
            ptx::tma_load_3d(sW, &tmW_hi, w_full, 0, 0, 0);
            ptx::tma_load_3d(sW + W_PLANE, &tmW_lo, w_full, 0, 0, 0);
            int a = 0; uint32_t aph = 0;
            for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x) {
                const int n = wk / tiles_per_img, t = wk - n * tiles_per_img;
                const int th = t / p.tiles_w, tw = t - th * p.tiles_w;
                ptx::mbar_wait(&a_empty[a], aph ^ 1);
                ptx::mbar_expect_tx(&a_full[a], 2 * A_BYTES);
                ptx::tma_load_4d(sA + a * A_SLOT, &tmA_hi, &a_full[a], 0, tw * TW - 1, th * TH - 1, n);
                ptx::tma_load_4d(sA + a * A_SLOT + A_PLANE, &tmA_lo, &a_full[a], 0, tw * TW - 1, th * TH - 1, n);
                if (++a == NA) { a = 0; aph ^= 1; }
            }
        }
    } else if (warp == 1) {
        const bool leader = ptx::elect_one();
        constexpr uint32_t d_hi = ptx::umma_desc_hi_k_sw128();
        const uint32_t id64 = ptx::umma_idesc_nofmt(128, 64) | p.fmt_bits, id32 = ptx::umma_idesc_nofmt(128, 32) | p.fmt_bits;
        const uint32_t w_lo0 = ptx::umma_desc_lo(ptx::smem_u32(sW));
        ptx::mbar_wait(w_full, 0);
        int a = 0; uint32_t aph = 0; int it = 0;
        for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x, ++it) {
            const int acc = it & 1;
            ptx::mbar_wait(&t_empty[acc], ((it >> 1) & 1) ^ 1);
            ptx::mbar_wait(&a_full[a], aph);
            ptx::tc_fence_after();
            const uint32_t a0 = ptx::smem_u32(sA + a * A_SLOT);
            if (leader) {
#pragma unroll
                for (int mb = 0; mb < 2; ++mb) {
                    const uint32_t d = tmem_base + acc * STAGE_COLS + mb * 64;
                    const uint32_t ah = ptx::umma_desc_lo(a0 + mb * 128 * 128), al = ptx::umma_desc_lo(a0 + A_PLANE + mb * 128 * 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        ptx::mma_ss2h(d, ah + ((k * 32) >> 4), d_hi, w_lo0 + ((k * 32) >> 4), d_hi, id64, k ? 1u : 0u);        // main | hi x w_lo
                        ptx::mma_ss2h(d + 32, al + ((k * 32) >> 4), d_hi, w_lo0 + ((k * 32) >> 4), d_hi, id32, 1u);            // + lo x w_hi
                    }
                }
                ptx::tc_commit(&a_empty[a]);
                ptx::tc_commit(&t_full[acc]);
            }
            __syncwarp();
            if (++a == NA) { a = 0; aph ^= 1; }
        }
    } else {
        const int e = warp - EPI_FIRST, q = warp & 3, mb = e >> 2;      // warp & 3 = the TMEM lane quarter this warp may read
        const int row = mb * 128 + q * 32 + lane;                        // box pixel of this thread
        const int et = threadIdx.x - EPI_FIRST * 32;                     // 0..255
        const float mul = p.acc_mul * slot_inv(p.acc_mul_slot);
        const size_t HW = static_cast<size_t>(p.H) * p.W;
        int it = 0;
        for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x, ++it) {
            const int acc = it & 1;
            ptx::mbar_wait(&t_full[acc], (it >> 1) & 1);
            ptx::tc_fence_after();
            uint32_t m0[16], m1[16], c0[16], c1[16];
            const uint32_t ta = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * STAGE_COLS + mb * 64;
            ptx::tmem_ld16(ta, m0); ptx::tmem_ld16(ta + 16, m1);
            ptx::tmem_ld16(ta + 32, c0); ptx::tmem_ld16(ta + 48, c1);
            ptx::tmem_ld_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&t_empty[acc]);
            // tables alternate between tiles: the one written now was last read two tiles ago, and every thread has passed the
            // barrier of the tile in between since then (one barrier per tile)
            float* tab = sP + acc * P_TABLE;
            if (row < BOX_PIX) {
                float* dst = tab + row * P_LD;
#pragma unroll
                for (int i = 0; i < 16; ++i) dst[i] = (__uint_as_float(m0[i]) + __uint_as_float(c0[i])) * mul;
#pragma unroll
                for (int i = 0; i < W_ROWS - 16; ++i) dst[16 + i] = (__uint_as_float(m1[i]) + __uint_as_float(c1[i])) * mul;
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            {
                // all 256 threads gather: threads 0..127 channels 0 and 1 of their pixel, threads 128..255 channel 2
                const int n = wk / tiles_per_img, t = wk - n * tiles_per_img;
                const int th = t / p.tiles_w, tw = t - th * p.tiles_w;
                const int m = et & (TH * TW - 1), part = et >> 7;
                const int hl = m / TW, wl = m - hl * TW;
                const int h = th * TH + hl, w = tw * TW + wl;
                if (h < p.H && w < p.W) {
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc) {
                        const int c = part ? 2 : cc;
                        if ((part == 0 || cc == 0) && c < p.nchw_c) {
                            float s = 0.f;
#pragma unroll
                            for (int r = 0; r < 3; ++r)
#pragma unroll
                                for (int sx = 0; sx < 3; ++sx) s += tab[((hl + r) * BOX_W + wl + sx) * P_LD + c * 9 + r * 3 + sx];
                            p.out[(static_cast<size_t>(n) * p.nchw_c + c) * HW + static_cast<size_t>(h) * p.W + w] = s;
                        }
                    }
                }
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, 2 * STAGE_COLS);
}

}  // namespace

// x: the gradient of the conv1_1 tap [N][H][W][64], two 16-bit planes; w: packed backward weight, planes [rows >= 3][9][64]
// (ops.PackedWeight mode 2); out: [N][Cnchw <= 3][H][W] fp32.  acc_mul / acc_mul_slot as in conv_tc_run.
bool conv_imgrad_tc_eligible(int ks, int Cin, int Cout_pad, int x_planes, int w_planes, int w_rows_total, int Cnchw) {
    static const bool off = []() { const char* e = getenv("STRANSFER_IMGRAD_GATHER"); return e && e[0] == '0'; }();
    return !off && ks == 3 && Cin == 64 && Cout_pad == 16 && Cnchw >= 1 && Cnchw <= 3 && w_rows_total >= 3 &&
           pc_planes(x_planes) == 2 && pc_planes(w_planes) == 2 && pc_f16(x_planes) == pc_f16(w_planes);
}

int conv_imgrad_tc_run(const bf16* x_hi, const bf16* x_lo, int x_planes, int N, int H, int W,
                       const bf16* w_hi, const bf16* w_lo, int w_rows_total, float* out, int Cnchw,
                       float acc_mul, const float* acc_mul_slot, cudaStream_t stream) {
    ST_CHECK_ARG(x_hi && x_lo && w_hi && w_lo && out, "conv_imgrad_tc: null pointer");
    ImgradParams p;
    p.N = N; p.H = H; p.W = W; p.tiles_h = (H + TH - 1) / TH; p.tiles_w = (W + TW - 1) / TW;
    p.work = N * p.tiles_h * p.tiles_w; p.nchw_c = Cnchw;
    p.fmt_bits = ptx::umma_fmt_bits(pc_f16(x_planes));
    p.acc_mul = acc_mul; p.acc_mul_slot = acc_mul_slot; p.out = out;
    CUtensorMap ah, al, wh, wl;
    const uint64_t ad[4] = {64, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    const uint32_t ab[4] = {64, (uint32_t)BOX_W, (uint32_t)BOX_H, 1};
    const uint64_t wd[3] = {64, 9, (uint64_t)w_rows_total};
    const uint32_t wb[3] = {64, 9, 3};
    int rc;
    if ((rc = st_make_tmap(&ah, x_hi, 4, ad, ab, 128, 1))) return rc;
    if ((rc = st_make_tmap(&al, x_lo, 4, ad, ab, 128, 1))) return rc;
    if ((rc = st_make_tmap(&wh, w_hi, 3, wd, wb, 128, 1))) return rc;
    if ((rc = st_make_tmap(&wl, w_lo, 3, wd, wb, 128, 1))) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_imgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        if (e != cudaSuccess) { st_set_error("conv_imgrad_tc smem attribute: %s", cudaGetErrorString(e)); return ST_ERR_CUDA; }
        attr_set = true;
    }
    const int grid = p.work < st_num_sms() ? p.work : st_num_sms();
    st_launch(grid, THREADS, SMEM, stream)(conv_imgrad_tc_kernel, ah, al, wh, wl, p);
    ST_CHECK_LAUNCH("conv_imgrad_tc_kernel");
    return ST_OK;
}
