// tcgen05 implicit-GEMM convolution with a run-time tap geometry: the shapes of the transform net
// that the stride-1 / 64-channel kernel of conv_tc.cu does not take
//   * 3x3 stride-2 forward            (network.py:533-547: 32 -> 64 and 64 -> 128, "down-sampling"),
//   * their backward-data (a transposed convolution), split into the 4 output-parity classes
//     (dx[2i+a][2j+b] only receives the taps r = a+1 (mod 2), s = b+1 (mod 2): 1, 2, 2 and 4 taps,
//     no multiplications by the inserted zeros),
//   * 3x3 stride-1 convs whose K side has 32 channels (backward-data of the 64 -> 32 layer,
//     network.py:594-599) with 64-byte-swizzled operand rows.
// Same structure as conv_tc_kernel (warp 0 TMA producer, warp 1 MMA issuer, warps 4-11 epilogue,
// A ring / B ring / two TMEM stages, main + correction accumulators); what differs is that the list
// of A boxes per K chunk, the taps each box feeds (row offset into the box, weight tap index) and
// the output pixel mapping come from a small table in the kernel parameters:
//   A box   = ROWS x 16 pixels x KC channels at (w0*cs + dw, h0*cs + dh); for the stride-2 forward
//             the tensor map has a traversal stride of 2 pixels, so a box holds every second pixel
//             and every second row (one box per horizontal tap and row parity),
//   output  = pixel (h*os + oa, w*os + ob) of an OH x OW tensor.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.cuh"
#include "conv_tc_epilogue.cuh"
#include <string.h>

namespace {

constexpr int X_MAX_BOX = 6, X_MAX_TAP = 9, X_MAX_CLS = 4;
constexpr int X_ROWS = TILE_H + 2;                 // pixel rows of the largest A box of the 3x3 modes
constexpr int X_ROWS9 = TILE_H + 8;                // ... of the vertical 9-tap mode

struct XBox { int dw, dh, ntap; int row_off[X_MAX_TAP]; int widx[X_MAX_TAP]; };
struct XClass { int nbox, oa, ob, taps; XBox box[X_MAX_BOX]; };
struct XGeom {
    int ncls, per_cls, cs, os, OH, OW;              // classes, work items per class, coordinate / output scales
    int a_tx;                                       // bytes one A box load delivers (all planes)
    XClass cls[X_MAX_CLS];
};

template <int BN, int PL, int KC, int ROWS = X_ROWS> struct XCfg {
    static constexpr int ROWB = KC * 2;                                  // operand row: 64 or 128 bytes
    static constexpr int A_PLANE = ROWS * TILE_W * ROWB;
    static constexpr int B_PLANE = BN * ROWB;
    static constexpr int A_SLOT = PL * A_PLANE, B_SLOT = PL * B_PLANE;
    static constexpr int BAR_BYTES = 256;
    static constexpr int SMEM_MAX = 232448;
    static constexpr int NA = (PL * A_PLANE <= 40960) ? 3 : 2;
    static constexpr int NB_FIT = (SMEM_MAX - 1024 - BAR_BYTES - NA * A_SLOT) / B_SLOT;
    static constexpr int NB = NB_FIT > 6 ? 6 : NB_FIT;
    static_assert(NB >= 2, "weight ring needs two slots");
    static constexpr int SMEM = 1024 + NA * A_SLOT + NB * B_SLOT + BAR_BYTES;
    static constexpr int NACC = PL == 1 ? 1 : (PL == 3 && BN <= 64) ? 3 : 2;
    static constexpr int STAGE_COLS = NACC * BN;
    static constexpr int TMEM_COLS = (2 * STAGE_COLS <= 32) ? 32 : (2 * STAGE_COLS <= 64) ? 64 :
                                     (2 * STAGE_COLS <= 128) ? 128 : (2 * STAGE_COLS <= 256) ? 256 : 512;
    // K-major, swizzled: SBO = 8 rows, layout 2 = 128-byte swizzle, 4 = 64-byte swizzle
    static constexpr uint32_t DESC_HI = static_cast<uint32_t>((8 * ROWB) >> 4) | (1u << 14) | ((ROWB == 128 ? 2u : 4u) << 29);
};

struct XWork { int cls, nb, tw, th, n; };
// classes are the slowest index (heaviest first), so every CTA of the persistent grid gets the same mix
__device__ __forceinline__ XWork x_decode(int wk, const ConvTcParams& p, int per_cls) {
    XWork w;
    // cls < 4: compare instead of dividing by the run-time per_cls
    w.cls = (wk >= per_cls) + (wk >= 2 * per_cls) + (wk >= 3 * per_cls);
    uint32_t t = static_cast<uint32_t>(wk - w.cls * per_cls), nb, tw, th, n;
    p.fd_nb.divmod(t, t, nb);
    p.fd_tw.divmod(t, t, tw);
    p.fd_th.divmod(t, n, th);
    w.nb = static_cast<int>(nb); w.tw = static_cast<int>(tw); w.th = static_cast<int>(th); w.n = static_cast<int>(n);
    return w;
}

template <int BN, int PL, int KC, int ROWS>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_tcx_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                const __grid_constant__ CUtensorMap tmW_hi, const __grid_constant__ CUtensorMap tmW_lo,
                const ConvTcParams p, const __grid_constant__ XGeom geo) {
    pdl_trigger();
    using C = XCfg<BN, PL, KC, ROWS>;
    constexpr int NA = C::NA, NB = C::NB;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + NA * C::A_SLOT;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + NB * C::B_SLOT);
    uint64_t* a_full = bars;
    uint64_t* a_empty = a_full + NA;
    uint64_t* b_full = a_empty + NA;
    uint64_t* b_empty = b_full + NB;
    uint64_t* t_full = b_empty + NB;
    uint64_t* t_empty = t_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kchunks = p.Cin / KC;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA_hi); ptx::prefetch_tmap(&tmW_hi);
        if (PL > 1) { ptx::prefetch_tmap(&tmA_lo); ptx::prefetch_tmap(&tmW_lo); }
        for (int i = 0; i < NA; ++i) { ptx::mbar_init(&a_full[i], 1); ptx::mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < NB; ++i) { ptx::mbar_init(&b_full[i], 1); ptx::mbar_init(&b_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&t_full[i], 1); ptx::mbar_init(&t_empty[i], EPI_WARPS); }
        ptx::fence_barrier_init();
    }
    if (warp == 1) { ptx::tmem_alloc(tmem_slot, C::TMEM_COLS); ptx::tmem_relinquish(); }
    pdl_wait();                              // from here on the previous kernel's results are visible
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < EPI_FIRST_WARP) {
    ptx::reg_dec<REGS_CONTROL>();          // the whole first warpgroup: producer, MMA issuer and two idle warps
    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int a = 0, b = 0; uint32_t aph = 0, bph = 0;
            for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x) {
                const XWork w = x_decode(wk, p, geo.per_cls);
                const XClass& g = geo.cls[w.cls];
                const int h0 = w.th * TILE_H * geo.cs, w0 = w.tw * TILE_W * geo.cs;
                const int wrow = w.n * p.w_img_rows + w.nb * BN;
                for (int c = 0; c < kchunks; ++c) {
                    for (int bx = 0; bx < g.nbox; ++bx) {
                        const XBox& B = g.box[bx];
                        ptx::mbar_wait(&a_empty[a], aph ^ 1);
                        ptx::mbar_expect_tx(&a_full[a], geo.a_tx);
                        uint8_t* dst = sA + a * C::A_SLOT;
                        ptx::tma_load_4d(dst, &tmA_hi, &a_full[a], c * KC, w0 + B.dw, h0 + B.dh, w.n);
                        if (PL > 1) ptx::tma_load_4d(dst + C::A_PLANE, &tmA_lo, &a_full[a], c * KC, w0 + B.dw, h0 + B.dh, w.n);
                        if (PL > 2) ptx::tma_load_4d(dst + 2 * C::A_PLANE, &tmA_lo, &a_full[a], c * KC, w0 + B.dw, h0 + B.dh, w.n + p.N);
                        if (++a == NA) { a = 0; aph ^= 1; }
                        for (int t = 0; t < B.ntap; ++t) {
                            ptx::mbar_wait(&b_empty[b], bph ^ 1);
                            ptx::mbar_expect_tx(&b_full[b], C::B_SLOT);
                            uint8_t* bd = sB + b * C::B_SLOT;
                            ptx::tma_load_3d(bd, &tmW_hi, &b_full[b], c * KC, B.widx[t], wrow);
                            if (PL > 1) ptx::tma_load_3d(bd + C::B_PLANE, &tmW_lo, &b_full[b], c * KC, B.widx[t], wrow);
                            if (PL > 2) ptx::tma_load_3d(bd + 2 * C::B_PLANE, &tmW_lo, &b_full[b], c * KC, B.widx[t], wrow + p.w_rows_total);
                            if (++b == NB) { b = 0; bph ^= 1; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (warp-uniform walk, one elected lane issues) =====================
        const bool leader = ptx::elect_one();
        const uint32_t fmt = p.fmt_bits;
        constexpr uint32_t desc_hi = C::DESC_HI;
        int a = 0, b = 0; uint32_t aph = 0, bph = 0;
        int it = 0;
        for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x, ++it) {
            const int acc = it & 1;
            const XClass& g = geo.cls[wk / geo.per_cls];
            ptx::mbar_wait(&t_empty[acc], ((it >> 1) & 1) ^ 1);
            ptx::tc_fence_after();
            const uint32_t d_main = tmem_base + acc * C::STAGE_COLS;
            uint32_t acc_main = 0;
            for (int c = 0; c < kchunks; ++c) {
                for (int bx = 0; bx < g.nbox; ++bx) {
                    const XBox& B = g.box[bx];
                    ptx::mbar_wait(&a_full[a], aph);
                    ptx::tc_fence_after();
                    const uint32_t a_lo0 = ptx::umma_desc_lo(ptx::smem_u32(sA + a * C::A_SLOT));
                    for (int t = 0; t < B.ntap; ++t) {
                        ptx::mbar_wait(&b_full[b], bph);
                        ptx::tc_fence_after();
                        const uint32_t b_lo0 = ptx::umma_desc_lo(ptx::smem_u32(sB + b * C::B_SLOT));
                        const uint32_t a_lo = a_lo0 + ((B.row_off[t] * TILE_W * C::ROWB) >> 4);
#pragma unroll
                        for (int k = 0; k < KC / 16; ++k) {
                            // planes of B are consecutive N rows in shared memory, so one MMA of N = 2*BN / 3*BN
                            // multiplies a plane of A with several planes of B (A is read once, not per pass):
                            //   D0 = a0 w0 | D1 = a0 w1 + a1 w0 | D2 = a0 w2 + a1 w1 + a2 w0   (adjacent TMEM columns)
                            const uint32_t ak = a_lo + ((k * 32) >> 4), bk = b_lo0 + ((k * 32) >> 4);
                            if (PL == 1) {
                                if (leader) ptx::mma_bf16_ss2(d_main, ak, bk, desc_hi, (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt), acc_main);
                            } else if (PL == 2) {
                                if (leader) {
                                    ptx::mma_bf16_ss2(d_main, ak, bk, desc_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 2 * BN) | fmt), acc_main);
                                    ptx::mma_bf16_ss2(d_main + BN, ak + (C::A_PLANE >> 4), bk, desc_hi, (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt), 1u);
                                }
                            } else if (C::NACC == 3) {
                                if (leader) {
                                    ptx::mma_bf16_ss2(d_main, ak, bk, desc_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 3 * BN) | fmt), acc_main);
                                    ptx::mma_bf16_ss2(d_main + BN, ak + (C::A_PLANE >> 4), bk, desc_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 2 * BN) | fmt), 1u);
                                    ptx::mma_bf16_ss2(d_main + 2 * BN, ak + ((2 * C::A_PLANE) >> 4), bk, desc_hi, (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt), 1u);
                                }
                            } else {
                                // 3 planes, BN = 128: N = 256 is the widest MMA; the 2^-16 terms share the D1 accumulator
                                if (leader) {
                                    const uint32_t id1 = (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt);
                                    ptx::mma_bf16_ss2(d_main, ak, bk, desc_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 2 * BN) | fmt), acc_main);
                                    ptx::mma_bf16_ss2(d_main + BN, ak, bk + ((2 * C::B_PLANE) >> 4), desc_hi, id1, 1u);
                                    ptx::mma_bf16_ss2(d_main + BN, ak + (C::A_PLANE >> 4), bk + (C::B_PLANE >> 4), desc_hi, id1, 1u);
                                    ptx::mma_bf16_ss2(d_main + BN, ak + ((2 * C::A_PLANE) >> 4), bk, desc_hi, id1, 1u);
                                    ptx::mma_bf16_ss2(d_main + BN, ak + (C::A_PLANE >> 4), bk, desc_hi, id1, 1u);
                                }
                            }
                            acc_main = 1;
                        }
                        if (leader) ptx::tc_commit(&b_empty[b]);
                        __syncwarp();
                        if (++b == NB) { b = 0; bph ^= 1; }
                    }
                    if (leader) ptx::tc_commit(&a_empty[a]);
                    if (++a == NA) { a = 0; aph ^= 1; }
                }
            }
            if (leader) ptx::tc_commit(&t_full[acc]);
            __syncwarp();
        }
    }
    } else {
        // ===================== epilogue (warps 4..11) =====================
        ptx::reg_inc<REGS_EPILOGUE>();
        auto dec = [&](int wk, int m) {
            const XWork w = x_decode(wk, p, geo.per_cls);
            const XClass& g = geo.cls[w.cls];
            const int oh = (w.th * TILE_H + m / TILE_W) * geo.os + g.oa, ow = (w.tw * TILE_W + m % TILE_W) * geo.os + g.ob;
            TcPix px;
            px.nb = w.nb;
            px.valid = (oh < geo.OH) && (ow < geo.OW);
            px.pix = (static_cast<size_t>(w.n) * geo.OH + oh) * geo.OW + ow;
            return px;
        };
        tc_epilogue_warps<BN, C::NACC>(p, p.bias, tmem_base, C::STAGE_COLS, t_full, t_empty, warp, lane, dec);
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
}

template <int BN, int PL, int KC, int ROWS = X_ROWS>
int xlaunch(const CUtensorMap& ah, const CUtensorMap& al, const CUtensorMap& wh, const CUtensorMap& wl,
            const ConvTcParams& p, const XGeom& g, cudaStream_t stream) {
    using C = XCfg<BN, PL, KC, ROWS>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_tcx_kernel<BN, PL, KC, ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { st_set_error("conv_tcx smem attribute: %s", cudaGetErrorString(e)); return ST_ERR_CUDA; }
        attr_set = true;
    }
    const int grid = p.work < st_num_sms() ? p.work : st_num_sms();
    st_launch(grid, NUM_THREADS, C::SMEM, stream)(conv_tcx_kernel<BN, PL, KC, ROWS>, ah, al, wh, wl, p, g);
    ST_CHECK_LAUNCH("conv_tcx_kernel");
    return ST_OK;
}

template <int BN, int KC>
int xlaunch_planes(int planes, const CUtensorMap& ah, const CUtensorMap& al, const CUtensorMap& wh, const CUtensorMap& wl,
                   const ConvTcParams& p, const XGeom& g, cudaStream_t stream) {
    if (planes == 3) return xlaunch<BN, 3, KC>(ah, al, wh, wl, p, g, stream);
    if (planes == 2) return xlaunch<BN, 2, KC>(ah, al, wh, wl, p, g, stream);
    return xlaunch<BN, 1, KC>(ah, al, wh, wl, p, g, stream);
}


void x_add_box(XClass& c, int dw, int dh) { XBox& b = c.box[c.nbox++]; b.dw = dw; b.dh = dh; b.ntap = 0; }
void x_add_tap(XClass& c, int row_off, int widx) { XBox& b = c.box[c.nbox - 1]; b.row_off[b.ntap] = row_off; b.widx[b.ntap] = widx; ++b.ntap; ++c.taps; }

}  // namespace

// mode 0: 3x3 stride 1 pad 1 (weight tap r*3+s reads pixel + (r-1, s-1))
// mode 1: 3x3 stride 2 pad 1 forward
// mode 2: backward-data of mode 1 (weights packed rows = ci, cols = co, unflipped taps), Ho = 2H, Wo = 2W
// mode 3: vertical 9-tap conv (pad 4), 32 -> 32 channels, fp32 output [pixel][32] (first stage of the
//         32 -> 3 9x9 layer: weight row co*9 + s, tap r; see conv9_reduce_tc_run)
bool conv_tcx_eligible(int mode, int ks, int Cin, int Cout, int H, int W, int Ho, int Wo) {
    if (mode == 3) return ks == 9 && Cin == 32 && Cout == 32 && Ho == H && Wo == W;
    if (ks != 3 || Cin % 32 != 0 || !(Cout % 128 == 0 || Cout == 64 || Cout == 32)) return false;
    if (mode == 0) return Ho == H && Wo == W;
    if (mode == 1) return Ho == (H - 1) / 2 + 1 && Wo == (W - 1) / 2 + 1;
    return Ho == 2 * H && Wo == 2 * W;
}

// x_planes / w_planes / out_planes / add_code are plane CODES (common.cuh); acc_mul, acc_mul_slot, add_mul_slot: see ConvTcParams
int conv_tcx_run(int mode, const bf16* x_hi, const bf16* x_lo, int x_planes, int N, int H, int W, int Cin,
                 const bf16* w_hi, const bf16* w_lo, int w_planes, int w_rows_total, int Cout, int Ho, int Wo,
                 const float* bias, const bf16* mask, const bf16* add_hi, const bf16* add_lo, int add_code,
                 bf16* out_hi, bf16* out_lo, bf16* relu_hi, bf16* relu_lo, int out_planes, float* out_f32,
                 float acc_mul, const float* acc_mul_slot, const float* add_mul_slot, cudaStream_t stream) {
    ST_CHECK_ARG(pc_f16(x_planes) == pc_f16(w_planes), "conv_tcx: activation and weight planes must have the same 16-bit format");
    const int BN = (Cout % 128 == 0) ? 128 : Cout;
    const int KC = (Cin % 64 == 0) ? 64 : 32;
    const int IH = mode == 1 ? Ho : H, IW = mode == 1 ? Wo : W;      // tile iteration space
    XGeom g; memset(&g, 0, sizeof(g));
    g.OH = Ho; g.OW = Wo; g.cs = mode == 1 ? 2 : 1; g.os = mode == 2 ? 2 : 1;
    int rows;
    if (mode == 0) {
        g.ncls = 1; rows = TILE_H + 2;
        for (int s = 0; s < 3; ++s) {
            x_add_box(g.cls[0], s - 1, -1);
            for (int r = 0; r < 3; ++r) x_add_tap(g.cls[0], r, r * 3 + s);
        }
    } else if (mode == 3) {
        g.ncls = 1; rows = X_ROWS9;
        x_add_box(g.cls[0], 0, -4);
        for (int r = 0; r < 9; ++r) x_add_tap(g.cls[0], r, r);
    } else if (mode == 1) {
        // input row 2*oh + r - 1: r = 1 reads even rows, r = 0 / 2 the odd rows oh - 1 / oh of one strided box
        g.ncls = 1; rows = TILE_H + 1;
        for (int s = 0; s < 3; ++s) {
            x_add_box(g.cls[0], s - 1, 0);  x_add_tap(g.cls[0], 0, 3 + s);
            x_add_box(g.cls[0], s - 1, -1); x_add_tap(g.cls[0], 0, s); x_add_tap(g.cls[0], 1, 6 + s);
        }
    } else {
        // dx[2i+a][2j+b] = sum over r = a+1 (mod 2), s = b+1 (mod 2) of dy[i + (a+1-r)/2][j + (b+1-s)/2] * w[r][s]
        g.ncls = 4; rows = TILE_H + 1;
        const int order[4][2] = {{1, 1}, {1, 0}, {0, 1}, {0, 0}};
        for (int k = 0; k < 4; ++k) {
            const int a = order[k][0], b = order[k][1];
            XClass& c = g.cls[k]; c.oa = a; c.ob = b;
            for (int s = (b + 1) & 1; s < 3; s += 2) {
                x_add_box(c, (b + 1 - s) / 2, 0);
                for (int r = (a + 1) & 1; r < 3; r += 2) x_add_tap(c, (a + 1 - r) / 2, r * 3 + s);
            }
        }
    }
    ConvTcParams p;
    p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout;
    p.tiles_h = (IH + TILE_H - 1) / TILE_H; p.tiles_w = (IW + TILE_W - 1) / TILE_W;
    p.n_blocks = Cout / BN;
    g.per_cls = N * p.tiles_h * p.tiles_w * p.n_blocks; p.work = g.per_cls * g.ncls;
    p.fd_nb.init(p.n_blocks); p.fd_tw.init(p.tiles_w); p.fd_th.init(p.tiles_h); p.fd_W.init(p.W > 0 ? p.W : 1); p.fd_H.init(p.H > 0 ? p.H : 1);
    p.w_img_rows = 0; p.w_rows_total = w_rows_total;
    const int xp = pc_planes(x_planes), wp = pc_planes(w_planes);
    const int a_planes = x_lo ? (xp >= 3 ? 3 : (xp >= 2 ? 2 : 1)) : 1;
    const int w_planes_ = w_lo ? (wp >= 3 ? 3 : (wp >= 2 ? 2 : 1)) : 1;
    const int planes = a_planes < w_planes_ ? a_planes : w_planes_;
    g.a_tx = planes * rows * TILE_W * KC * 2;
    p.fmt_bits = ptx::umma_fmt_bits(pc_f16(x_planes));
    p.out_planes = out_planes; p.add_code = add_code;
    p.acc_mul = acc_mul; p.acc_mul_slot = acc_mul_slot; p.add_mul_slot = add_mul_slot;
    p.bias = bias; p.mask = mask; p.add_hi = add_hi; p.add_lo = add_lo;
    p.out_hi = out_hi; p.out_lo = out_lo; p.relu_hi = relu_hi; p.relu_lo = relu_lo;
    p.out_nchw = nullptr; p.nchw_c = 0; p.out_f32 = out_f32; p.dbg = 0;
    p.add_route = nullptr; p.route_Ho = p.route_Wo = 0;
    p.pool_hi = p.pool_lo = nullptr; p.pool_idx = nullptr; p.pool_lane_h = 0; p.pool_Ho = p.pool_Wo = 0;

    CUtensorMap ah, al, wh, wl;
    const int ps = mode == 1 ? 2 : 1;
    uint64_t ad[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t ad2[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)N * (planes == 3 ? 2 : 1)};
    uint32_t ab[4] = {(uint32_t)KC, (uint32_t)(TILE_W * ps), (uint32_t)(rows * ps), 1};
    uint64_t wd[3] = {(uint64_t)Cin, 9, (uint64_t)w_rows_total};
    uint64_t wd2[3] = {(uint64_t)Cin, 9, (uint64_t)w_rows_total * (planes == 3 ? 2 : 1)};
    uint32_t wb[3] = {(uint32_t)KC, 1, (uint32_t)BN};
    int rc;
    if ((rc = st_make_tmap(&ah, x_hi, 4, ad, ab, KC * 2, ps))) return rc;
    if ((rc = st_make_tmap(&wh, w_hi, 3, wd, wb, KC * 2))) return rc;
    al = ah; wl = wh;
    if (planes > 1 && (rc = st_make_tmap(&al, x_lo, 4, ad2, ab, KC * 2, ps))) return rc;
    if (planes > 1 && (rc = st_make_tmap(&wl, w_lo, 3, wd2, wb, KC * 2))) return rc;
    if (mode == 3) {
        if (planes == 3) return xlaunch<32, 3, 32, X_ROWS9>(ah, al, wh, wl, p, g, stream);
        if (planes == 2) return xlaunch<32, 2, 32, X_ROWS9>(ah, al, wh, wl, p, g, stream);
        return xlaunch<32, 1, 32, X_ROWS9>(ah, al, wh, wl, p, g, stream);
    }
    if (KC == 64) {
        if (BN == 128) return xlaunch_planes<128, 64>(planes, ah, al, wh, wl, p, g, stream);
        if (BN == 64) return xlaunch_planes<64, 64>(planes, ah, al, wh, wl, p, g, stream);
        if (BN == 32) return xlaunch_planes<32, 64>(planes, ah, al, wh, wl, p, g, stream);
    } else {
        if (BN == 128) return xlaunch_planes<128, 32>(planes, ah, al, wh, wl, p, g, stream);
        if (BN == 64) return xlaunch_planes<64, 32>(planes, ah, al, wh, wl, p, g, stream);
        if (BN == 32) return xlaunch_planes<32, 32>(planes, ah, al, wh, wl, p, g, stream);
    }
    st_set_error("conv_tcx: unsupported tile configuration BN=%d KC=%d", BN, KC);
    return ST_ERR_UNSUPPORTED;
}
