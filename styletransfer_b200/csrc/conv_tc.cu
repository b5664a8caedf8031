// tcgen05 / TMEM / TMA implicit-GEMM convolution for sm_100a (stride 1, 3x3 pad 1
// or 1x1), NHWC split-bf16 activations.  This one kernel serves
//   * conv forward of the VGG prefix and of the transform net's 3x3 layers
//     (reference: nn.Conv2d call sites network.py:246-275, :468-481, :582-599),
//   * their backward-data (dgrad = the same conv with flipped/transposed weights,
//     ReLU mask and the loss gradient of the tap fused into the epilogue),
//   * the Gram backward dF = F * D (a 1x1 conv with per-image weights D,
//     reference: autograd of network.py:92-123).
//
// GEMM view: M = 128 output pixels (3x3: a 16 x 8 patch), N = BN output channels,
// K = taps * Cin.  ONE TMA box load of an (16+2) x (8+2) pixel x 64 channel halo
// patch per plane feeds all nine taps (see Cfg).  Zero padding comes from TMA
// out-of-bounds fill.
//
// Warp roles (384 threads, persistent, one CTA per SM):
//   warp 0  TMA producer      warp 1  tcgen05.mma issuer (+ TMEM alloc)
//   warps 4-11 epilogue (two per TMEM lane quarter; warps 2-3 idle, registers handed over by setmaxnreg): tcgen05.ld -> bias / mask / add -> split-bf16 stores
// Pipelines: A ring (halo boxes), B ring (weight tiles), 2 TMEM accumulators.
#include "common.cuh"
#include "ptx.cuh"
#ifdef ST_CONV_TRACE
__device__ long long g_conv_trace[3 * 64 * 8] = {};
#define TC_TRACE(role, tile, ev) do { if (blockIdx.x == 0 && (tile) < 64) g_conv_trace[((role) * 64 + (tile)) * 8 + (ev)] = clock64(); } while (0)
#endif
#include "conv_tc_epilogue.cuh"
#include <mutex>
#include <unordered_map>
#include <string>
#include <string.h>

extern int g_conv_debug;

// Per-tile timeline of CTA 0 (tools/trace_conv.py; the library is built with -DST_CONV_TRACE for that run only):
// g_conv_trace[role][tile][event] = clock64().  roles: 0 producer, 1 MMA issuer, 2 epilogue warp 4.
#ifdef ST_CONV_TRACE
#define TRACE(role, tile, ev) do { if (blockIdx.x == 0 && (tile) < 64) g_conv_trace[((role) * 64 + (tile)) * 8 + (ev)] = clock64(); } while (0)
extern "C" int st_conv_trace_read(long long* out) {
    return cudaMemcpyFromSymbol(out, g_conv_trace, sizeof(g_conv_trace)) == cudaSuccess ? 0 : -2;
}
#else
#define TRACE(role, tile, ev) do { } while (0)
#endif

namespace {

constexpr int KCH = 64;

// KS: 1 or 3 taps per side; BN: output channels per tile; PL: 16-bit planes of BOTH operands
// (1 -> 1 MMA per product, 2 -> 3, 3 -> 6); RES: resident-weights schedule (see below).
//
// Geometry.  3x3: an output tile is 16 rows x 8 pixels (M row m = h*8 + w) and ONE TMA box of (16+2) x (8+2) pixels x
// 64 channels per plane and channel chunk feeds all nine taps: tap (r, s) is the same box read from row (r*10 + s) on,
// with a stride of 10 rows (1280 B) between the 8-row groups of the operand -- tcgen05.mma applies the 128-byte
// swizzle by absolute shared-memory address, exactly as TMA wrote it, so neither the start row nor the group stride
// has to be a multiple of the 1024-byte swizzle period (ptx::umma_desc_hi_k_sw128_sbo).  Per 128 output pixels that
// is 45 KB of L2 -> shared-memory traffic for two planes instead of the 120 KB of one 10 x 16 box per horizontal tap.
// 1x1: tile 8 rows x 16 pixels, dense box, group stride 1024 B.
//
// Schedules.  General: an A slot holds all planes of one (tile, chunk) box; the weight tiles of the taps stream
// through the B ring.  RES (PL = 2, one channel chunk, one channel block, shared weights): the nine weight tiles stay
// in shared memory for the life of the CTA, which leaves room for three PLANE-sized A slots; a tile first runs the
// 36 wide MMAs of its hi plane (a0 x [w0|w1]), then the 36 of its lo plane (a1 x w0), so plane slots are filled and
// released in one fixed order and the loads run up to a whole tile ahead of the MMAs.
template <int KS, int BN, int PL, bool RES> struct Cfg {
    static constexpr int TH = KS == 3 ? 16 : 8, TW = KS == 3 ? 8 : 16;
    static constexpr int BOX_W = TW + KS - 1, BOX_H = TH + KS - 1;
    static constexpr int A_BYTES = BOX_W * BOX_H * 128;               // one plane of one box (64 channels per pixel row)
    static constexpr int A_PLANE = (A_BYTES + 1023) & ~1023;          // slots stay 1024-byte aligned
    static constexpr int A_SBO = KS == 3 ? BOX_W * 128 : 1024;
    static constexpr int B_PLANE = BN * 128;
    static constexpr int A_SLOT = RES ? A_PLANE : PL * A_PLANE, B_SLOT = PL * B_PLANE;
    static constexpr int A_TX = RES ? A_BYTES : PL * A_BYTES;
    static constexpr int BIAS_MAX = 512;                    // output channels whose bias is staged in shared memory
    static constexpr int BAR_BYTES = 256 + BIAS_MAX * 4;    // barriers + TMEM slot, then the bias
    static constexpr int SMEM_MAX = 232448;                 // 227 KB opt-in limit
    static constexpr int TAPS = KS * KS;
    // 1x1 launches (the Gram backward) are bound by their epilogue: their stores go through a shared-memory tile per
    // epilogue warp (tc_group, TS).  Measured on the 3x3 general-schedule launches too (where the tile costs one stage of
    // the weight ring): slower, Gatys closure 0.674 -> 0.690 ms, fast_st step 8.57 -> 8.98 ms -- they hide their epilogue
    // under the MMA phase, so 3x3 keeps the direct stores.
    static constexpr bool TS = KS == 1 && BN >= 64 && PL <= 2;
    static constexpr int STG_BYTES = TS ? EPI_WARPS * TS_WARP_BYTES : 0;
    static constexpr int NA = RES ? 3 : (KS == 1 && PL < 3) ? (BN == 128 ? 3 : 4) : 2;
    static constexpr int NB_FIT = (SMEM_MAX - 1024 - BAR_BYTES - STG_BYTES - NA * A_SLOT) / B_SLOT;
    static constexpr int NB = RES ? TAPS : (NB_FIT > 9 ? 9 : NB_FIT);
    static_assert(NB >= 2 && NB <= NB_FIT, "weight ring does not fit");
    static_assert(!RES || (PL == 2 && KS == 3), "resident schedule: two planes, 3x3");
    static constexpr int SMEM = 1024 + NA * A_SLOT + NB * B_SLOT + BAR_BYTES + STG_BYTES;
    // PL > 1: the a0*w0 products go to a "main" accumulator, all correction terms to a second
    // one, summed in fp32 by the epilogue.  The tensor core's accumulate truncates; keeping the
    // ~2^-8 smaller corrections out of the main chain divides that error by the pass count.
    static constexpr int NACC = PL == 1 ? 1 : (PL == 3 && BN <= 64) ? 3 : 2;
    static constexpr int STAGE_COLS = NACC * BN;
    static constexpr int TMEM_COLS = (2 * STAGE_COLS <= 32) ? 32 : (2 * STAGE_COLS <= 64) ? 64 :
                                     (2 * STAGE_COLS <= 128) ? 128 : (2 * STAGE_COLS <= 256) ? 256 : 512;
    static_assert(2 * STAGE_COLS <= 512, "accumulators exceed TMEM");
    static_assert(2 * NA + 2 * NB + 4 <= 256 / 8 - 1, "barrier area");
};

template <int KS, int BN, int PL, bool RES, bool FUSED>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
               const __grid_constant__ CUtensorMap tmW_hi, const __grid_constant__ CUtensorMap tmW_lo,
               const ConvTcParams p) {
    pdl_trigger();
    using C = Cfg<KS, BN, PL, RES>;
    constexpr int NA = C::NA, NB = C::NB;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + NA * C::A_SLOT;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + NB * C::B_SLOT);
    uint64_t* a_full = bars;                 // [NA]
    uint64_t* a_empty = a_full + NA;         // [NA]
    uint64_t* b_full = a_empty + NA;         // [NB]
    uint64_t* b_empty = b_full + NB;         // [NB]
    uint64_t* t_full = b_empty + NB;         // [2]
    uint64_t* t_empty = t_full + 2;          // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);
    float* sbias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);
    uint8_t* ts_stage = reinterpret_cast<uint8_t*>(bars) + C::BAR_BYTES;       // C::STG_BYTES, 256-byte aligned

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int pad = KS / 2;
    const int kchunks = p.Cin / KCH;
    const bool bias_staged = p.bias && p.Cout <= C::BIAS_MAX;

    // set-up that touches no global memory runs while the previous kernel of the stream drains (pdl_trigger above)
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
    // the bias of every output channel, read by the epilogue as shared-memory broadcasts
    if (bias_staged)
        for (int i = threadIdx.x; i < p.Cout; i += NUM_THREADS) sbias[i] = p.bias[i];
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
                uint32_t t, unb, utw, uth, un;
                p.fd_nb.divmod(static_cast<uint32_t>(wk), t, unb);
                p.fd_tw.divmod(t, t, utw);
                p.fd_th.divmod(t, un, uth);
                const int nb = static_cast<int>(unb), tw = static_cast<int>(utw), th = static_cast<int>(uth), n = static_cast<int>(un);
                const int h0 = th * C::TH - pad, w0 = tw * C::TW - pad;
                const int wrow = n * p.w_img_rows + nb * BN;
                if (RES) {
                    if (wk == static_cast<int>(blockIdx.x)) {
                        // the nine weight tiles: loaded once, never released
#pragma unroll
                        for (int tap = 0; tap < C::TAPS; ++tap) {
                            ptx::mbar_expect_tx(&b_full[tap], C::B_SLOT);
                            ptx::tma_load_3d(sB + tap * C::B_SLOT, &tmW_hi, &b_full[tap], 0, tap, wrow);
                            ptx::tma_load_3d(sB + tap * C::B_SLOT + C::B_PLANE, &tmW_lo, &b_full[tap], 0, tap, wrow);
                        }
                    }
#pragma unroll
                    for (int pl = 0; pl < 2; ++pl) {
                        ptx::mbar_wait(&a_empty[a], aph ^ 1);
                        TRACE(0, (wk - blockIdx.x) / gridDim.x, pl);
                        if (p.dbg & 4) { ptx::mbar_arrive(&a_full[a]); if (++a == NA) { a = 0; aph ^= 1; } continue; }
                        ptx::mbar_expect_tx(&a_full[a], C::A_TX);
                        ptx::tma_load_4d(sA + a * C::A_SLOT, pl ? &tmA_lo : &tmA_hi, &a_full[a], 0, w0, h0, n);
                        if (++a == NA) { a = 0; aph ^= 1; }
                    }
                    continue;
                }
                for (int c = 0; c < kchunks; ++c) {
                    ptx::mbar_wait(&a_empty[a], aph ^ 1);
                    if (c == 0) TRACE(0, (wk - blockIdx.x) / gridDim.x, 0);
                    ptx::mbar_expect_tx(&a_full[a], C::A_TX);
                    uint8_t* dst = sA + a * C::A_SLOT;
                    // planes 1 and 2 of a tensor are equally spaced: the "lo" map spans both
                    ptx::tma_load_4d(dst, &tmA_hi, &a_full[a], c * KCH, w0, h0, n);
                    if (PL > 1) ptx::tma_load_4d(dst + C::A_PLANE, &tmA_lo, &a_full[a], c * KCH, w0, h0, n);
                    if (PL > 2) ptx::tma_load_4d(dst + 2 * C::A_PLANE, &tmA_lo, &a_full[a], c * KCH, w0, h0, n + p.N);
                    if (++a == NA) { a = 0; aph ^= 1; }
#pragma unroll 1
                    for (int tap = 0; tap < C::TAPS; ++tap) {
                        ptx::mbar_wait(&b_empty[b], bph ^ 1);
                        ptx::mbar_expect_tx(&b_full[b], C::B_SLOT);
                        uint8_t* bd = sB + b * C::B_SLOT;
                        ptx::tma_load_3d(bd, &tmW_hi, &b_full[b], c * KCH, tap, wrow);
                        if (PL > 1) ptx::tma_load_3d(bd + C::B_PLANE, &tmW_lo, &b_full[b], c * KCH, tap, wrow);
                        if (PL > 2) ptx::tma_load_3d(bd + 2 * C::B_PLANE, &tmW_lo, &b_full[b], c * KCH, tap, wrow + p.w_rows_total);
                        if (++b == NB) { b = 0; bph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        // the whole warp walks the pipeline state (warp-uniform -> uniform registers); one elected
        // lane issues the tcgen05 instructions
        {
            const bool leader = ptx::elect_one();
            const uint32_t fmt = p.fmt_bits;
            constexpr uint32_t a_hi = ptx::umma_desc_hi_k_sw128_sbo(C::A_SBO), b_hi = ptx::umma_desc_hi_k_sw128();
            int a = 0, b = 0; uint32_t aph = 0, bph = 0;
            int it = 0;
            // barriers are polled one step AHEAD (mbar_test before a group of MMAs is issued, see ptx.cuh); the flags
            // say that the barrier the next step needs has already been seen complete
            bool t_ready = false, a_ready = false, b_ready = false;
            for (int wk = blockIdx.x; wk < p.work; wk += gridDim.x, ++it) {
                const int acc = it & 1;
                const bool more = wk + static_cast<int>(gridDim.x) < p.work;
                if (!t_ready) ptx::mbar_wait(&t_empty[acc], ((it >> 1) & 1) ^ 1);
                t_ready = false;
                ptx::tc_fence_after();
                const uint32_t d_main = tmem_base + acc * C::STAGE_COLS;
                uint32_t acc_main = 0;
                TRACE(1, it, 0);
                if (RES) {
                    // rolled on purpose: a fully unrolled tile is > 1000 SASS instructions of descriptor arithmetic that
                    // spill uniform registers and miss the instruction cache
                    if (it == 0) {
#pragma unroll 1
                        for (int tap = 0; tap < C::TAPS; ++tap) ptx::mbar_wait(&b_full[tap], 0);
                        ptx::tc_fence_after();
                    }
#pragma unroll 1
                    for (int pl = 0; pl < 2; ++pl) {
                        if (!a_ready) ptx::mbar_wait(&a_full[a], aph);
                        a_ready = false;
                        ptx::tc_fence_after();
                        TRACE(1, it, 1 + 2 * pl);
                        const uint32_t idesc = ((p.dbg & 8) ? ptx::umma_idesc_nofmt(BLOCK_M, 16) : ptx::umma_idesc_nofmt(BLOCK_M, pl == 0 ? 2 * BN : BN)) | fmt;
                        const uint32_t d = pl == 0 ? d_main : d_main + BN;
                        uint32_t a_lo = ptx::umma_desc_lo(ptx::smem_u32(sA + a * C::A_SLOT));
                        uint32_t b_lo = ptx::umma_desc_lo(ptx::smem_u32(sB));
                        const int an = (a + 1 == NA) ? 0 : a + 1;
                        const uint32_t anph = (a + 1 == NA) ? aph ^ 1 : aph;
#pragma unroll 1
                        for (int r = 0; r < KS; ++r, a_lo += (C::BOX_W * 128) >> 4, b_lo += (KS * C::B_SLOT) >> 4) {
                            if (r == KS - 1 && (pl == 0 || more)) {
                                a_ready = ptx::mbar_test(&a_full[an], anph);
                                if (pl == 1) t_ready = ptx::mbar_test(&t_empty[acc ^ 1], (((it + 1) >> 1) & 1) ^ 1);
                            }
                            if (leader) {
                                // one filter row = 12 MMAs per elected-lane region; the commits that follow the last row
                                // ride in the same region (every instruction between two groups of MMAs is tensor-pipe idle time)
                                if (!(p.dbg & 2)) {
#pragma unroll
                                    for (int sx = 0; sx < KS; ++sx)
#pragma unroll
                                        for (int k = 0; k < KCH / 16; ++k)
                                            ptx::mma_ss2h(d + ((p.dbg & 16) ? (k & 3) * 16 : 0), a_lo + ((sx * 128 + k * 32) >> 4), a_hi, b_lo + ((sx * C::B_SLOT + k * 32) >> 4), b_hi, idesc,
                                                          (pl == 1 || sx > 0 || k > 0) ? 1u : acc_main);
                                }
                                if (r == KS - 1) {
                                    ptx::tc_commit(&a_empty[a]);
                                    if (pl == 1) ptx::tc_commit(&t_full[acc]);
                                }
                            }
                            acc_main = 1;
                        }
                        TRACE(1, it, 2 + 2 * pl);
                        a = an; aph = anph;
                    }
                    continue;
                }
                for (int c = 0; c < kchunks; ++c) {
                    if (!a_ready) ptx::mbar_wait(&a_full[a], aph);
                    a_ready = false;
                    ptx::tc_fence_after();
                    if (c == 0) TRACE(1, it, 1);
                    const uint32_t a_lo0 = ptx::umma_desc_lo(ptx::smem_u32(sA + a * C::A_SLOT));
                    const int an = (a + 1 == NA) ? 0 : a + 1;
                    const uint32_t anph = (a + 1 == NA) ? aph ^ 1 : aph;
#pragma unroll 1
                    for (int tap = 0; tap < C::TAPS; ++tap) {
                        if (!b_ready) ptx::mbar_wait(&b_full[b], bph);
                        b_ready = false;
                        ptx::tc_fence_after();
                        const uint32_t b_lo0 = ptx::umma_desc_lo(ptx::smem_u32(sB + b * C::B_SLOT));
                        const int r = tap / KS, s = tap - r * KS;
                        const uint32_t a_lo = a_lo0 + (((r * C::BOX_W + s) * 128) >> 4);
                        // what the NEXT step needs: the next weight tile; after the last tap of a chunk also the next
                        // activation slot, after the last chunk of a tile also the other accumulator stage
                        const int bn = (b + 1 == NB) ? 0 : b + 1;
                        const uint32_t bnph = (b + 1 == NB) ? bph ^ 1 : bph;
                        const bool last_tap = tap == C::TAPS - 1, last_chunk = c == kchunks - 1;
                        if (!(last_tap && last_chunk) || more) {
                            b_ready = ptx::mbar_test(&b_full[bn], bnph);
                            if (last_tap) {
                                a_ready = ptx::mbar_test(&a_full[an], anph);
                                if (last_chunk) t_ready = ptx::mbar_test(&t_empty[acc ^ 1], (((it + 1) >> 1) & 1) ^ 1);
                            }
                        }
                        // planes of B are consecutive N rows in shared memory, so one MMA of N = 2*BN / 3*BN
                        // multiplies a plane of A with several planes of B (A is read once, not per pass):
                        //   D0 = a0 w0 | D1 = a0 w1 + a1 w0 | D2 = a0 w2 + a1 w1 + a2 w0   (adjacent TMEM columns)
                        if (leader) {
#pragma unroll
                            for (int k = 0; k < KCH / 16; ++k) {
                                const uint32_t ak = a_lo + ((k * 32) >> 4), bk = b_lo0 + ((k * 32) >> 4);
                                const uint32_t acc0 = k > 0 ? 1u : acc_main;
                                if (PL == 1) {
                                    ptx::mma_ss2h(d_main, ak, a_hi, bk, b_hi, (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt), acc0);
                                } else if (PL == 2) {
                                    ptx::mma_ss2h(d_main, ak, a_hi, bk, b_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 2 * BN) | fmt), acc0);
                                    ptx::mma_ss2h(d_main + BN, ak + (C::A_PLANE >> 4), a_hi, bk, b_hi, (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt), 1u);
                                } else if (C::NACC == 3) {
                                    ptx::mma_ss2h(d_main, ak, a_hi, bk, b_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 3 * BN) | fmt), acc0);
                                    ptx::mma_ss2h(d_main + BN, ak + (C::A_PLANE >> 4), a_hi, bk, b_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 2 * BN) | fmt), 1u);
                                    ptx::mma_ss2h(d_main + 2 * BN, ak + ((2 * C::A_PLANE) >> 4), a_hi, bk, b_hi, (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt), 1u);
                                } else {
                                    // 3 planes, BN = 128: N = 256 is the widest MMA; the 2^-16 terms share the D1 accumulator
                                    const uint32_t id1 = (ptx::umma_idesc_nofmt(BLOCK_M, BN) | fmt);
                                    ptx::mma_ss2h(d_main, ak, a_hi, bk, b_hi, (ptx::umma_idesc_nofmt(BLOCK_M, 2 * BN) | fmt), acc0);
                                    ptx::mma_ss2h(d_main + BN, ak, a_hi, bk + ((2 * C::B_PLANE) >> 4), b_hi, id1, 1u);
                                    ptx::mma_ss2h(d_main + BN, ak + (C::A_PLANE >> 4), a_hi, bk + (C::B_PLANE >> 4), b_hi, id1, 1u);
                                    ptx::mma_ss2h(d_main + BN, ak + ((2 * C::A_PLANE) >> 4), a_hi, bk, b_hi, id1, 1u);
                                    ptx::mma_ss2h(d_main + BN, ak + (C::A_PLANE >> 4), a_hi, bk, b_hi, id1, 1u);
                                }
                            }
                            ptx::tc_commit(&b_empty[b]);
                            if (last_tap) {
                                ptx::tc_commit(&a_empty[a]);
                                if (last_chunk) ptx::tc_commit(&t_full[acc]);
                            }
                        }
                        acc_main = 1;
                        b = bn; bph = bnph;
                    }
                    a = an; aph = anph;
                    if (c == kchunks - 1) TRACE(1, it, 2);
                }
            }
        }
    }
    } else {
        // ===================== epilogue (warps 4..11) =====================
        ptx::reg_inc<REGS_EPILOGUE>();
        auto dec = [&](int wk, int m) {
            TcPix px;
            uint32_t t, unb, utw, uth, un;
            p.fd_nb.divmod(static_cast<uint32_t>(wk), t, unb);
            p.fd_tw.divmod(t, t, utw);
            p.fd_th.divmod(t, un, uth);
            px.nb = static_cast<int>(unb);
            const int tw = static_cast<int>(utw), th = static_cast<int>(uth), n = static_cast<int>(un);
            const int h = th * C::TH + m / C::TW, w = tw * C::TW + m % C::TW;
            px.valid = (h < p.H) && (w < p.W);
            px.pix = (static_cast<size_t>(n) * p.H + h) * p.W + w;
            return px;
        };
        tc_epilogue_warps<BN, C::NACC, FUSED ? 1 : 0, C::TS>(p, bias_staged ? sbias : p.bias, tmem_base, C::STAGE_COLS, t_full, t_empty, warp, lane, dec, ts_stage);
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
}

// ---------------------------------------------------------------------------
// host: tensor maps
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

struct MapKey {
    const void* ptr; uint64_t d[4]; uint32_t box[4]; int rank;
    bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
        size_t h = 1469598103934665603ull;
        for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
        return h;
    }
};
std::mutex g_map_mu;
std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;

}  // namespace

// bf16 tensor, dims innermost first, contiguous; 128B (or 64B) swizzle, zero OOB fill.
// Shared with wgrad_tc.cu (declared in tmap.cuh).
int st_make_tmap(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint32_t* box, int swizzle_bytes, int pixel_stride) {
    MapKey key; memset(&key, 0, sizeof(key));
    key.ptr = ptr; key.rank = rank | (swizzle_bytes << 8) | (pixel_stride << 16);
    for (int i = 0; i < rank; ++i) { key.d[i] = dims[i]; key.box[i] = box[i]; }
    std::lock_guard<std::mutex> lk(g_map_mu);
    auto itf = g_maps.find(key);
    if (itf != g_maps.end()) { *out = itf->second; return ST_OK; }
    EncodeTiledFn enc = get_encode();
    if (!enc) { st_set_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)"); return ST_ERR_CUDA; }
    cuuint64_t gd[4], gs[3]; cuuint32_t bx[4], es[4];
    uint64_t stride = 2;
    for (int i = 0; i < rank; ++i) {
        gd[i] = dims[i]; bx[i] = box[i]; es[i] = (i == 1 || i == 2) ? pixel_stride : 1;
        stride *= dims[i];
        if (i < rank - 1) gs[i] = stride;
    }
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(ptr), gd, gs, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { st_set_error("cuTensorMapEncodeTiled failed: %d", (int)r); return ST_ERR_CUDA; }
    if (g_maps.size() > 8192) g_maps.clear();
    g_maps.emplace(key, *out);
    return ST_OK;
}

namespace {
inline int make_map(CUtensorMap* out, const void* ptr, int rank, const uint64_t* dims, const uint32_t* box) {
    return st_make_tmap(out, ptr, rank, dims, box, 128, 1);
}

template <int KS, int BN, int PL, bool RES, bool FUSED>
int launch_one(const CUtensorMap& ah, const CUtensorMap& al, const CUtensorMap& wh, const CUtensorMap& wl,
               const ConvTcParams& p, cudaStream_t stream) {
    using C = Cfg<KS, BN, PL, RES>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel<KS, BN, PL, RES, FUSED>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
        if (e != cudaSuccess) { st_set_error("conv_tc smem attribute: %s", cudaGetErrorString(e)); return ST_ERR_CUDA; }
        attr_set = true;
    }
    const int grid = p.work < st_num_sms() ? p.work : st_num_sms();
    st_launch(grid, NUM_THREADS, C::SMEM, stream)(conv_tc_kernel<KS, BN, PL, RES, FUSED>, ah, al, wh, wl, p);
    ST_CHECK_LAUNCH("conv_tc_kernel");
    return ST_OK;
}
// the fused epilogue (ReLU mask / addend) and the plain one live in separate kernels
template <int KS, int BN, int PL, bool RES>
int launch(const CUtensorMap& ah, const CUtensorMap& al, const CUtensorMap& wh, const CUtensorMap& wl,
           const ConvTcParams& p, cudaStream_t stream) {
    if (p.mask || p.add_hi) return launch_one<KS, BN, PL, RES, true>(ah, al, wh, wl, p, stream);
    return launch_one<KS, BN, PL, RES, false>(ah, al, wh, wl, p, stream);
}

template <int KS, int BN>
int launch_planes(int planes, const CUtensorMap& ah, const CUtensorMap& al, const CUtensorMap& wh, const CUtensorMap& wl,
                  const ConvTcParams& p, cudaStream_t stream) {
    if constexpr (KS == 3 && BN <= 64) {
        // one K chunk, one channel block, shared weights: the nine weight tiles stay in shared memory
        if (planes == 2 && p.w_resident) return launch<KS, BN, 2, true>(ah, al, wh, wl, p, stream);
    }
    if constexpr (!(KS == 3 && BN == 128)) {
        if (planes == 3) return launch<KS, BN, 3, false>(ah, al, wh, wl, p, stream);
    }
    if (planes == 3) { st_set_error("conv_tc: 3 planes with BN = 128 (3x3) is dispatched as two BN = 64 blocks"); return ST_ERR_UNSUPPORTED; }
    if (planes == 2) return launch<KS, BN, 2, false>(ah, al, wh, wl, p, stream);
    return launch<KS, BN, 1, false>(ah, al, wh, wl, p, stream);
}

}  // namespace

// 1x1 launches whose operands are read at <= 2 planes store through the per-warp shared-memory tiles (Cfg::TS), which
// hold two output planes: a three-plane output of such a launch (not produced by the product) goes to the generic kernel.
bool conv_tc_outputs_ok(int ks, int Cout, int x_planes, int w_planes, int out_planes) {
    const int p = pc_planes(x_planes) < pc_planes(w_planes) ? pc_planes(x_planes) : pc_planes(w_planes);
    return ks != 1 || Cout < 64 || p >= 3 || pc_planes(out_planes) <= 2;
}
// Eligibility of the tcgen05 path (the generic kernel covers everything else).
bool conv_tc_eligible(int ks, int stride, int Cin, int Cout) {
    return (ks == 3 || ks == 1) && stride == 1 && Cin % 64 == 0 && (Cout % 128 == 0 || Cout == 64 || Cout == 32 || (Cout == 16 && ks == 3));
}

// x_planes / w_planes / out_planes / add_code are plane CODES (common.cuh); acc_mul, acc_mul_slot, add_mul_slot: see ConvTcParams
int conv_tc_run(const bf16* x_hi, const bf16* x_lo, int x_planes, int N, int H, int W, int Cin,
                const bf16* w_hi, const bf16* w_lo, int w_planes, int w_rows_total, int w_img_rows, int ks, int Cout,
                const float* bias, const bf16* mask, const bf16* add_hi, const bf16* add_lo, int add_code,
                bf16* out_hi, bf16* out_lo, bf16* relu_hi, bf16* relu_lo, int out_planes, float* out_nchw, int Cnchw,
                float acc_mul, const float* acc_mul_slot, const float* add_mul_slot, const unsigned char* add_route,
                bf16* pool_hi, bf16* pool_lo, unsigned char* pool_idx, cudaStream_t stream) {
    const int taps = ks * ks;
    ST_CHECK_ARG(!pool_idx || (ks == 3 && pool_hi && !mask && !add_hi && H >= 2 && W >= 2 && static_cast<long long>(N) * H * W < (1ll << 31)),
                 "conv_tc: the fused relu + max-pool output needs an unfused 3x3 launch");
    ST_CHECK_ARG(!add_route || (add_hi && pc_planes(add_code) <= 2), "conv_tc: a routed addend needs the pooled gradient in add_hi (<= 2 planes)");
    ST_CHECK_ARG(!add_route || static_cast<long long>(N) * H * W < (1ll << 31), "conv_tc: a routed addend indexes pixels with 32 bits (N*H*W < 2^31)");
    ST_CHECK_ARG(!out_nchw || (Cnchw >= 1 && Cnchw <= 4 && !mask && !add_hi), "conv_tc: the NCHW output takes <= 4 channels and no mask / addend");
    ST_CHECK_ARG(pc_f16(x_planes) == pc_f16(w_planes), "conv_tc: activation and weight planes must have the same 16-bit format");
    // both operands are read at the same plane count (the lower one if they differ)
    const int xp = pc_planes(x_planes), wp = pc_planes(w_planes);
    const int a_planes = x_lo ? (xp >= 3 ? 3 : (xp >= 2 ? 2 : 1)) : 1;
    const int w_planes_ = w_lo ? (wp >= 3 ? 3 : (wp >= 2 ? 2 : 1)) : 1;
    const int planes = a_planes < w_planes_ ? a_planes : w_planes_;
    // 3x3 with three planes: two A slots of three 23 KB planes leave no room for 128-row weight tiles -> 64-channel blocks
    int BN = (Cout % 128 == 0) ? ((ks == 3 && planes == 3) ? 64 : 128) : Cout;
    const int TH = ks == 3 ? 16 : 8, TW = ks == 3 ? 8 : 16;
    // small problems (two images of 64 x 64 per GPU under data parallelism): when 128-channel tiles leave more than half
    // of the SMs without a work item, 64-channel tiles double the items and halve the weights each CTA has to fetch
    if (BN == 128 && N * ((H + TH - 1) / TH) * ((W + TW - 1) / TW) * (Cout / 128) * 2 <= st_num_sms()) BN = 64;
    ConvTcParams p;
    p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout;
    p.tiles_h = (H + TH - 1) / TH; p.tiles_w = (W + TW - 1) / TW;
    p.n_blocks = Cout / BN; p.work = N * p.tiles_h * p.tiles_w * p.n_blocks;
    p.fd_nb.init(p.n_blocks); p.fd_tw.init(p.tiles_w); p.fd_th.init(p.tiles_h); p.fd_W.init(W); p.fd_H.init(H);
    p.w_img_rows = w_img_rows; p.w_rows_total = w_rows_total;
    p.w_resident = (ks == 3 && BN <= 64 && planes == 2 && Cin == KCH && p.n_blocks == 1 && w_img_rows == 0) ? 1 : 0;
    p.fmt_bits = ptx::umma_fmt_bits(pc_f16(x_planes));
    p.out_planes = out_planes; p.add_code = add_code;
    p.acc_mul = acc_mul; p.acc_mul_slot = acc_mul_slot; p.add_mul_slot = add_mul_slot;
    p.bias = bias; p.mask = mask; p.add_hi = add_hi; p.add_lo = add_lo;
    p.add_route = add_route; p.route_Ho = H / 2; p.route_Wo = W / 2;
    p.pool_hi = pool_hi; p.pool_lo = pool_lo; p.pool_idx = pool_idx; p.pool_lane_h = TW; p.pool_Ho = H / 2; p.pool_Wo = W / 2;
    p.out_hi = out_hi; p.out_lo = out_lo; p.relu_hi = relu_hi; p.relu_lo = relu_lo;
    p.out_nchw = out_nchw; p.nchw_c = Cnchw; p.out_f32 = nullptr;
    p.dbg = g_conv_debug;

    // the third plane follows the second at the same stride: the "lo" map simply spans both
    CUtensorMap ah, al, wh, wl;
    uint64_t ad[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t ad2[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)N * (planes == 3 ? 2 : 1)};
    uint32_t ab[4] = {KCH, (uint32_t)(TW + ks - 1), (uint32_t)(TH + ks - 1), 1};
    uint64_t wd[3] = {(uint64_t)Cin, (uint64_t)taps, (uint64_t)w_rows_total};
    uint64_t wd2[3] = {(uint64_t)Cin, (uint64_t)taps, (uint64_t)w_rows_total * (planes == 3 ? 2 : 1)};
    uint32_t wb[3] = {KCH, 1, (uint32_t)BN};
    int rc;
    if ((rc = make_map(&ah, x_hi, 4, ad, ab))) return rc;
    if ((rc = make_map(&wh, w_hi, 3, wd, wb))) return rc;
    al = ah; wl = wh;
    if (planes > 1 && (rc = make_map(&al, x_lo, 4, ad2, ab))) return rc;
    if (planes > 1 && (rc = make_map(&wl, w_lo, 3, wd2, wb))) return rc;

    if (ks == 3) {
        if (BN == 128) return launch_planes<3, 128>(planes, ah, al, wh, wl, p, stream);
        if (BN == 64) return launch_planes<3, 64>(planes, ah, al, wh, wl, p, stream);
        if (BN == 32) return launch_planes<3, 32>(planes, ah, al, wh, wl, p, stream);
        if (BN == 16) return launch_planes<3, 16>(planes, ah, al, wh, wl, p, stream);
    } else {
        if (BN == 128) return launch_planes<1, 128>(planes, ah, al, wh, wl, p, stream);
        if (BN == 64) return launch_planes<1, 64>(planes, ah, al, wh, wl, p, stream);
        if (BN == 32) return launch_planes<1, 32>(planes, ah, al, wh, wl, p, stream);
    }
    st_set_error("conv_tc: unsupported tile configuration ks=%d BN=%d", ks, BN);
    return ST_ERR_UNSUPPORTED;
}
