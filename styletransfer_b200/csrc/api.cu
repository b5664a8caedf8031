// extern "C" entry points of libstransfer_b200.so (see include/stransfer_b200.h).
#include "common.cuh"
#include "../../include/stransfer_b200.h"
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

static thread_local char g_err[512] = "";
void st_set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
static int g_conv_path = 0;
// programmatic dependent launch for every kernel of the library (common.cuh): STRANSFER_PDL=1 turns it on.  Measured
// neutral on B200 (Gatys 1294 vs 1314 it/s, fast_st at 2 images per GPU 2.210 vs 2.221 ms per step; all GPU tests pass
// either way): the graph-replayed steps are bound by the kernels' own fixed costs, not by the gaps between them.
int g_st_pdl = []() { const char* e = getenv("STRANSFER_PDL"); return (e && e[0] == '1') ? 1 : 0; }();
long long g_st_launches = 0;

// internal kernels (other translation units)
bool conv_tc_eligible(int ks, int stride, int Cin, int Cout);
bool conv_imgrad_tc_eligible(int ks, int Cin, int Cout_pad, int x_planes, int w_planes, int w_rows_total, int Cnchw);
int conv_imgrad_tc_run(const bf16*, const bf16*, int, int, int, int, const bf16*, const bf16*, int, float*, int, float, const float*, cudaStream_t);
bool conv_tc_outputs_ok(int ks, int Cout, int x_planes, int w_planes, int out_planes);
int conv_tc_run(const bf16*, const bf16*, int, int, int, int, int, const bf16*, const bf16*, int, int, int, int, int,
                const float*, const bf16*, const bf16*, const bf16*, int, bf16*, bf16*, bf16*, bf16*, int, float*, int,
                float, const float*, const float*, const unsigned char*, bf16*, bf16*, unsigned char*, cudaStream_t);
int conv_gen_run(const bf16*, const bf16*, int, int, int, int, int, int, const bf16*, const bf16*, int, int, int, int, int,
                 int, int, int, int, int, const float*, const bf16*, const bf16*, const bf16*, int,
                 bf16*, bf16*, bf16*, bf16*, int, float*, int, float, const float*, const float*, cudaStream_t);
int conv_image_fwd(const float*, const float*, const float*, bf16*, bf16*, bf16*, bf16*, int, int, int, int, int, cudaStream_t);
int conv_image_dgrad(const bf16*, const bf16*, int, const float*, float*, int, int, int, int, cudaStream_t);
int conv9_expand_run(const float*, const float*, int, int, const float*, const float*, int, bf16*, bf16*, int, int, int, int, int, cudaStream_t);
int conv9_reduce_run(const bf16*, const bf16*, int, const float*, const float*, float*, int, int, int, cudaStream_t);
size_t conv9_wgrad_workspace_bytes(int CS);
size_t conv9_reduce_workspace_bytes(int N, int H, int W);
int conv9_reduce_tc_run(const bf16*, const bf16*, int, const float*, const float*, float*, void*, size_t, int, int, int, cudaStream_t);
int conv9_wgrad_run(const bf16*, const bf16*, int, const float*, const float*, const float*, const float*, int, int, int, float*, float*, float*, size_t, int, int, int, int, cudaStream_t);
bool conv_tcx_eligible(int mode, int ks, int Cin, int Cout, int H, int W, int Ho, int Wo);
int conv_tcx_run(int mode, const bf16*, const bf16*, int, int, int, int, int, const bf16*, const bf16*, int, int, int, int, int,
                 const float*, const bf16*, const bf16*, const bf16*, int, bf16*, bf16*, bf16*, bf16*, int, float*,
                 float, const float*, const float*, cudaStream_t);
bool gram_tc_eligible(int C, bool two_planes);
size_t gram_tc_workspace_bytes(int B, int HW, int C);
int gram_tc_finish_blocks(int B, int HW, int C);
int gram_tc_run(const bf16*, bool, int, int, int, float, float*, void*, size_t, const float*, float*, float*, float*, float*, int*, cudaStream_t);
bool wgrad_tc_eligible(int ks, int stride, int pad, int up, int CinPad, int Cdy, bool two_planes);
size_t wgrad_tc_workspace_bytes(int N, int H, int W, int CinPad, int Cdy, int ks);
int wgrad_tc_run(const bf16*, const bf16*, bool, const float*, int, int, int, int, int, int, int, int, float*, void*, size_t, cudaStream_t);
size_t contract_workspace_bytes(int N, int Ho, int Wo, int Cdy, int Ktot, int groups);
int contract_run(const bf16*, const bf16*, int, bool, const float*, const bf16*, const bf16*, bool, int, int, int, int, int, int, int, int, int, int,
                 int, int, int, float, float*, void*, size_t, cudaStream_t);
int ew_nchw_to_nhwc(const float*, const float*, bf16*, bf16*, int, int, int, int, int, int, int, cudaStream_t);
int ew_nhwc_to_nchw(const bf16*, const bf16*, int, const float*, float*, int, int, int, int, int, cudaStream_t);
int ew_pack_weight(const float*, bf16*, bf16*, int, int, int, int, int, int, int, cudaStream_t);
int ew_relu_maxpool_fwd(const bf16*, const bf16*, bf16*, bf16*, int, int, int, int, int, cudaStream_t);
int ew_relu_maxpool_bwd(const bf16*, const bf16*, const bf16*, const bf16*, const bf16*, const bf16*, bf16*, bf16*, int, int, int, int, int, cudaStream_t);
int ew_relu_maxpool_fwd_idx(const bf16*, const bf16*, bf16*, bf16*, unsigned char*, int, int, int, int, int, cudaStream_t);
int ew_relu_maxpool_bwd_idx(const unsigned char*, const bf16*, const bf16*, const bf16*, const bf16*, bf16*, bf16*, int, int, int, int, cudaStream_t);
int ew_mse_fwd(const bf16*, const bf16*, int, const bf16*, const bf16*, int, size_t, int, int, float*, float*, cudaStream_t);
int ew_mse_bwd(const bf16*, const bf16*, int, const bf16*, const bf16*, int, size_t, int, float, const float*, const float*, int, bf16*, bf16*, cudaStream_t);
int ew_style_fwd(const float*, const float*, int, int, float*, float*, float*, float*, cudaStream_t);
int ew_style_seed(const float*, int, int, float, const float*, bf16*, bf16*, int, const float*, float*, cudaStream_t);
int ew_style_seed_multi(const StyleSeedTable&, int, cudaStream_t);
int ew_absmax_slot(const float*, size_t, float, float*, float*, cudaStream_t);
int ew_loss_total(const float*, const float*, int, const float*, float, const float*, float, float*, cudaStream_t);
int ew_image_u8_to_nchw(const unsigned char*, float*, int, int, int, const float*, const float*, cudaStream_t);
int ew_image_nchw_to_u8(const float*, unsigned char*, int, int, int, const float*, const float*, cudaStream_t);
int ew_channel_sum(const bf16*, const bf16*, int, const float*, int, int, int, float*, float*, cudaStream_t);
size_t in_workspace_floats(int N, int C);
int in_forward(const bf16*, const bf16*, const float*, const float*, float, bf16*, bf16*, int, float*, float*, int, int, int, int, int, int, float*, cudaStream_t);
int in_backward(const bf16*, const bf16*, const bf16*, const bf16*, const float*, const float*, const float*, const float*,
                bf16*, bf16*, int, float*, float*, float*, float*, int, int, int, int, int, int, int, float*, cudaStream_t);
int tv_forward(const float*, size_t, int, int, float, float*, float*, cudaStream_t);
int tv_backward(const float*, size_t, int, int, float, const float*, float*, cudaStream_t);
int temporal_forward(const float*, const float*, const float*, const float*, size_t, float, float*, float*, cudaStream_t);
int temporal_backward(const float*, const float*, size_t, float, const float*, const float*, float*, cudaStream_t);
int temporal_warp_forward(const float*, const float*, const float*, const float*, const float*, int, int, int, int, float, float*, float*, cudaStream_t);
int temporal_warp_backward(const float*, const float*, const float*, int, int, int, int, float, const float*, const float*, float*, cudaStream_t);

// plane codes (common.cuh): 1..3 bf16 planes, or 1..2 fp16 planes
static bool plane_code_ok(int code) { const int n = pc_planes(code); return n >= 1 && n <= (pc_f16(code) ? 2 : 3) && (code & ~31) == 0; }
static int one_plane_if_no_lo(int code, const void* lo) { return lo ? code : ((code & ST_FMT_F16) | 1); }

#define BF(p) static_cast<bf16*>(p)
#define CBF(p) static_cast<const bf16*>(p)
#define STRM(s) static_cast<cudaStream_t>(s)

extern "C" {

int st_arch(void) { return 100; }
const char* st_last_error(void) { return g_err; }
int g_conv_debug = 0;        // experiment switches of conv_tc (bits 8.. of st_set_conv_path): see ConvTcParams::dbg
int st_set_conv_path(int path) { g_conv_path = path & 0xff; g_conv_debug = path >> 8; return ST_OK; }
static int g_last_conv_kernel = 0;
int st_last_conv_kernel(void) { return g_last_conv_kernel; }
long long st_launch_count(void) { return g_st_launches; }

int st_nchw_to_nhwc(const float* a, const float* b, void* hi, void* lo, int planes, int N, int C1, int C2, int H, int W, int Cpad, void* stream) {
    ST_CHECK_ARG(a && hi && N > 0 && H > 0 && W > 0, "st_nchw_to_nhwc: null pointer or empty shape");
    ST_CHECK_ARG(Cpad % 8 == 0 && C1 + C2 <= Cpad && (C2 == 0 || b), "st_nchw_to_nhwc: Cpad must be a multiple of 8 and >= C1+C2");
    ST_CHECK_ARG(plane_code_ok(planes), "st_nchw_to_nhwc: bad plane code %d", planes);
    return ew_nchw_to_nhwc(a, b, BF(hi), BF(lo), one_plane_if_no_lo(planes, lo), N, C1, C2, H, W, Cpad, STRM(stream));
}
int st_nhwc_to_nchw(const void* hi, const void* lo, int planes, const float* slot, float* out, int N, int C, int H, int W, int Cpad, void* stream) {
    ST_CHECK_ARG(hi && out && N > 0 && H > 0 && W > 0, "st_nhwc_to_nchw: null pointer or empty shape");
    ST_CHECK_ARG(Cpad % 8 == 0 && C <= Cpad, "st_nhwc_to_nchw: Cpad must be a multiple of 8 and >= C");
    return ew_nhwc_to_nchw(CBF(hi), CBF(lo), one_plane_if_no_lo(planes, lo), slot, out, N, C, H, W, Cpad, STRM(stream));
}
int st_image_u8_to_nchw(const unsigned char* in, float* out, int N, int H, int W, const float* mean3, const float* std3, void* stream) {
    ST_CHECK_ARG(in && out && mean3 && std3 && N > 0 && H > 0 && W > 0, "st_image_u8_to_nchw: null pointer or empty shape");
    return ew_image_u8_to_nchw(in, out, N, H, W, mean3, std3, STRM(stream));
}
int st_image_nchw_to_u8(const float* in, unsigned char* out, int N, int H, int W, const float* mean3, const float* std3, void* stream) {
    ST_CHECK_ARG(in && out && mean3 && std3 && N > 0 && H > 0 && W > 0, "st_image_nchw_to_u8: null pointer or empty shape");
    return ew_image_nchw_to_u8(in, out, N, H, W, mean3, std3, STRM(stream));
}
int st_pack_conv_weight(const float* w, void* hi, void* lo, int planes, int Cout, int Cin, int ks, int rows_pad, int cols_pad, int mode, void* stream) {
    ST_CHECK_ARG(w && hi, "st_pack_conv_weight: null pointer");
    ST_CHECK_ARG(mode >= 0 && mode <= 2 && cols_pad % 8 == 0, "st_pack_conv_weight: bad mode or cols_pad %% 8 != 0");
    const int rows = mode == 0 ? Cout : Cin, cols = mode == 0 ? Cin : Cout;
    ST_CHECK_ARG(rows_pad >= rows && cols_pad >= cols, "st_pack_conv_weight: padded sizes smaller than the weight");
    ST_CHECK_ARG(plane_code_ok(planes), "st_pack_conv_weight: bad plane code %d", planes);
    return ew_pack_weight(w, BF(hi), BF(lo), one_plane_if_no_lo(planes, lo), Cout, Cin, ks, rows_pad, cols_pad, mode, STRM(stream));
}

static int conv2d_impl(const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int Cin, int up,
              const void* w_hi, const void* w_lo, int w_planes, int w_rows_total, int w_img_rows,
              int ks, int stride, int pad, int transposed, int Ho, int Wo, int Cout_pad, int Cbias,
              const float* bias, const void* mask, const void* add_hi, const void* add_lo, int add_planes,
              void* out_hi, void* out_lo, void* relu_hi, void* relu_lo, int out_planes, float* out_nchw, int Cnchw,
              float acc_mul, const float* acc_mul_slot, const float* add_mul_slot, const unsigned char* add_route,
              void* pool_hi, void* pool_lo, unsigned char* pool_idx, void* stream) {
    ST_CHECK_ARG(x_hi && w_hi, "st_conv2d: null input or weight");
    ST_CHECK_ARG(N > 0 && H > 0 && W > 0 && Ho > 0 && Wo > 0, "st_conv2d: empty shape");
    ST_CHECK_ARG(Cin % 8 == 0 && Cout_pad % 8 == 0, "st_conv2d: channel counts must be multiples of 8 (got %d, %d)", Cin, Cout_pad);
    ST_CHECK_ARG(ks >= 1 && ks <= 11 && (ks & 1) && stride >= 1, "st_conv2d: unsupported kernel size/stride");
    ST_CHECK_ARG(out_hi || relu_hi || out_nchw || pool_idx, "st_conv2d: no output requested");
    ST_CHECK_ARG(!(add_lo && !add_hi), "st_conv2d: add_lo without add_hi");
    ST_CHECK_ARG(plane_code_ok(x_planes) && plane_code_ok(w_planes) && plane_code_ok(out_planes) && (!add_hi || plane_code_ok(add_planes)),
                 "st_conv2d: bad plane code (x %d, w %d, out %d, add %d)", x_planes, w_planes, out_planes, add_planes);
    x_planes = one_plane_if_no_lo(x_planes, x_lo); w_planes = one_plane_if_no_lo(w_planes, w_lo);
    add_planes = add_hi ? one_plane_if_no_lo(add_planes, add_lo) : 1;
    // the tensor cores multiply two operands of ONE 16-bit format; mixed formats go to the converting mma.sync kernel
    const bool same_fmt = pc_f16(x_planes) == pc_f16(w_planes);
    // an fp32 NCHW output (the image gradient) is taken by the 16-channel tile, unfused launches only
    const bool nchw_ok = !out_nchw || (Cout_pad == 16 && Cnchw <= 4 && !mask && !add_hi);
    const bool tc = (g_conv_path == 0 || g_conv_path == 2) && same_fmt && !up && !transposed && pad == ks / 2 && Ho == H && Wo == W && nchw_ok &&
                    w_rows_total >= Cout_pad && conv_tc_eligible(ks, stride, Cin, Cout_pad) && (Cout_pad != 16 || out_nchw) &&
                    conv_tc_outputs_ok(ks, Cout_pad, x_planes, w_planes, out_planes);
    g_last_conv_kernel = tc ? 1 : 3;
    // backward-data of conv1_1 to the image: "project, then gather" kernel (conv_imgrad_tc.cu) instead of the 16-channel tile
    if (tc && g_conv_path == 0 && out_nchw && !out_hi && !relu_hi && !bias && w_img_rows == 0 && x_lo && w_lo &&
        conv_imgrad_tc_eligible(ks, Cin, Cout_pad, x_planes, w_planes, w_rows_total, Cnchw)) {
        g_last_conv_kernel = 4;
        return conv_imgrad_tc_run(CBF(x_hi), CBF(x_lo), x_planes, N, H, W, CBF(w_hi), CBF(w_lo), w_rows_total, out_nchw, Cnchw,
                                  acc_mul, acc_mul_slot, STRM(stream));
    }
    if (tc)
        return conv_tc_run(CBF(x_hi), CBF(x_lo), x_planes, N, H, W, Cin, CBF(w_hi), CBF(w_lo), w_planes, w_rows_total, w_img_rows, ks, Cout_pad,
                           bias, CBF(mask), CBF(add_hi), CBF(add_lo), add_planes, BF(out_hi), BF(out_lo), BF(relu_hi), BF(relu_lo), out_planes, out_nchw, Cnchw,
                           acc_mul, acc_mul_slot, add_mul_slot, add_route, BF(pool_hi), BF(pool_lo), pool_idx, STRM(stream));
    if (pool_idx) {
        st_set_error("st_conv2d_pool: the fused relu + max-pool output needs the stride-1 3x3 tensor-core path (ks %d stride %d Cin %d Cout %d)",
                     ks, stride, Cin, Cout_pad);
        return ST_ERR_UNSUPPORTED;
    }
    if (add_route) {
        st_set_error("st_conv2d_routed: a routed addend needs the stride-1 tensor-core path (ks %d stride %d Cin %d Cout %d, same 16-bit format %d)",
                     ks, stride, Cin, Cout_pad, int(same_fmt));
        return ST_ERR_UNSUPPORTED;
    }
    // run-time-geometry tcgen05 kernel: stride-2 3x3 (forward / transposed) and 32-channel K chunks
    const int xmode = transposed ? 2 : (stride == 2 ? 1 : 0);
    const bool tcx = g_conv_path == 0 && same_fmt && !up && pad == 1 && stride <= 2 && (!transposed || stride == 2) && !out_nchw && w_img_rows == 0 &&
                     (!bias || Cbias == Cout_pad) && w_rows_total >= Cout_pad && conv_tcx_eligible(xmode, ks, Cin, Cout_pad, H, W, Ho, Wo);
    if (tcx) g_last_conv_kernel = 2;
    if (tcx)
        return conv_tcx_run(xmode, CBF(x_hi), CBF(x_lo), x_planes, N, H, W, Cin, CBF(w_hi), CBF(w_lo), w_planes, w_rows_total, Cout_pad, Ho, Wo,
                            bias, CBF(mask), CBF(add_hi), CBF(add_lo), add_planes, BF(out_hi), BF(out_lo), BF(relu_hi), BF(relu_lo), out_planes, nullptr,
                            acc_mul, acc_mul_slot, add_mul_slot, STRM(stream));
    // the mma.sync kernel multiplies bf16 planes: at most 2 source planes of a bf16 operand of a backward launch, as before
    auto gen_code = [](int code) { return pc_f16(code) ? code : (pc_planes(code) >= 3 ? 3 : pc_planes(code)); };
    if (w_img_rows != 0) {
        // per-image weights on the generic path: one launch per image
        for (int n = 0; n < N; ++n) {
            const size_t xin = static_cast<size_t>(n) * H * W * Cin, xo = static_cast<size_t>(n) * Ho * Wo * Cout_pad;
            const size_t wo = static_cast<size_t>(n) * w_img_rows * ks * ks * Cin;
            const int xc = pc_f16(x_planes) ? x_planes : (pc_planes(x_planes) >= 3 ? 2 : x_planes);
            const int wc = pc_f16(w_planes) ? w_planes : (pc_planes(w_planes) >= 3 ? 2 : w_planes);
            const int oc = pc_f16(out_planes) ? out_planes : (pc_planes(out_planes) >= 3 ? 2 : out_planes);
            int rc = conv_gen_run(CBF(x_hi) + xin, x_lo ? CBF(x_lo) + xin : nullptr, xc, 1, H, W, Cin, up,
                                  CBF(w_hi) + wo, w_lo ? CBF(w_lo) + wo : nullptr, wc, ks, stride, pad, transposed, Ho, Wo, Cout_pad, Cout_pad, Cbias,
                                  bias, mask ? CBF(mask) + xo : nullptr, add_hi ? CBF(add_hi) + xo : nullptr, add_lo ? CBF(add_lo) + xo : nullptr, add_planes,
                                  out_hi ? BF(out_hi) + xo : nullptr, out_lo ? BF(out_lo) + xo : nullptr,
                                  relu_hi ? BF(relu_hi) + xo : nullptr, relu_lo ? BF(relu_lo) + xo : nullptr, oc,
                                  out_nchw ? out_nchw + static_cast<size_t>(n) * Cnchw * Ho * Wo : nullptr, Cnchw,
                                  acc_mul, acc_mul_slot, add_mul_slot, STRM(stream));
            if (rc) return rc;
        }
        return ST_OK;
    }
    return conv_gen_run(CBF(x_hi), CBF(x_lo), gen_code(x_planes), N, H, W, Cin, up, CBF(w_hi), CBF(w_lo), gen_code(w_planes), ks, stride, pad, transposed,
                        Ho, Wo, Cout_pad, Cout_pad, Cbias, bias, CBF(mask), CBF(add_hi), CBF(add_lo), add_planes,
                        BF(out_hi), BF(out_lo), BF(relu_hi), BF(relu_lo), out_planes, out_nchw, Cnchw,
                        acc_mul, acc_mul_slot, add_mul_slot, STRM(stream));
}

int st_conv2d(const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int Cin, int up,
              const void* w_hi, const void* w_lo, int w_planes, int w_rows_total, int w_img_rows,
              int ks, int stride, int pad, int transposed, int Ho, int Wo, int Cout_pad, int Cbias,
              const float* bias, const void* mask, const void* add_hi, const void* add_lo, int add_planes,
              void* out_hi, void* out_lo, void* relu_hi, void* relu_lo, int out_planes, float* out_nchw, int Cnchw,
              float acc_mul, const float* acc_mul_slot, const float* add_mul_slot, void* stream) {
    return conv2d_impl(x_hi, x_lo, x_planes, N, H, W, Cin, up, w_hi, w_lo, w_planes, w_rows_total, w_img_rows, ks, stride, pad, transposed,
                       Ho, Wo, Cout_pad, Cbias, bias, mask, add_hi, add_lo, add_planes, out_hi, out_lo, relu_hi, relu_lo, out_planes,
                       out_nchw, Cnchw, acc_mul, acc_mul_slot, add_mul_slot, nullptr, nullptr, nullptr, nullptr, stream);
}
int st_conv2d_routed(const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int Cin,
                     const void* w_hi, const void* w_lo, int w_planes, int w_rows_total, int w_img_rows, int ks, int Cout_pad,
                     const void* mask, const void* gy_hi, const void* gy_lo, int gy_planes, const unsigned char* route,
                     void* out_hi, void* out_lo, int out_planes, float acc_mul, const float* acc_mul_slot, const float* gy_mul_slot,
                     void* stream) {
    ST_CHECK_ARG(gy_hi && route && out_hi, "st_conv2d_routed: null pooled gradient, routing map or output");
    ST_CHECK_ARG(H >= 2 && W >= 2, "st_conv2d_routed: the 2x2 pool needs H, W >= 2 (got %d x %d)", H, W);
    return conv2d_impl(x_hi, x_lo, x_planes, N, H, W, Cin, 0, w_hi, w_lo, w_planes, w_rows_total, w_img_rows, ks, 1, ks / 2, 0,
                       H, W, Cout_pad, 0, nullptr, mask, gy_hi, gy_lo, gy_planes, out_hi, out_lo, nullptr, nullptr, out_planes,
                       nullptr, 0, acc_mul, acc_mul_slot, gy_mul_slot, route, nullptr, nullptr, nullptr, stream);
}

int st_conv2d_pool(const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int Cin,
                   const void* w_hi, const void* w_lo, int w_planes, int w_rows_total, int ks, int Cout_pad, int Cbias, const float* bias,
                   void* out_hi, void* out_lo, int out_planes, void* pool_hi, void* pool_lo, unsigned char* pool_idx,
                   float acc_mul, const float* acc_mul_slot, void* stream) {
    ST_CHECK_ARG(pool_hi && pool_idx, "st_conv2d_pool: null pooled output or routing map");
    ST_CHECK_ARG(pc_planes(out_planes) < 2 || pool_lo, "st_conv2d_pool: the pooled output needs its lo plane");
    ST_CHECK_ARG(H >= 2 && W >= 2, "st_conv2d_pool: the 2x2 pool needs H, W >= 2 (got %d x %d)", H, W);
    return conv2d_impl(x_hi, x_lo, x_planes, N, H, W, Cin, 0, w_hi, w_lo, w_planes, w_rows_total, 0, ks, 1, ks / 2, 0,
                       H, W, Cout_pad, Cbias, bias, nullptr, nullptr, nullptr, 1, out_hi, out_lo, nullptr, nullptr, out_planes,
                       nullptr, 0, acc_mul, acc_mul_slot, nullptr, nullptr, pool_hi, pool_lo, pool_idx, stream);
}

int st_conv3x3_image_fwd(const float* x, const float* w, const float* bias, void* out_hi, void* out_lo,
                         void* relu_hi, void* relu_lo, int planes, int N, int H, int W, int Cout, void* stream) {
    ST_CHECK_ARG(x && w && (out_hi || relu_hi), "st_conv3x3_image_fwd: null pointer");
    ST_CHECK_ARG(N > 0 && H > 0 && W > 0 && Cout % 8 == 0 && Cout > 0 && Cout <= 256, "st_conv3x3_image_fwd: bad shape (Cout %% 8, <= 256)");
    ST_CHECK_ARG(plane_code_ok(planes) && (pc_planes(planes) == 1 || ((!out_hi || out_lo) && (!relu_hi || relu_lo))), "st_conv3x3_image_fwd: planes/lo mismatch");
    return conv_image_fwd(x, w, bias, BF(out_hi), BF(out_lo), BF(relu_hi), BF(relu_lo), planes, N, H, W, Cout, STRM(stream));
}
int st_conv3x3_image_dgrad(const void* g_hi, const void* g_lo, int g_planes, const float* w, float* gx,
                           int N, int H, int W, int Cout, void* stream) {
    ST_CHECK_ARG(g_hi && w && gx, "st_conv3x3_image_dgrad: null pointer");
    ST_CHECK_ARG(N > 0 && H > 0 && W > 0 && Cout % 8 == 0 && Cout > 0 && Cout <= 256, "st_conv3x3_image_dgrad: bad shape (Cout %% 8, <= 256)");
    return conv_image_dgrad(CBF(g_hi), CBF(g_lo), one_plane_if_no_lo(g_planes, g_lo), w, gx, N, H, W, Cout, STRM(stream));
}

int st_conv9_expand(const float* s0, const float* s1, int sc, int CS, const float* w, const float* bias, int flip,
                    void* out_hi, void* out_lo, int planes, int N, int H, int W, void* stream) {
    ST_CHECK_ARG(s0 && w && out_hi && (CS == 3 || (CS == 6 && s1)), "st_conv9_expand: null pointer or CS not 3/6");
    ST_CHECK_ARG(N > 0 && H > 0 && W > 0 && sc >= 3 && plane_code_ok(planes) && (pc_planes(planes) == 1 || out_lo), "st_conv9_expand: bad shape/planes");
    return conv9_expand_run(s0, s1, sc, CS, w, bias, flip, BF(out_hi), BF(out_lo), planes, N, H, W, g_conv_path == 0, STRM(stream));
}
size_t st_conv9_reduce_workspace_bytes(int N, int H, int W) { return (N > 0 && H > 0 && W > 0) ? conv9_reduce_workspace_bytes(N, H, W) : 0; }
int st_conv9_reduce(const void* x_hi, const void* x_lo, int planes, const float* w, const float* bias, float* out,
                    void* workspace, size_t workspace_bytes, int N, int H, int W, void* stream) {
    ST_CHECK_ARG(x_hi && w && out && N > 0 && H > 0 && W > 0, "st_conv9_reduce: null pointer or empty shape");
    // tensor-core path: planes one plane stride apart (the TMA map spans them), workspace given
    const bool planes_ok = !x_lo || CBF(x_lo) == CBF(x_hi) + static_cast<size_t>(N) * H * W * 32;
    if (workspace && g_conv_path == 0 && planes_ok)
        return conv9_reduce_tc_run(CBF(x_hi), CBF(x_lo), one_plane_if_no_lo(planes, x_lo), w, bias, out, workspace, workspace_bytes, N, H, W, STRM(stream));
    return conv9_reduce_run(CBF(x_hi), CBF(x_lo), one_plane_if_no_lo(planes, x_lo), w, bias, out, N, H, W, STRM(stream));
}
size_t st_conv9_wgrad_workspace_bytes(int CS) { return (CS == 3 || CS == 6) ? conv9_wgrad_workspace_bytes(CS) : 0; }
int st_conv9_wgrad(const void* b_hi, const void* b_lo, int b_planes, const float* big_slot, const float* small_slot,
                   const float* s0, const float* s1, int sc, int CS, int mode,
                   float* dw, float* db, void* workspace, size_t workspace_bytes, int N, int H, int W, void* stream) {
    ST_CHECK_ARG(b_hi && s0 && dw && workspace && (CS == 3 || (CS == 6 && s1)) && (mode == 0 || mode == 1), "st_conv9_wgrad: bad arguments");
    ST_CHECK_ARG(mode == 0 || CS == 3, "st_conv9_wgrad: layer 22 has 3 output channels");
    ST_CHECK_ARG(N > 0 && H > 0 && W > 0 && sc >= 3, "st_conv9_wgrad: empty shape");
    ST_CHECK_ARG(plane_code_ok(b_planes), "st_conv9_wgrad: bad plane code %d", b_planes);
    return conv9_wgrad_run(CBF(b_hi), CBF(b_lo), one_plane_if_no_lo(b_planes, b_lo), big_slot, small_slot, s0, s1, sc, CS, mode, dw, db,
                           static_cast<float*>(workspace), workspace_bytes, N, H, W, g_conv_path == 0, STRM(stream));
}

size_t st_conv2d_wgrad_workspace_bytes(int N, int Ho, int Wo, int Cdy, int CinPad, int ks) {
    size_t a = contract_workspace_bytes(N, Ho, Wo, Cdy, ks * ks * CinPad, 1);
    size_t b = static_cast<size_t>(1024) * Cdy * sizeof(float);
    size_t c = (ks == 3 && CinPad % 64 == 0) ? wgrad_tc_workspace_bytes(N, Ho, Wo, CinPad, Cdy, ks) : 0;
    if (b > a) a = b;
    return a > c ? a : c;
}
int st_conv2d_wgrad(const void* dy_hi, const void* dy_lo, int dy_planes, const float* dy_slot, int Cdy,
                    const void* x_hi, const void* x_lo, int x_planes,
                    int N, int H, int W, int CinPad, int up, int Ho, int Wo, int ks, int stride, int pad,
                    int Cout, int Cin, float* dw, float* db, void* workspace, size_t workspace_bytes, void* stream) {
    ST_CHECK_ARG(dy_hi && x_hi && dw && workspace, "st_conv2d_wgrad: null pointer");
    ST_CHECK_ARG(Cdy % 8 == 0 && CinPad % 8 == 0 && Cout <= Cdy && Cin <= CinPad, "st_conv2d_wgrad: bad channel counts");
    ST_CHECK_ARG(plane_code_ok(dy_planes) && plane_code_ok(x_planes), "st_conv2d_wgrad: bad plane code");
    ST_CHECK_ARG(workspace_bytes >= st_conv2d_wgrad_workspace_bytes(N, Ho, Wo, Cdy, CinPad, ks), "st_conv2d_wgrad: workspace too small");
    int rc;
    // tcgen05 path: stride-1 3x3, 64-channel multiples, both tensors with their planes one plane stride apart and in
    // ONE 16-bit format (the tensor cores cannot mix fp16 and bf16 operands)
    const size_t x_plane = static_cast<size_t>(N) * H * W * CinPad, d_plane = static_cast<size_t>(N) * Ho * Wo * Cdy;
    const bool planes_ok = dy_lo && x_lo && pc_planes(dy_planes) >= 2 && pc_planes(x_planes) >= 2 && pc_f16(dy_planes) == pc_f16(x_planes) &&
                           CBF(x_lo) == CBF(x_hi) + x_plane && CBF(dy_lo) == CBF(dy_hi) + d_plane;
    if (g_conv_path == 0 && Ho == H && Wo == W && wgrad_tc_eligible(ks, stride, pad, up, CinPad, Cdy, planes_ok))
        rc = wgrad_tc_run(CBF(dy_hi), CBF(x_hi), pc_f16(x_planes), dy_slot, N, H, W, CinPad, Cdy, ks, Cout, Cin, dw, workspace, workspace_bytes, STRM(stream));
    else
        rc = contract_run(CBF(dy_hi), pc_planes(dy_planes) >= 2 ? CBF(dy_lo) : nullptr, Cdy, pc_f16(dy_planes), dy_slot,
                          CBF(x_hi), pc_planes(x_planes) >= 2 ? CBF(x_lo) : nullptr, pc_f16(x_planes), N, H, W, CinPad, up, Ho, Wo, ks, stride, pad,
                          0, Cout, Cin, 1.f, dw, workspace, workspace_bytes, STRM(stream));
    if (rc) return rc;
    if (db) {
        ST_CHECK_ARG(Cdy <= 256, "st_conv2d_wgrad: bias gradient supports up to 256 channels");
        rc = ew_channel_sum(CBF(dy_hi), CBF(dy_lo), one_plane_if_no_lo(dy_planes, dy_lo), dy_slot, N * Ho * Wo, Cdy, Cout, db, static_cast<float*>(workspace), STRM(stream));
    }
    return rc;
}

int st_relu_maxpool2x2_fwd(const void* x_hi, const void* x_lo, void* y_hi, void* y_lo, int planes, int N, int H, int W, int C, void* stream) {
    ST_CHECK_ARG(x_hi && y_hi && C % 8 == 0 && N > 0, "st_relu_maxpool2x2_fwd: null pointer or C %% 8 != 0");
    return ew_relu_maxpool_fwd(CBF(x_hi), CBF(x_lo), BF(y_hi), BF(y_lo), one_plane_if_no_lo(planes, x_lo), N, H, W, C, STRM(stream));
}
int st_relu_maxpool2x2_bwd(const void* x_hi, const void* x_lo, const void* gy_hi, const void* gy_lo,
                           const void* add_hi, const void* add_lo, void* gx_hi, void* gx_lo, int x_planes, int N, int H, int W, int C, void* stream) {
    ST_CHECK_ARG(x_hi && gy_hi && gx_hi && C % 8 == 0 && N > 0, "st_relu_maxpool2x2_bwd: null pointer or C %% 8 != 0");
    return ew_relu_maxpool_bwd(CBF(x_hi), CBF(x_lo), CBF(gy_hi), CBF(gy_lo), CBF(add_hi), CBF(add_lo), BF(gx_hi), BF(gx_lo), one_plane_if_no_lo(x_planes, x_lo), N, H, W, C, STRM(stream));
}

int st_relu_maxpool2x2_fwd_idx(const void* x_hi, const void* x_lo, void* y_hi, void* y_lo, unsigned char* arg, int planes,
                               int N, int H, int W, int C, void* stream) {
    ST_CHECK_ARG(x_hi && y_hi && arg && C % 8 == 0 && N > 0 && H >= 2 && W >= 2, "st_relu_maxpool2x2_fwd_idx: null pointer, C %% 8 != 0 or H/W < 2");
    return ew_relu_maxpool_fwd_idx(CBF(x_hi), CBF(x_lo), BF(y_hi), BF(y_lo), arg, one_plane_if_no_lo(planes, x_lo), N, H, W, C, STRM(stream));
}
int st_relu_maxpool2x2_bwd_idx(const unsigned char* arg, const void* gy_hi, const void* gy_lo, const void* add_hi, const void* add_lo,
                               void* gx_hi, void* gx_lo, int N, int H, int W, int C, void* stream) {
    ST_CHECK_ARG(arg && gy_hi && gx_hi && C % 8 == 0 && N > 0, "st_relu_maxpool2x2_bwd_idx: null pointer or C %% 8 != 0");
    return ew_relu_maxpool_bwd_idx(arg, CBF(gy_hi), CBF(gy_lo), CBF(add_hi), CBF(add_lo), BF(gx_hi), BF(gx_lo), N, H, W, C, STRM(stream));
}

size_t st_gram_workspace_bytes(int B, int HW, int C) {
    const size_t a = contract_workspace_bytes(B, HW, 1, C, C, B);
    const size_t b = gram_tc_eligible(C, true) ? gram_tc_workspace_bytes(B, HW, C) : 0;
    return a > b ? a : b;
}
int st_gram_fwd(const void* f_hi, const void* f_lo, int f_planes, int B, int HW, int C, float* G, void* workspace, size_t workspace_bytes, void* stream) {
    ST_CHECK_ARG(f_hi && G && workspace && B > 0 && HW > 0, "st_gram_fwd: null pointer or empty shape");
    ST_CHECK_ARG(C % 8 == 0 && plane_code_ok(f_planes), "st_gram_fwd: C %% 8 != 0 or bad plane code");
    const float scale = 1.0f / (static_cast<float>(C) * static_cast<float>(HW));
    // tcgen05 path: 64 / 128 / 256 channels, planes one plane stride apart
    const bool planes_ok = f_lo && pc_planes(f_planes) >= 2 && CBF(f_lo) == CBF(f_hi) + static_cast<size_t>(B) * HW * C;
    if (g_conv_path == 0 && gram_tc_eligible(C, planes_ok))
        return gram_tc_run(CBF(f_hi), pc_f16(f_planes), B, HW, C, scale, G, workspace, workspace_bytes,
                           nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, STRM(stream));
    const bf16* lo = pc_planes(f_planes) >= 2 ? CBF(f_lo) : nullptr;
    return contract_run(CBF(f_hi), lo, C, pc_f16(f_planes), nullptr, CBF(f_hi), lo, pc_f16(f_planes), B, HW, 1, C, 0, HW, 1, 1, 1, 0,
                        1, C, C, scale, G, workspace, workspace_bytes, STRM(stream));
}
// Gram + StyleLoss.forward in two launches (tcgen05 Gram kernel, then ONE finish launch that also forms diff, the loss
// and max|diff|); shapes the tensor-core Gram kernel does not take run the separate launches.
// workspace layout: [Gram split partials | G (B*C*C floats) | style partials (2 * 2048 floats at least)]
static size_t gram_style_part_floats(int B, int HW, int C) {
    const int blocks = gram_tc_eligible(C, true) ? gram_tc_finish_blocks(B, HW, C) : 0;
    return 2 * static_cast<size_t>(blocks > 2048 ? blocks : 2048);
}
size_t st_gram_style_workspace_bytes(int B, int HW, int C) {
    const size_t gram = (st_gram_workspace_bytes(B, HW, C) + 255) & ~static_cast<size_t>(255);
    return gram + (static_cast<size_t>(B) * C * C + gram_style_part_floats(B, HW, C)) * sizeof(float);
}
int st_gram_style_fwd(const void* f_hi, const void* f_lo, int f_planes, int B, int HW, int C, const float* T,
                      float* diff, float* loss, float* diff_amax, int* ticket, void* workspace, size_t workspace_bytes, void* stream) {
    ST_CHECK_ARG(f_hi && T && diff && loss && ticket && workspace && B > 0 && HW > 0, "st_gram_style_fwd: null pointer or empty shape");
    ST_CHECK_ARG(C % 8 == 0 && plane_code_ok(f_planes), "st_gram_style_fwd: C %% 8 != 0 or bad plane code");
    ST_CHECK_ARG(workspace_bytes >= st_gram_style_workspace_bytes(B, HW, C), "st_gram_style_fwd: workspace too small");
    const size_t gram_bytes = (st_gram_workspace_bytes(B, HW, C) + 255) & ~static_cast<size_t>(255);
    float* G = reinterpret_cast<float*>(static_cast<char*>(workspace) + gram_bytes);
    float* part = G + static_cast<size_t>(B) * C * C;
    const bool planes_ok = f_lo && pc_planes(f_planes) >= 2 && CBF(f_lo) == CBF(f_hi) + static_cast<size_t>(B) * HW * C;
    if (g_conv_path == 0 && gram_tc_eligible(C, planes_ok)) {
        const float scale = 1.0f / (static_cast<float>(C) * static_cast<float>(HW));
        return gram_tc_run(CBF(f_hi), pc_f16(f_planes), B, HW, C, scale, nullptr, workspace, gram_bytes,
                           T, diff, loss, diff_amax, part, ticket, STRM(stream));
    }
    int rc = st_gram_fwd(f_hi, f_lo, f_planes, B, HW, C, G, workspace, gram_bytes, stream);
    if (rc) return rc;
    return ew_style_fwd(G, T, B, C, diff, loss, diff_amax, part, STRM(stream));
}
int st_style_mse_fwd(const float* G, const float* T, int B, int C, float* diff, float* loss, float* diff_amax, float* ws, void* stream) {
    ST_CHECK_ARG(G && T && diff && loss && ws, "st_style_mse_fwd: null pointer");
    return ew_style_fwd(G, T, B, C, diff, loss, diff_amax, ws, STRM(stream));
}
int st_style_mse_bwd_seed(const float* diff, int B, int C, int HW, const float* gw, void* d_hi, void* d_lo, int d_planes,
                          const float* diff_amax, float* d_slot, void* stream) {
    ST_CHECK_ARG(diff && gw && d_hi && plane_code_ok(d_planes), "st_style_mse_bwd_seed: null pointer or bad plane code");
    ST_CHECK_ARG(!pc_f16(d_planes) || (diff_amax && d_slot), "st_style_mse_bwd_seed: fp16 planes need diff_amax and a scale slot");
    // d/dF of mean_{b,i,j}(G-T)^2 with G = F^T F/(C*HW): 4/(B*C^2) * 1/(C*HW) * F (G-T)   
    const double scale = 4.0 / (static_cast<double>(B) * C * C) / (static_cast<double>(C) * HW);
    return ew_style_seed(diff, B, C, static_cast<float>(scale), gw, BF(d_hi), pc_planes(d_planes) >= 2 ? BF(d_lo) : nullptr, d_planes,
                         diff_amax, d_slot, STRM(stream));
}
// the seeds of up to ST_MAX_STYLE_NODES style nodes in ONE launch (they depend on the loss gradient only, not on the chain)
int st_style_mse_bwd_seed_multi(int n, const float* const* diff, const int* B, const int* C, const int* HW, const float* const* gw,
                                void* const* d_hi, void* const* d_lo, const int* d_planes, const float* const* diff_amax,
                                float* const* d_slot, void* stream) {
    ST_CHECK_ARG(n >= 1 && n <= ST_MAX_STYLE_NODES && diff && B && C && HW && gw && d_hi && d_lo && d_planes && diff_amax && d_slot,
                 "st_style_mse_bwd_seed_multi: 1..%d nodes, no null tables", ST_MAX_STYLE_NODES);
    StyleSeedTable t;
    t.n = n;
    int max_total = 0;
    for (int i = 0; i < n; ++i) {
        ST_CHECK_ARG(diff[i] && gw[i] && d_hi[i] && plane_code_ok(d_planes[i]), "st_style_mse_bwd_seed_multi: node %d: null pointer or bad plane code", i);
        ST_CHECK_ARG(!pc_f16(d_planes[i]) || (diff_amax[i] && d_slot[i]), "st_style_mse_bwd_seed_multi: node %d: fp16 planes need diff_amax and a scale slot", i);
        t.e[i].diff = diff[i]; t.e[i].total = B[i] * C[i] * C[i];
        t.e[i].scale = static_cast<float>(4.0 / (static_cast<double>(B[i]) * C[i] * C[i]) / (static_cast<double>(C[i]) * HW[i]));
        t.e[i].gw = gw[i]; t.e[i].dh = BF(d_hi[i]); t.e[i].dl = pc_planes(d_planes[i]) >= 2 ? BF(d_lo[i]) : nullptr;
        t.e[i].code = d_planes[i]; t.e[i].amax = diff_amax[i]; t.e[i].slot = d_slot[i];
        if (t.e[i].total > max_total) max_total = t.e[i].total;
    }
    return ew_style_seed_multi(t, max_total, STRM(stream));
}
int st_loss_total(const float* losses, const float* weights, int n, const float* extra0, float scale0,
                  const float* extra1, float scale1, float* out, void* stream) {
    ST_CHECK_ARG(losses && weights && out && n > 0, "st_loss_total: null pointer or no loss nodes");
    return ew_loss_total(losses, weights, n, extra0, scale0, extra1, scale1, out, STRM(stream));
}
int st_absmax_slot(const float* x, size_t n, float target_log2, float* slot, float* ws, void* stream) {
    ST_CHECK_ARG(x && slot && ws && n > 0, "st_absmax_slot: null pointer or empty array");
    return ew_absmax_slot(x, n, target_log2, slot, ws, STRM(stream));
}
int st_mse_fwd(const void* x_hi, const void* x_lo, int x_planes, const void* t_hi, const void* t_lo, int t_planes, size_t numel,
               int relu, int feature_form, float* loss, float* ws, void* stream) {
    ST_CHECK_ARG(x_hi && t_hi && loss && ws && numel % 8 == 0 && numel > 0, "st_mse_fwd: null pointer or numel %% 8 != 0");
    ST_CHECK_ARG(plane_code_ok(x_planes) && plane_code_ok(t_planes), "st_mse_fwd: bad plane code");
    return ew_mse_fwd(CBF(x_hi), CBF(x_lo), one_plane_if_no_lo(x_planes, x_lo), CBF(t_hi), CBF(t_lo), one_plane_if_no_lo(t_planes, t_lo),
                      numel, relu, feature_form, loss, ws, STRM(stream));
}
int st_mse_bwd(const void* x_hi, const void* x_lo, int x_planes, const void* t_hi, const void* t_lo, int t_planes, size_t numel,
               int relu, float scale, const float* gw, const float* mul, int accumulate, void* g_hi, void* g_lo, void* stream) {
    ST_CHECK_ARG(x_hi && t_hi && gw && g_hi && numel % 8 == 0 && numel > 0, "st_mse_bwd: null pointer or numel %% 8 != 0");
    ST_CHECK_ARG(plane_code_ok(x_planes) && plane_code_ok(t_planes), "st_mse_bwd: bad plane code");
    return ew_mse_bwd(CBF(x_hi), CBF(x_lo), one_plane_if_no_lo(x_planes, x_lo), CBF(t_hi), CBF(t_lo), one_plane_if_no_lo(t_planes, t_lo),
                      numel, relu, scale, gw, mul, accumulate, BF(g_hi), BF(g_lo), STRM(stream));
}
int st_tv_fwd(const float* x, int N, int C, int H, int W, float factor, float* loss, float* ws, void* stream) {
    ST_CHECK_ARG(x && loss && ws && N > 0 && C > 0 && H > 0 && W > 0, "st_tv_fwd: null pointer or empty shape");
    return tv_forward(x, static_cast<size_t>(N) * C * H * W, H, W, factor, loss, ws, STRM(stream));
}
int st_tv_bwd(const float* x, int N, int C, int H, int W, float factor, const float* gw, float* g, void* stream) {
    ST_CHECK_ARG(x && gw && g && N > 0 && C > 0 && H > 0 && W > 0, "st_tv_bwd: null pointer or empty shape");
    return tv_backward(x, static_cast<size_t>(N) * C * H * W, H, W, factor, gw, g, STRM(stream));
}
int st_temporal_fwd(const float* oc, const float* os, const float* cc, const float* cs, size_t numel, float tw,
                    float* out3, float* ws, void* stream) {
    ST_CHECK_ARG(oc && os && cc && cs && out3 && ws && numel > 0, "st_temporal_fwd: null pointer or empty shape");
    return temporal_forward(oc, os, cc, cs, numel, tw, out3, ws, STRM(stream));
}
int st_temporal_warp_fwd(const float* oc, const float* os, const float* cc, const float* cs, const float* flow, int N, int C, int H,
                         int W, float tw, float* out3, float* ws, void* stream) {
    ST_CHECK_ARG(oc && os && cc && cs && flow && out3 && ws && N > 0 && C > 0 && H > 0 && W > 0, "st_temporal_warp_fwd: null pointer or empty shape");
    return temporal_warp_forward(oc, os, cc, cs, flow, N, C, H, W, tw, out3, ws, STRM(stream));
}
int st_temporal_warp_bwd(const float* os, const float* cs, const float* flow, int N, int C, int H, int W, float tw,
                         const float* stats3, const float* gw, float* g, void* stream) {
    ST_CHECK_ARG(os && cs && flow && stats3 && gw && g && N > 0 && C > 0 && H > 0 && W > 0, "st_temporal_warp_bwd: null pointer or empty shape");
    return temporal_warp_backward(os, cs, flow, N, C, H, W, tw, stats3, gw, g, STRM(stream));
}
int st_temporal_bwd(const float* os, const float* cs, size_t numel, float tw, const float* stats3, const float* gw, float* g, void* stream) {
    ST_CHECK_ARG(os && cs && stats3 && gw && g && numel > 0, "st_temporal_bwd: null pointer or empty shape");
    return temporal_backward(os, cs, numel, tw, stats3, gw, g, STRM(stream));
}

size_t st_instnorm_workspace_bytes(int N, int C) { return (in_workspace_floats(N, C) + 2 * static_cast<size_t>(N) * C) * sizeof(float); }
static bool in_channels_ok(int C) { return C >= 8 && C <= 256 && (C & (C - 1)) == 0; }
int st_instnorm_fwd(const void* x_hi, const void* x_lo, const float* gamma, const float* beta, float eps,
                    void* y_hi, void* y_lo, int planes, float* mean, float* rstd, int N, int H, int W, int C, int relu, int up,
                    void* workspace, void* stream) {
    ST_CHECK_ARG(x_hi && gamma && beta && y_hi && mean && rstd && workspace, "st_instnorm_fwd: null pointer");
    ST_CHECK_ARG(in_channels_ok(C), "st_instnorm_fwd: C must be a power of two in [8,256] (got %d)", C);
    ST_CHECK_ARG(plane_code_ok(planes), "st_instnorm_fwd: bad plane code %d", planes);
    return in_forward(CBF(x_hi), CBF(x_lo), gamma, beta, eps, BF(y_hi), BF(y_lo), one_plane_if_no_lo(planes, x_lo), mean, rstd, N, H, W, C, relu, up,
                      static_cast<float*>(workspace), STRM(stream));
}
int st_instnorm_bwd(const void* x_hi, const void* x_lo, const void* gy_hi, const void* gy_lo,
                    const float* mean, const float* rstd, const float* gamma, const float* beta,
                    void* dx_hi, void* dx_lo, int dx_planes, float* dx_slot, float* dgamma, float* dbeta, float* dsum, int x_planes,
                    int N, int H, int W, int C, int relu, int up, void* workspace, void* stream) {
    ST_CHECK_ARG(x_hi && gy_hi && mean && rstd && gamma && beta && dx_hi && dgamma && dbeta && workspace, "st_instnorm_bwd: null pointer");
    ST_CHECK_ARG(in_channels_ok(C), "st_instnorm_bwd: C must be a power of two in [8,256] (got %d)", C);
    ST_CHECK_ARG(plane_code_ok(dx_planes) && plane_code_ok(x_planes), "st_instnorm_bwd: bad plane code");
    ST_CHECK_ARG(!pc_f16(dx_planes) || dx_slot, "st_instnorm_bwd: fp16 gradient planes need a scale slot");
    return in_backward(CBF(x_hi), CBF(x_lo), CBF(gy_hi), CBF(gy_lo), mean, rstd, gamma, beta, BF(dx_hi), BF(dx_lo),
                       one_plane_if_no_lo(dx_planes, dx_lo), pc_f16(dx_planes) ? dx_slot : nullptr, dgamma, dbeta, dsum,
                       one_plane_if_no_lo(x_planes, x_lo), N, H, W, C, relu, up, static_cast<float*>(workspace), STRM(stream));
}

}  // extern "C"
