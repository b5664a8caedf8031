/* libstransfer_b200.so -- C ABI of the B200 (sm_100a) kernels behind
 * tupini07/StyleTransfer's `stransfer.network` hot path.
 *
 * The reference has no FFI of its own: every numeric op is a stock torch call.
 * Each entry point below replaces the torch call(s) cited beside it
 * (file:line into /root/reference/stransfer/network.py); INTEGRATION.md shows
 * the ctypes binding + torch.autograd.Function a maintainer would add there.
 *
 * Conventions
 *  - every function returns 0 on success or a negative ST_ERR_* code;
 *    st_last_error() returns a thread-local message for the last failure.
 *  - all data pointers are DEVICE pointers owned by the caller; nothing is
 *    allocated inside; scratch comes from a caller-provided workspace whose
 *    size the matching *_workspace_bytes() query returns.
 *  - every launch is stream-ordered on `stream` (a cudaStream_t); no hidden
 *    synchronisation, so the calls can be captured in a CUDA graph.
 *  - activations use the "split 16-bit NHWC" format: channels-last, two 16-bit
 *    planes hi = rn16(x), lo = rn16(x - hi).  A NULL lo plane selects a single
 *    plane (1 tensor-core pass per product instead of 3).  Every `planes`
 *    argument is a PLANE CODE: bits 0..3 = number of planes, bit 4
 *    (ST_PLANES_F16) = the planes hold fp16 instead of bf16:
 *      bf16, 3 planes: a THIRD plane bf16(x - hi - lo) lives one plane stride
 *        after lo (at lo + (lo - hi)): ~24 mantissa bits, 6 tensor-core passes
 *        per product;
 *      fp16, 2 planes: 22 mantissa bits in 3 passes; values saturate at
 *        +-65504 and keep an absolute precision of 2^-25 below 2^-3.
 *    Forward activations need > 16 bits: ReLU masks / max-pool argmaxes flip on
 *    1e-5 errors and each flip costs a whole gradient entry.
 *    The tensor cores multiply two operands of ONE format (fp16 x bf16 is an
 *    illegal tcgen05.mma): where the operands of a product differ, the mma.sync
 *    kernel that converts on load is used instead.
 *  - a tensor whose values may leave the fp16 range (gradients) is stored
 *    multiplied by a power of two held in a 2-float device SCALE SLOT
 *    {s, 1/s}; its producer writes the slot, its consumers take the slot and
 *    multiply their results by slot[1].  fp16-packed conv weights carry the
 *    fixed factor 256 (acc_mul = 1/256 at the consumer).
 *    Channel counts are multiples of 8 (images are padded 3/6 -> 8 with zeros).
 *  - scalars that come from autograd (loss weights = grad_output entries) are
 *    passed as device pointers (`gw`) so that no host sync is needed.
 */
#ifndef STRANSFER_B200_H
#define STRANSFER_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ST_OK 0
#define ST_ERR_INVALID (-1)
#define ST_ERR_CUDA (-2)
#define ST_ERR_UNSUPPORTED (-3)
#define ST_PLANES_F16 16     /* plane-code bit: fp16 planes */

/* build info: compute capability the kernels were compiled for (100) */
int st_arch(void);
const char* st_last_error(void);
/* 0 = auto (tcgen05 where eligible), 1 = force the generic mma.sync conv (testing), 2 = auto, but the image gradient of
 * conv1_1 on the 16-channel tile of conv_tc instead of conv_imgrad_tc (testing) */
int st_set_conv_path(int path);
/* test hook: which kernel the last st_conv2d call launched: 1 = conv_tc (tcgen05, stride 1),
 * 2 = conv_tcx (tcgen05, run-time geometry), 3 = conv_gen (mma.sync), 4 = conv_imgrad_tc (tcgen05, the image gradient
 * of conv1_1 in project-then-gather form) */
int st_last_conv_kernel(void);
/* number of kernels this library has launched so far in this process */
long long st_launch_count(void);

/* ---- layout / packing ------------------------------------------------- */
/* img_utils/torch.cat boundary: NCHW fp32 (optionally two tensors concatenated on
 * channels, network.py:984-986) -> split-bf16 NHWC with Cpad channels. */
int st_nchw_to_nhwc(const float* a, const float* b, void* hi, void* lo, int planes,
                    int N, int C1, int C2, int H, int W, int Cpad, void* stream);
/* slot (optional): scale slot of the tensor; the values are multiplied by slot[1] */
int st_nhwc_to_nchw(const void* hi, const void* lo, int planes, const float* slot, float* out,
                    int N, int C, int H, int W, int Cpad, void* stream);
/* Image boundary on the device (reference img_utils.py:33-42 ToTensor + Normalize, and :105-109 de-normalise + clamp
 * of imshow): u8 HWC [N][H][W][3] (device) <-> normalised NCHW fp32 [N][3][H][W]; mean3 / std3 are HOST arrays of
 * three floats (constants.py:16-17).  Separately rounded IEEE operations: bit-identical to the reference's CPU code. */
int st_image_u8_to_nchw(const unsigned char* in, float* out, int N, int H, int W, const float* mean3, const float* std3, void* stream);
int st_image_nchw_to_u8(const float* in, unsigned char* out, int N, int H, int W, const float* mean3, const float* std3, void* stream);
/* nn.Conv2d.weight [Cout][Cin][ks][ks] fp32 -> packed [rows_pad][ks*ks][cols_pad] split-bf16.
 * mode 0 forward; 1 backward-data (gather form, any stride); 2 backward-data as a
 * forward conv with flipped taps (stride 1). */
int st_pack_conv_weight(const float* w, void* hi, void* lo, int planes, int Cout, int Cin, int ks,
                        int rows_pad, int cols_pad, int mode, void* stream);

/* ---- convolution ------------------------------------------------------ */
/* nn.Conv2d forward (network.py:246-275 VGG, :468-481, :525-609 transform net) and
 * its autograd backward-data.  y = conv(x, w) [+bias] ; y *= (mask > 0) ; y += add ;
 * out = y ; relu_out = relu(y).  `up`=1 reads x through a nearest x2 upsample
 * (nn.Upsample network.py:580-594).  `transposed`=1 evaluates the backward-data
 * gather (weights packed with mode 1).  w_img_rows != 0 selects per-image weights
 * (rows n*w_img_rows.. of the packed matrix): used for the Gram backward dF = F*D.
 * out_nchw (optional) receives the first Cnchw channels as NCHW fp32.
 * x_planes / w_planes / add_planes / out_planes: plane codes (see the conventions above);
 * `out` and `relu_out` carry out_planes; `mask` is read from its hi plane only (sign test,
 * either format).  Before the bias the accumulator is multiplied by
 * acc_mul * (acc_mul_slot ? acc_mul_slot[1] : 1) (1/256 for fp16-packed weights; the scale slot of a
 * scaled operand); the addend is multiplied by add_mul_slot[1] when that slot is given. */
int st_conv2d(const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int Cin, int up,
              const void* w_hi, const void* w_lo, int w_planes, int w_rows_total, int w_img_rows,
              int ks, int stride, int pad, int transposed,
              int Ho, int Wo, int Cout_pad, int Cbias,
              const float* bias, const void* mask, const void* add_hi, const void* add_lo, int add_planes,
              void* out_hi, void* out_lo, void* relu_hi, void* relu_lo, int out_planes,
              float* out_nchw, int Cnchw, float acc_mul, const float* acc_mul_slot, const float* add_mul_slot,
              void* stream);

/* st_conv2d (stride 1, pad ks/2, no bias) whose addend is the gradient gy of the OUTPUT of the 2x2 max-pool that
 * follows this tensor, [N][H/2][W/2][Cout_pad], routed on the fly: out(h, w, c) = conv [* (mask > 0)] +
 * (route[n][h/2][w/2][c] == (h&1)*2 + (w&1) ? gy[n][h/2][w/2][c] * gy_mul_slot[1] : 0), route = the byte map of
 * st_relu_maxpool2x2_fwd_idx (4 = window maximum <= 0; rows / columns beyond 2*(H/2), 2*(W/2) receive no gradient).
 * It replaces st_relu_maxpool2x2_bwd_idx + st_conv2d(add=...) in the backward of a VGG block that ends in a pool and
 * carries a style node: torch autograd's MaxPool2d / ReLU backward of the pieces reference network.py:257-314 builds,
 * followed by the Gram backward of StyleLoss (network.py:92-108, :127-168).  Tensor-core path only: returns
 * ST_ERR_UNSUPPORTED where st_conv2d would have taken another kernel (the caller then runs the two-launch form). */
int st_conv2d_routed(const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int Cin,
                     const void* w_hi, const void* w_lo, int w_planes, int w_rows_total, int w_img_rows, int ks, int Cout_pad,
                     const void* mask, const void* gy_hi, const void* gy_lo, int gy_planes, const unsigned char* route,
                     void* out_hi, void* out_lo, int out_planes, float acc_mul, const float* acc_mul_slot, const float* gy_mul_slot,
                     void* stream);

/* st_conv2d (3x3, stride 1, pad 1, shared weights) that also writes the output of the ReLU + MaxPool2d(2, 2) behind it:
 * pool = relu(maxpool2x2(conv + bias)) [N][H/2][W/2][Cout_pad] in the planes of `out`, and one routing byte per pooled
 * element (window position 0..3 of the first maximum in PyTorch's scan order, 4 = maximum <= 0) for
 * st_relu_maxpool2x2_bwd_idx / st_conv2d_routed.  `out` (the pre-ReLU tap) is optional.  Replaces st_conv2d +
 * st_relu_maxpool2x2_fwd_idx for the convs in front of VGG pieces 2 and 5 (nn.Conv2d -> nn.ReLU -> nn.MaxPool2d of the
 * pieces reference network.py:257-314 builds): the pool no longer re-reads the full-resolution tap.  Tensor-core path
 * only: ST_ERR_UNSUPPORTED where st_conv2d would take another kernel (the caller then runs the two launches). */
int st_conv2d_pool(const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int Cin,
                   const void* w_hi, const void* w_lo, int w_planes, int w_rows_total, int ks, int Cout_pad, int Cbias, const float* bias,
                   void* out_hi, void* out_lo, int out_planes, void* pool_hi, void* pool_lo, unsigned char* pool_idx,
                   float acc_mul, const float* acc_mul_slot, void* stream);

/* First VGG layer (torchvision vgg19.features[0], 3 -> Cout channels, 3x3, zero padding 1;
 * run by StyleNetwork.run_through_pieces network.py:316-340) in exact fp32 on the CUDA cores:
 * x is the NCHW fp32 image (module boundary format), w [Cout][3][3][3] and bias [Cout] fp32;
 * out = pre-ReLU tap, relu_out = relu(tap) (either may be NULL), split-bf16 NHWC with `planes`
 * planes.  Cout: multiple of 8, <= 256. */
int st_conv3x3_image_fwd(const float* x, const float* w, const float* bias, void* out_hi, void* out_lo,
                         void* relu_hi, void* relu_lo, int planes, int N, int H, int W, int Cout, void* stream);
/* its backward-data to the image (autograd of network.py:451): g = gradient of the tap
 * (split-bf16 NHWC, g_planes planes), gx NCHW fp32 [N][3][H][W]. */
int st_conv3x3_image_dgrad(const void* g_hi, const void* g_lo, int g_planes, const float* w, float* gx,
                           int N, int H, int W, int Cout, void* stream);

/* ---- the transform net's 9x9 layers (network.py:525-530 / :850-855: 3|6 -> 32; :604-609: 32 -> 3)
 * as exact-fp32 direct convolutions on the CUDA cores (one side has 3 / 6 channels).
 * "Small" tensors are NCHW fp32 (module boundary): channels 0..2 are read from s0, 3..5 from s1,
 * both with `sc` channels per sample (a 6-channel tensor x: s0 = x, s1 = x + 3*H*W, sc = 6).
 * "Big" tensors have 32 channels, split-bf16 NHWC. */
/* small -> big.  flip = 0: forward of layer 0, w [32][CS][9][9], bias [32].
 *                flip = 1: backward-data of layer 22, small = dy, w [CS][32][9][9] (the layer's weight). */
int st_conv9_expand(const float* s0, const float* s1, int sc, int CS, const float* w, const float* bias, int flip,
                    void* out_hi, void* out_lo, int planes, int N, int H, int W, void* stream);
/* big -> small: forward of layer 22, w [3][32][9][9], bias [3], out NCHW fp32 [N][3][H][W].
 * workspace (st_conv9_reduce_workspace_bytes) holds the fp32 partial outputs of the vertical stage of the
 * tensor-core path; with workspace == NULL the CUDA-core kernel runs. */
size_t st_conv9_reduce_workspace_bytes(int N, int H, int W);
int st_conv9_reduce(const void* x_hi, const void* x_lo, int planes, const float* w, const float* bias, float* out,
                    void* workspace, size_t workspace_bytes, int N, int H, int W, void* stream);
/* backward-weight (+ bias).  mode 0: layer 0, big = dy, small = the layer's input -> dw [32][CS][9][9], db [32];
 *                            mode 1: layer 22, big = the layer's input, small = dy -> dw [CS][32][9][9], db [CS] */
/* big_slot (optional): scale slot of the big tensor (a scaled fp16 gradient); small_slot (optional): when the big
 * tensor has fp16 planes the small fp32 tensor is multiplied by small_slot[0] on its way into fp16 operands
 * (st_absmax_slot of a gradient image); dw / db come out unscaled. */
size_t st_conv9_wgrad_workspace_bytes(int CS);
int st_conv9_wgrad(const void* b_hi, const void* b_lo, int b_planes, const float* big_slot, const float* small_slot,
                   const float* s0, const float* s1, int sc, int CS, int mode,
                   float* dw, float* db, void* workspace, size_t workspace_bytes, int N, int H, int W, void* stream);

/* autograd backward-weight of nn.Conv2d: dw [Cout][Cin][ks][ks] fp32 (+ db [Cout]).  dy_planes / x_planes: plane
 * codes; dy_slot (optional): scale slot of dy (undone in dw / db). */
size_t st_conv2d_wgrad_workspace_bytes(int N, int Ho, int Wo, int Cdy, int CinPad, int ks);
int st_conv2d_wgrad(const void* dy_hi, const void* dy_lo, int dy_planes, const float* dy_slot, int Cdy,
                    const void* x_hi, const void* x_lo, int x_planes, int N, int H, int W, int CinPad, int up,
                    int Ho, int Wo, int ks, int stride, int pad, int Cout, int Cin,
                    float* dw, float* db, void* workspace, size_t workspace_bytes, void* stream);

/* ---- VGG pieces -------------------------------------------------------- */
/* nn.ReLU + nn.MaxPool2d(2,2) of VGG pieces 2 and 5 (network.py:257-314) */
int st_relu_maxpool2x2_fwd(const void* x_hi, const void* x_lo, void* y_hi, void* y_lo, int planes,
                           int N, int H, int W, int C, void* stream);
/* backward: gx = route(gy) * relu'(x) + add */
int st_relu_maxpool2x2_bwd(const void* x_hi, const void* x_lo, const void* gy_hi, const void* gy_lo,
                           const void* add_hi, const void* add_lo, void* gx_hi, void* gx_lo, int x_planes,
                           int N, int H, int W, int C, void* stream);
/* The same pair with the routing decision saved by the forward: arg [N][H/2][W/2][C] bytes = position 0..3 of the
 * first maximum of relu(x) in the window, 4 = nothing routed (relu'(x) = 0).  The backward reads 1 byte per pooled
 * element instead of every plane of the full-resolution tap. */
int st_relu_maxpool2x2_fwd_idx(const void* x_hi, const void* x_lo, void* y_hi, void* y_lo, unsigned char* arg, int planes,
                               int N, int H, int W, int C, void* stream);
int st_relu_maxpool2x2_bwd_idx(const unsigned char* arg, const void* gy_hi, const void* gy_lo, const void* add_hi,
                               const void* add_lo, void* gx_hi, void* gx_lo, int N, int H, int W, int C, void* stream);

/* ---- losses ------------------------------------------------------------ */
/* StyleLoss.gram_matrix (network.py:92-108): G[b] = F_b^T F_b / (C*H*W), fp32 [B][C][C] */
size_t st_gram_workspace_bytes(int B, int HW, int C);
/* tuning hook: cap on the pixel split-K of the Gram kernel (0 = automatic); returns the old value */
int st_set_gram_max_splits(int max_splits);
int st_gram_fwd(const void* f_hi, const void* f_lo, int f_planes, int B, int HW, int C, float* G,
                void* workspace, size_t workspace_bytes, void* stream);
/* StyleLoss.gram_matrix + StyleLoss.forward (network.py:92-123) for one node in TWO launches: the Gram kernel and one
 * finish launch that reduces the split partials and also writes diff = G - T, loss = mean(diff^2) and max|diff|
 * (three launches fewer per style node than st_gram_fwd + st_style_mse_fwd).  ticket: a device int that is zero before
 * the first call; the kernel hands it back at zero (calls that share it must be stream-ordered). */
size_t st_gram_style_workspace_bytes(int B, int HW, int C);
int st_gram_style_fwd(const void* f_hi, const void* f_lo, int f_planes, int B, int HW, int C, const float* T,
                      float* diff, float* loss, float* diff_amax, int* ticket, void* workspace, size_t workspace_bytes,
                      void* stream);
/* StyleLoss.forward (network.py:110-123): loss = mean((G - T)^2), T [C][C] broadcast;
 * diff = G - T is kept for the backward; diff_amax (optional) receives max |diff|.  ws: >= 4096 floats */
int st_style_mse_fwd(const float* G, const float* T, int B, int C, float* diff, float* loss, float* diff_amax,
                     float* ws, void* stream);
/* backward seed: D = (*gw) * 4/(B*C^2*C*HW) * diff, packed [B*C][C] with plane code d_planes;
 * dF = F * D is then st_conv2d(ks=1, w_img_rows=C).  fp16 planes: D is stored multiplied by the power of two
 * that maps max|D| (from diff_amax) to 2^10 and d_slot receives {s, 1/s} (acc_mul_slot of that conv). */
int st_style_mse_bwd_seed(const float* diff, int B, int C, int HW, const float* gw, void* d_hi, void* d_lo,
                          int d_planes, const float* diff_amax, float* d_slot, void* stream);
/* st_style_mse_bwd_seed for every style node of a loss network in one launch (n <= 16): tables of n entries each */
int st_style_mse_bwd_seed_multi(int n, const float* const* diff, const int* B, const int* C, const int* HW,
                                const float* const* gw, void* const* d_hi, void* const* d_lo, const int* d_planes,
                                const float* const* diff_amax, float* const* d_slot, void* stream);
/* The value a training closure returns (network.py:704-722, :1020): total = sum_i losses[i] * weights[i]
 * + extra0 * scale0 + extra1 * scale1 (total-variation and temporal terms; either may be null), one launch instead of
 * the stack / sum / mul / add chain of the loss getters (network.py:358-364). */
int st_loss_total(const float* losses, const float* weights, int n, const float* extra0, float scale0,
                  const float* extra1, float scale1, float* out, void* stream);
/* slot = {s, 1/s}, s the power of two that maps max|x| of a flat fp32 array to 2^target_log2 (the scale under which
 * a gradient image enters fp16 tensor-core operands).  ws: >= 1024 floats */
int st_absmax_slot(const float* x, size_t n, float target_log2, float* slot, float* ws, void* stream);
/* ContentLoss (network.py:155-164) / FeatureReconstructionLoss (network.py:186-201):
 * relu=1 compares relu(x) with relu(t); feature_form=1 returns mse^2/numel. */
int st_mse_fwd(const void* x_hi, const void* x_lo, int x_planes, const void* t_hi, const void* t_lo, int t_planes,
               size_t numel, int relu, int feature_form, float* loss, float* ws, void* stream);
/* g (+)= scale * (*gw) * (mul ? *mul : 1) * (f(x) - f(t)) * f'(x); g: two bf16 planes */
int st_mse_bwd(const void* x_hi, const void* x_lo, int x_planes, const void* t_hi, const void* t_lo, int t_planes,
               size_t numel, int relu, float scale, const float* gw, const float* mul, int accumulate,
               void* g_hi, void* g_lo, void* stream);
/* total variation (network.py:621-641) on NCHW fp32; ws >= 2048 floats */
int st_tv_fwd(const float* x, int N, int C, int H, int W, float factor, float* loss, float* ws, void* stream);
int st_tv_bwd(const float* x, int N, int C, int H, int W, float factor, const float* gw, float* g, void* stream);
/* temporal loss (network.py:885-903) on NCHW fp32: out3 = {loss, ||cs-os||, ||cc-oc||}; ws >= 2048 floats */
int st_temporal_fwd(const float* old_content, const float* old_stylized, const float* cur_content,
                    const float* cur_stylized, size_t numel, float tw, float* out3, float* ws, void* stream);
int st_temporal_bwd(const float* old_stylized, const float* cur_stylized, size_t numel, float tw,
                    const float* stats3, const float* gw, float* g, void* stream);
/* EXTENSION without reference code (the reference's temporal loss has no optical flow; BASELINE.json
 * words the loss as flow-warped): the previous content / stylised frames are warped by flow [N][2][H][W] (dx, dy in
 * pixels; bilinear, zeros outside = torch grid_sample(align_corners=True) in pixel units) on the fly:
 * loss = tw * ||cs - warp(os)|| / (||cc - warp(oc)|| + 1); out3 / stats3 / ws as st_temporal_fwd. */
int st_temporal_warp_fwd(const float* old_content, const float* old_stylized, const float* cur_content,
                         const float* cur_stylized, const float* flow, int N, int C, int H, int W, float tw,
                         float* out3, float* ws, void* stream);
int st_temporal_warp_bwd(const float* old_stylized, const float* cur_stylized, const float* flow, int N, int C, int H, int W,
                         float tw, const float* stats3, const float* gw, float* g, void* stream);

/* ---- InstanceNorm2d(affine) (network.py:474, :483, :531-600) ------------- */
size_t st_instnorm_workspace_bytes(int N, int C);
/* y = gamma*(x-mean)*rstd+beta, optional fused ReLU, optional x2 nearest upsample on write
 * (the nn.Upsample that follows, network.py:580, :592); mean/rstd [N][C] are saved for backward */
int st_instnorm_fwd(const void* x_hi, const void* x_lo, const float* gamma, const float* beta, float eps,
                    void* y_hi, void* y_lo, int planes, float* mean, float* rstd,
                    int N, int H, int W, int C, int relu, int up, void* workspace, void* stream);
/* gy: two bf16 planes.  dsum (optional, [C]): sum of dx over images and pixels = bias gradient of the convolution in
 * front.  dx_planes: plane code of dx; with fp16 planes dx is stored multiplied by a power of two chosen from a bound
 * of |dx| that the statistics pass measures, and dx_slot receives {s, 1/s} (dx feeds the backward-data and
 * backward-weight tensor-core products of the convolution in front). */
int st_instnorm_bwd(const void* x_hi, const void* x_lo, const void* gy_hi, const void* gy_lo,
                    const float* mean, const float* rstd, const float* gamma, const float* beta,
                    void* dx_hi, void* dx_lo, int dx_planes, float* dx_slot, float* dgamma, float* dbeta, float* dsum,
                    int x_planes, int N, int H, int W, int C, int relu, int up, void* workspace, void* stream);

/* st_instnorm_fwd without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid it).  Same kernels and results as st_instnorm_fwd; the statistics partials
 * live in a static device array of ST_INSTNORM_STATIC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit.  Limit: N*C <= ST_INSTNORM_STATIC_MAX_NC (e.g. 8 frames of 128 channels), larger problems are rejected with
 * ST_ERR_INVALID and a message naming the limit.  The array is shared by every call and keeps the last call's
 * contents (nothing in it needs to start at zero): calls on one stream are ordered, calls that overlap on the GPU
 * race.  The first call on a device resolves the array's device address (cudaGetSymbolAddress, no allocation). */
#define ST_INSTNORM_STATIC_MAX_NC 1024
#define ST_INSTNORM_STATIC_SCRATCH_BYTES (ST_INSTNORM_STATIC_MAX_NC * 128 * 2 * 4)
int st_instnorm_fwd_static(const void* x_hi, const void* x_lo, const float* gamma, const float* beta, float eps,
                           void* y_hi, void* y_lo, int planes, float* mean, float* rstd,
                           int N, int H, int W, int C, int relu, int up, void* stream);

/* ---- optimizer steps on the device ---------- */
/* torch.optim.Adam.step over every parameter tensor of a transform net in one launch
 * (network.py:643-649, :765).  param / grad / exp_avg / exp_avg_sq are flat fp32 arrays of
 * sum(seg_numel) elements, one segment per parameter tensor in order; `steps` / `active` are
 * device int[nseg] (torch's per-parameter step count; active = the tensor has a gradient).
 * `chunks` is a device copy of the table st_adam_fill_chunks() writes on the host
 * (st_adam_chunk_count() entries of 16 bytes).  grad_scale multiplies the gradient first
 * (1/world of the data-parallel average).  packs (optional): device array of nseg 88-byte records
 *   { int64 seg_off; int32 cout, cin, ks, ndst; { void* hi; void* lo; int32 rows_pad, cols_pad, mode, code; } dst[2]; }
 * describing the packed copies (st_pack_conv_weight layouts, plane code `code`) of each convolution weight: the
 * kernel rewrites them from the updated fp32 value, so no packing launch follows an optimizer step. */
size_t st_adam_chunk_count(const long long* seg_numel, int nseg);
int st_adam_fill_chunks(const long long* seg_numel, int nseg, void* host_chunks);
int st_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, const void* chunks,
                 size_t nchunks, int* steps, const int* active, int nseg, double lr, double beta1, double beta2,
                 double eps, double weight_decay, double grad_scale, const void* packs, void* stream);
/* One iteration of torch.optim.LBFGS (no line search; network.py:435, :456) on the flat fp32
 * vector x [n] given the gradient and the loss of the closure at x (device pointers): history
 * update, two-loop recursion in coefficient space, x += t*d, torch's stopping rules on a device
 * flag (first_of_step = 1 for the first iteration of an optimizer.step call: clears the flag).
 * g_prev, dir: [n] scratch owned by the caller; S, Y: history rings [history][ld], ld >= n,
 * ld % 4 == 0; state: st_lbfgs_state_bytes() bytes, set up by st_lbfgs_init().  The first 32
 * doubles of state are readable diagnostics (k, head, n_iter, t, H_diag, ...; see optim.cu). */
size_t st_lbfgs_state_bytes(int history, size_t n);
/* tuning hook: 0 (default) = dots pass with 3 CTAs per SM, 1 = 2 CTAs per SM with software-pipelined loads; returns the old value */
int st_set_lbfgs_variant(int variant);
int st_lbfgs_init(void* state, int history, size_t n, double lr, double tolerance_grad, double tolerance_change, void* stream);
int st_lbfgs_iteration(float* x, const float* grad, const float* loss, float* g_prev, float* dir, float* S, float* Y,
                       size_t n, size_t ld, int history, void* state, int first_of_step, void* stream);

#ifdef __cplusplus
}
#endif
#endif
