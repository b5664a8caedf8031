// First VGG layer (conv1_1: 3 -> 64 channels, 3x3, zero padding 1) forward and its
// backward-data to the image (reference: torchvision VGG features[0] run by
// StyleNetwork.run_through_pieces, network.py:316-340, and autograd's backward to
// `input_image`, network.py:451).
//
// With K = 27 a tensor-core tile would be > 50 % padding and the layer is bound by writing
// its 64-channel output (768 B per pixel for tap + ReLU copy in 3 bf16 planes), so both
// directions are CUDA-core kernels in exact fp32: one thread owns 8 output channels of a pixel
// (forward) or 8 gradient channels of a tap (backward), 8 consecutive lanes cover one pixel,
// so that every global access of a warp is 4 full 128-byte lines.
//   forward : reads the image as NCHW fp32 (module boundary format, no layout pass),
//             writes the pre-ReLU tap and optionally relu(tap), split-bf16 NHWC.
//   backward: reads the tap gradient (split-bf16 NHWC), writes the image gradient NCHW fp32.
#include "common.cuh"

namespace {

constexpr int IMG_THREADS = 256;

// weights in shared memory as [tap = (c*3 + r)*3 + s][Cout] fp32
__global__ void __launch_bounds__(IMG_THREADS)
conv_image_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                      bf16* __restrict__ out_hi, bf16* __restrict__ out_lo, bf16* __restrict__ relu_hi, bf16* __restrict__ relu_lo,
                      int planes, int N, int H, int W, int Cout) {
    extern __shared__ float sw[];                 // [27][Cout] + [Cout]
    float* sb = sw + 27 * Cout;
    for (int i = threadIdx.x; i < 27 * Cout; i += IMG_THREADS) {
        const int co = i % Cout, tap = i / Cout;                  // w: [Cout][3][3][3] -> tap index = c*9 + r*3 + s
        sw[i] = w[co * 27 + tap];
    }
    for (int i = threadIdx.x; i < Cout; i += IMG_THREADS) sb[i] = bias ? bias[i] : 0.f;
    __syncthreads();
    const int groups = Cout / 8;
    const size_t HW = static_cast<size_t>(H) * W;
    const size_t total = static_cast<size_t>(N) * HW * groups;
    for (size_t idx = static_cast<size_t>(blockIdx.x) * IMG_THREADS + threadIdx.x; idx < total;
         idx += static_cast<size_t>(gridDim.x) * IMG_THREADS) {
        const int g = idx % groups; const size_t pix = idx / groups;
        const int n = pix / HW; const int rem = pix - static_cast<size_t>(n) * HW;
        const int h = rem / W, wq = rem - h * W;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = sb[g * 8 + j];
        const float* xn = x + static_cast<size_t>(n) * 3 * HW;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int ih = h + r - 1;
#pragma unroll
                for (int s = 0; s < 3; ++s) {
                    const int iw = wq + s - 1;
                    const bool ok = ih >= 0 && ih < H && iw >= 0 && iw < W;
                    const float v = ok ? __ldg(xn + c * HW + static_cast<size_t>(ih) * W + iw) : 0.f;
                    const float4 w0 = *reinterpret_cast<const float4*>(sw + ((c * 3 + r) * 3 + s) * Cout + g * 8);
                    const float4 w1 = *reinterpret_cast<const float4*>(sw + ((c * 3 + r) * 3 + s) * Cout + g * 8 + 4);
                    acc[0] = fmaf(v, w0.x, acc[0]); acc[1] = fmaf(v, w0.y, acc[1]);
                    acc[2] = fmaf(v, w0.z, acc[2]); acc[3] = fmaf(v, w0.w, acc[3]);
                    acc[4] = fmaf(v, w1.x, acc[4]); acc[5] = fmaf(v, w1.y, acc[5]);
                    acc[6] = fmaf(v, w1.z, acc[6]); acc[7] = fmaf(v, w1.w, acc[7]);
                }
            }
        }
        const size_t o = pix * Cout + g * 8;
        if (out_hi) store8p(out_hi, out_lo, planes, o, acc);
        if (relu_hi) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = fmaxf(acc[j], 0.f);
            store8p(relu_hi, relu_lo, planes, o, acc);
        }
    }
}

// gx[n][c][h][w] = sum_{co,r,s} g[n][h+1-r][w+1-s][co] * w[co][c][r][s]
// 8 lanes per output pixel, lane j owns gradient channels j*8 .. j*8+7 (+ 64 per extra round);
// the 3 channel sums are combined with warp shuffles.
__global__ void __launch_bounds__(IMG_THREADS)
conv_image_dgrad_kernel(const bf16* __restrict__ g_hi, const bf16* __restrict__ g_lo, int g_planes,
                        const float* __restrict__ w, float* __restrict__ gx, int N, int H, int W, int Cout) {
    extern __shared__ float sw[];                 // [9 taps][3 c][Cout]
    for (int i = threadIdx.x; i < 27 * Cout; i += IMG_THREADS) {
        const int co = i % Cout; int q = i / Cout; const int c = q % 3; const int tap = q / 3;   // tap = r*3 + s
        sw[i] = w[(co * 3 + c) * 9 + tap];
    }
    __syncthreads();
    const size_t HW = static_cast<size_t>(H) * W;
    const size_t total = static_cast<size_t>(N) * HW * 8;
    const size_t total_pad = (total + 31) / 32 * 32;          // whole warps stay in the loop for the shuffles
    for (size_t idx = static_cast<size_t>(blockIdx.x) * IMG_THREADS + threadIdx.x; idx < total_pad;
         idx += static_cast<size_t>(gridDim.x) * IMG_THREADS) {
        const int j = idx & 7; const size_t pix = idx >> 3;
        const bool live = idx < total;
        const int n = live ? pix / HW : 0; const int rem = live ? pix - static_cast<size_t>(n) * HW : 0;
        const int h = rem / W, wq = rem - h * W;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f;
        if (live) {
            for (int cb = j * 8; cb < Cout; cb += 64) {
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const int ih = h + 1 - r;
                    if (ih < 0 || ih >= H) continue;
#pragma unroll
                    for (int s = 0; s < 3; ++s) {
                        const int iw = wq + 1 - s;
                        if (iw < 0 || iw >= W) continue;
                        float v[8];
                        load8p(g_hi, g_lo, g_planes, ((static_cast<size_t>(n) * H + ih) * W + iw) * Cout + cb, v);
                        const float* wt = sw + (r * 3 + s) * 3 * Cout + cb;
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            a0 = fmaf(v[k], wt[k], a0);
                            a1 = fmaf(v[k], wt[Cout + k], a1);
                            a2 = fmaf(v[k], wt[2 * Cout + k], a2);
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            a0 += __shfl_xor_sync(0xffffffffu, a0, o);
            a1 += __shfl_xor_sync(0xffffffffu, a1, o);
            a2 += __shfl_xor_sync(0xffffffffu, a2, o);
        }
        if (live && j < 3) {
            const float v = j == 0 ? a0 : (j == 1 ? a1 : a2);
            gx[(static_cast<size_t>(n) * 3 + j) * HW + rem] = v;
        }
    }
}

inline int img_grid(size_t work) {
    size_t g = (work + IMG_THREADS - 1) / IMG_THREADS;
    const size_t cap = static_cast<size_t>(st_num_sms()) * 8;
    return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

int conv_image_fwd(const float* x, const float* w, const float* bias, bf16* out_hi, bf16* out_lo,
                   bf16* relu_hi, bf16* relu_lo, int planes, int N, int H, int W, int Cout, cudaStream_t st) {
    const size_t work = static_cast<size_t>(N) * H * W * (Cout / 8);
    conv_image_fwd_kernel<<<img_grid(work), IMG_THREADS, 28 * Cout * sizeof(float), st>>>(
        x, w, bias, out_hi, out_lo, relu_hi, relu_lo, planes, N, H, W, Cout);
    ST_CHECK_LAUNCH("conv_image_fwd_kernel");
    return ST_OK;
}

int conv_image_dgrad(const bf16* g_hi, const bf16* g_lo, int g_planes, const float* w, float* gx,
                     int N, int H, int W, int Cout, cudaStream_t st) {
    const size_t work = static_cast<size_t>(N) * H * W * 8;
    conv_image_dgrad_kernel<<<img_grid(work), IMG_THREADS, 27 * Cout * sizeof(float), st>>>(
        g_hi, g_lo, g_planes, w, gx, N, H, W, Cout);
    ST_CHECK_LAUNCH("conv_image_dgrad_kernel");
    return ST_OK;
}
