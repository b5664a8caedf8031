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
constexpr int IM_TH = 4, IM_TW = 32;               // output tile: 4 x 32 pixels
constexpr int IM_SH = IM_TH + 2;                   // staged rows
constexpr int IM_SW = IM_TW + 8;                   // staged fp32 image row: column x <-> iw = w0 + x - 4 (float4 aligned)
constexpr int IM_GW = IM_TW + 2;                   // staged gradient row (pixels)

// shared-memory channel permutation: channel ch = g*8 + half*4 + k lives at [half][g][k], so the 8
// lanes of a pixel read 128 contiguous bytes per LDS.128 (no bank conflicts)
__device__ __forceinline__ int perm64(int ch) { return ((ch >> 2) & 1) * 32 + (ch >> 3) * 4 + (ch & 3); }

// forward: 256 threads = 32 pixel quads (4 along w) x 8 channel groups; 4 px x 8 ch accumulators
__global__ void __launch_bounds__(IMG_THREADS)
conv_image_fwd4_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                      bf16* __restrict__ out_hi, bf16* __restrict__ out_lo, bf16* __restrict__ relu_hi, bf16* __restrict__ relu_lo,
                      int planes, int N, int H, int W, int tiles_h, int tiles_w) {
    pdl_prologue();
    __shared__ __align__(16) float sw[27 * 64];                      // [tap = (c*3 + r)*3 + s][perm64(co)]
    __shared__ __align__(16) float st[3 * IM_SH * IM_SW];
    for (int i = threadIdx.x; i < 27 * 64; i += IMG_THREADS) {
        const int co = i & 63, tap = i >> 6;                         // w: [64][3][3][3]
        sw[tap * 64 + perm64(co)] = w[co * 27 + tap];
    }
    const int g = threadIdx.x & 7, quad = threadIdx.x >> 3;
    const int qx = quad & 7, qy = quad >> 3;
    float b8[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) b8[j] = bias ? bias[g * 8 + j] : 0.f;
    const size_t HW = static_cast<size_t>(H) * W;
    const int tiles = N * tiles_h * tiles_w;
    // the staged image tile of the NEXT iteration is fetched into registers while this one is computed
    constexpr int ST_N = 3 * IM_SH * IM_SW, ST_PER = (ST_N + IMG_THREADS - 1) / IMG_THREADS;
    float pre[ST_PER];
    auto fetch = [&](int tile) {
        int q = tile; const int tw = q % tiles_w; q /= tiles_w; const int th = q % tiles_h; const int n = q / tiles_h;
#pragma unroll
        for (int k = 0; k < ST_PER; ++k) {
            const int i = threadIdx.x + k * IMG_THREADS;
            const int xx = i % IM_SW; int r = i / IM_SW; const int y = r % IM_SH; const int c = r / IM_SH;
            const int ih = th * IM_TH + y - 1, iw = tw * IM_TW + xx - 4;
            float v = 0.f;
            if (i < ST_N && ih >= 0 && ih < H && iw >= 0 && iw < W) v = __ldg(x + (static_cast<size_t>(n) * 3 + c) * HW + static_cast<size_t>(ih) * W + iw);
            pre[k] = v;
        }
    };
    if (blockIdx.x < tiles) fetch(blockIdx.x);
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int q = tile; const int tw = q % tiles_w; q /= tiles_w; const int th = q % tiles_h; const int n = q / tiles_h;
        const int h0 = th * IM_TH, w0 = tw * IM_TW;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < ST_PER; ++k) {
            const int i = threadIdx.x + k * IMG_THREADS;
            if (i < ST_N) st[i] = pre[k];
        }
        __syncthreads();
        if (tile + gridDim.x < tiles) fetch(tile + gridDim.x);
        float acc[4][8];
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[p][j] = b8[j];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const float* row = st + (c * IM_SH + qy + r) * IM_SW + qx * 4;
                float in[12];
#pragma unroll
                for (int v = 0; v < 3; ++v) {
                    const float4 t = *reinterpret_cast<const float4*>(row + v * 4);
                    in[v * 4] = t.x; in[v * 4 + 1] = t.y; in[v * 4 + 2] = t.z; in[v * 4 + 3] = t.w;
                }
#pragma unroll
                for (int sx = 0; sx < 3; ++sx) {
                    const float* wr = sw + ((c * 3 + r) * 3 + sx) * 64 + g * 4;
                    const float4 wa = *reinterpret_cast<const float4*>(wr);
                    const float4 wb = *reinterpret_cast<const float4*>(wr + 32);
                    const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[p][j] = fmaf(in[p + sx + 3], wv[j], acc[p][j]);
                }
            }
        }
        const int oh = h0 + qy;
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int ow = w0 + qx * 4 + p;
            if (oh < H && ow < W) {
                const size_t o = ((static_cast<size_t>(n) * H + oh) * W + ow) * 64 + g * 8;
                if (out_hi) store8p(out_hi, out_lo, planes, o, acc[p]);
                if (relu_hi) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[p][j] = fmaxf(acc[p][j], 0.f);
                    store8p(relu_hi, relu_lo, planes, o, acc[p]);
                }
            }
        }
    }
}

// large batches: 256 threads = 32 pixel blocks (2 rows x 4 pixels) x 8 channel groups; 8 px x 8 ch accumulators per thread.
// A weight vector fetched from shared memory feeds 8 pixels and an input row both output rows that touch it: 90 LDS.128
// per 1728 FMAs where the 4-pixel version issues 81 per 864.  Both are bound by instruction issue (ncu: 63 % issue
// slots busy, fma pipe 43 %): measured, the 8-pixel version wins with many tiles per block (16 x 256 x 256: fast_st step
// 8.75 -> 8.59 ms) and loses with few (one 512 x 512 image: 41.0 -> 43.6 us), so the launcher picks by tile count.
constexpr int FW_TH = 8, FW_SH = FW_TH + 2;        // forward tile: 8 x 32 pixels, staged rows
__global__ void __launch_bounds__(IMG_THREADS, 2)
conv_image_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                      bf16* __restrict__ out_hi, bf16* __restrict__ out_lo, bf16* __restrict__ relu_hi, bf16* __restrict__ relu_lo,
                      int planes, int N, int H, int W, int tiles_h, int tiles_w) {
    pdl_prologue();
    __shared__ __align__(16) float sw[27 * 64];                      // [tap = (c*3 + r)*3 + s][perm64(co)]
    __shared__ __align__(16) float st[3 * FW_SH * IM_SW];
    for (int i = threadIdx.x; i < 27 * 64; i += IMG_THREADS) {
        const int co = i & 63, tap = i >> 6;                         // w: [64][3][3][3]
        sw[tap * 64 + perm64(co)] = w[co * 27 + tap];
    }
    const int g = threadIdx.x & 7, quad = threadIdx.x >> 3;
    const int qx = quad & 7, qy = quad >> 3;                         // pixels qx*4 .. +3 of rows 2*qy, 2*qy + 1
    float b8[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) b8[j] = bias ? bias[g * 8 + j] : 0.f;
    const size_t HW = static_cast<size_t>(H) * W;
    const int tiles = N * tiles_h * tiles_w;
    // the staged image tile of the NEXT iteration is fetched into registers while this one is computed
    constexpr int ST_N = 3 * FW_SH * IM_SW, ST_PER = (ST_N + IMG_THREADS - 1) / IMG_THREADS;
    float pre[ST_PER];
    auto fetch = [&](int tile) {
        int q = tile; const int tw = q % tiles_w; q /= tiles_w; const int th = q % tiles_h; const int n = q / tiles_h;
#pragma unroll
        for (int k = 0; k < ST_PER; ++k) {
            const int i = threadIdx.x + k * IMG_THREADS;
            const int xx = i % IM_SW; int r = i / IM_SW; const int y = r % FW_SH; const int c = r / FW_SH;
            const int ih = th * FW_TH + y - 1, iw = tw * IM_TW + xx - 4;
            float v = 0.f;
            if (i < ST_N && ih >= 0 && ih < H && iw >= 0 && iw < W) v = __ldg(x + (static_cast<size_t>(n) * 3 + c) * HW + static_cast<size_t>(ih) * W + iw);
            pre[k] = v;
        }
    };
    if (blockIdx.x < tiles) fetch(blockIdx.x);
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int q = tile; const int tw = q % tiles_w; q /= tiles_w; const int th = q % tiles_h; const int n = q / tiles_h;
        const int h0 = th * FW_TH, w0 = tw * IM_TW;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < ST_PER; ++k) {
            const int i = threadIdx.x + k * IMG_THREADS;
            if (i < ST_N) st[i] = pre[k];
        }
        __syncthreads();
        if (tile + gridDim.x < tiles) fetch(tile + gridDim.x);
        float acc[2][4][8];
#pragma unroll
        for (int y = 0; y < 2; ++y)
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[y][p][j] = b8[j];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            // the four staged rows this thread's two output rows read, 12 values each (pixels qx*4 - 1 .. qx*4 + 4 at [3..8])
            float in[4][12];
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                const float* row = st + (c * FW_SH + 2 * qy + rr) * IM_SW + qx * 4;
#pragma unroll
                for (int v = 0; v < 3; ++v) {
                    const float4 t = *reinterpret_cast<const float4*>(row + v * 4);
                    in[rr][v * 4] = t.x; in[rr][v * 4 + 1] = t.y; in[rr][v * 4 + 2] = t.z; in[rr][v * 4 + 3] = t.w;
                }
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int sx = 0; sx < 3; ++sx) {
                    const float* wr = sw + ((c * 3 + r) * 3 + sx) * 64 + g * 4;
                    const float4 wa = *reinterpret_cast<const float4*>(wr);
                    const float4 wb = *reinterpret_cast<const float4*>(wr + 32);
                    const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
                    for (int y = 0; y < 2; ++y)
#pragma unroll
                        for (int p = 0; p < 4; ++p)
#pragma unroll
                            for (int j = 0; j < 8; ++j) acc[y][p][j] = fmaf(in[y + r][p + sx + 3], wv[j], acc[y][p][j]);
                }
            }
        }
#pragma unroll
        for (int y = 0; y < 2; ++y) {
            const int oh = h0 + 2 * qy + y;
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const int ow = w0 + qx * 4 + p;
                if (oh < H && ow < W) {
                    const size_t o = ((static_cast<size_t>(n) * H + oh) * W + ow) * 64 + g * 8;
                    if (out_hi) store8p(out_hi, out_lo, planes, o, acc[y][p]);
                    if (relu_hi) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[y][p][j] = fmaxf(acc[y][p][j], 0.f);
                        store8p(relu_hi, relu_lo, planes, o, acc[y][p]);
                    }
                }
            }
        }
    }
}

// gx[n][c][h][w] = sum_{co,r,s} g[n][h+1-r][w+1-s][co] * w[co][c][r][s]
// the gradient tile (+1 halo) is staged as fp32 [row][pixel][perm64(co)]; 256 threads = 32 strips of
// 4 pixels x 8 channel groups; each thread: 4 px x 3 outputs over its 8 channels, then a shuffle
// reduction over the 8 channel groups.
__global__ void __launch_bounds__(IMG_THREADS)
conv_image_dgrad_kernel(const bf16* __restrict__ g_hi, const bf16* __restrict__ g_lo, int g_planes,
                        const float* __restrict__ w, float* __restrict__ gx, int N, int H, int W, int tiles_h, int tiles_w) {
    pdl_prologue();
    extern __shared__ __align__(16) float smd[];
    float* sw = smd;                                   // [tap = r*3 + s][c][perm64(co)]
    float* st = smd + 27 * 64;                         // [IM_SH][IM_GW][64]
    for (int i = threadIdx.x; i < 27 * 64; i += IMG_THREADS) {
        const int co = i & 63; int q = i >> 6; const int c = q % 3; const int tap = q / 3;
        sw[(tap * 3 + c) * 64 + perm64(co)] = w[(co * 3 + c) * 9 + tap];
    }
    const int cg = threadIdx.x & 7, strip = threadIdx.x >> 3;
    const int sx = strip & 7, sy = strip >> 3;
    const size_t HW = static_cast<size_t>(H) * W;
    const int tiles = N * tiles_h * tiles_w;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int q = tile; const int tw = q % tiles_w; q /= tiles_w; const int th = q % tiles_h; const int n = q / tiles_h;
        const int h0 = th * IM_TH, w0 = tw * IM_TW;
        __syncthreads();
        for (int i = threadIdx.x; i < IM_SH * IM_GW * 8; i += IMG_THREADS) {
            const int v8 = i & 7; int r = i >> 3; const int xx = r % IM_GW; const int y = r / IM_GW;
            const int ih = h0 + y - 1, iw = w0 + xx - 1;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.f;
            if (ih >= 0 && ih < H && iw >= 0 && iw < W)
                load8p(g_hi, g_lo, g_planes, ((static_cast<size_t>(n) * H + ih) * W + iw) * 64 + v8 * 8, v);
            float* dst = st + (y * IM_GW + xx) * 64 + v8 * 4;
            *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(dst + 32) = make_float4(v[4], v[5], v[6], v[7]);
        }
        __syncthreads();
        float acc[4][3];
#pragma unroll
        for (int p = 0; p < 4; ++p) { acc[p][0] = 0.f; acc[p][1] = 0.f; acc[p][2] = 0.f; }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            // output row sy reads gradient row sy + 1 - r (+1 halo) ; output col p reads col p + 1 - s (+1 halo)
            const float* row = st + ((sy + 2 - r) * IM_GW + sx * 4) * 64 + cg * 4;
            float in[6][8];
#pragma unroll
            for (int xx = 0; xx < 6; ++xx) {
                const float4 a = *reinterpret_cast<const float4*>(row + xx * 64);
                const float4 b = *reinterpret_cast<const float4*>(row + xx * 64 + 32);
                in[xx][0] = a.x; in[xx][1] = a.y; in[xx][2] = a.z; in[xx][3] = a.w;
                in[xx][4] = b.x; in[xx][5] = b.y; in[xx][6] = b.z; in[xx][7] = b.w;
            }
#pragma unroll
            for (int s = 0; s < 3; ++s) {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float* wr = sw + ((r * 3 + s) * 3 + c) * 64 + cg * 4;
                    const float4 wa = *reinterpret_cast<const float4*>(wr);
                    const float4 wb = *reinterpret_cast<const float4*>(wr + 32);
                    const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int k = 0; k < 8; ++k) acc[p][c] = fmaf(in[p + 2 - s][k], wv[k], acc[p][c]);
                }
            }
        }
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float v = acc[p][c];
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                acc[p][c] = v;
            }
        // lane cg < 4 writes pixel p = cg of the strip, lanes 4..6 idle; 3 channels each
        const int oh = h0 + sy, ow = w0 + sx * 4 + (cg & 3);
        if (cg < 4 && oh < H && ow < W) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float v = 0.f;
#pragma unroll
                for (int p = 0; p < 4; ++p) v = (p == cg) ? acc[p][c] : v;
                gx[(static_cast<size_t>(n) * 3 + c) * HW + static_cast<size_t>(oh) * W + ow] = v;
            }
        }
    }
}

inline int img_grid(int tiles) {
    const int cap = st_num_sms() * 4;
    return tiles < cap ? tiles : cap;
}

}  // namespace

int conv_image_fwd(const float* x, const float* w, const float* bias, bf16* out_hi, bf16* out_lo,
                   bf16* relu_hi, bf16* relu_lo, int planes, int N, int H, int W, int Cout, cudaStream_t st) {
    ST_CHECK_ARG(Cout == 64, "conv_image_fwd: Cout must be 64 (got %d)", Cout);
    const int th8 = (H + FW_TH - 1) / FW_TH, tw = (W + IM_TW - 1) / IM_TW;
    const int tiles8 = N * th8 * tw, cap = 2 * st_num_sms();          // two resident blocks per SM (register budget)
    if (tiles8 >= 4 * cap) {
        st_launch(cap, IMG_THREADS, 0, st)(conv_image_fwd_kernel, x, w, bias, out_hi, out_lo, relu_hi, relu_lo, planes, N, H, W, th8, tw);
    } else {
        const int th = (H + IM_TH - 1) / IM_TH;
        st_launch(img_grid(N * th * tw), IMG_THREADS, 0, st)(conv_image_fwd4_kernel, x, w, bias, out_hi, out_lo, relu_hi, relu_lo, planes, N, H, W, th, tw);
    }
    ST_CHECK_LAUNCH("conv_image_fwd_kernel");
    return ST_OK;
}

int conv_image_dgrad(const bf16* g_hi, const bf16* g_lo, int g_planes, const float* w, float* gx,
                     int N, int H, int W, int Cout, cudaStream_t st) {
    ST_CHECK_ARG(Cout == 64, "conv_image_dgrad: Cout must be 64 (got %d)", Cout);
    const int th = (H + IM_TH - 1) / IM_TH, tw = (W + IM_TW - 1) / IM_TW;
    const int smem = (27 * 64 + IM_SH * IM_GW * 64) * sizeof(float);
    static bool set = false;
    if (!set) { cudaFuncSetAttribute(conv_image_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); set = true; }
    st_launch(img_grid(N * th * tw), IMG_THREADS, smem, st)(conv_image_dgrad_kernel, g_hi, g_lo, g_planes, w, gx, N, H, W, th, tw);
    ST_CHECK_LAUNCH("conv_image_dgrad_kernel");
    return ST_OK;
}
