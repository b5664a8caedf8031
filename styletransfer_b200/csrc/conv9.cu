// The transform net's two 9x9 convolutions (network.py:525-530 / :850-855: 3 or 6 -> 32 channels;
// :604-609: 32 -> 3 channels), forward, backward-data and backward-weight.
//
// One side of these layers has 3 (6) channels: as implicit GEMMs they have N = 3 or K-chunks of
// 3, i.e. tensor-core tiles that are > 80 % padding, and the mma.sync gather kernel spent 12 ms
// of a 29 ms training step on them.  They are written here as register-tiled direct convolutions
// on the CUDA cores in exact fp32 (the choice SURVEY.md K5 leaves to the builder):
//   expand    S -> B : out[p][32] = sum_{c,r,s} S[c][p + (r,s) - 4] * Wt[(c,r,s)][32]
//                      (layer-0 forward; layer-22 backward-data with flipped taps)
//   reduce    B -> S : out[co<3][p] = sum_{r,s,ci} B[p + (r,s) - 4][ci] * W[co][ci][r][s]
//                      (layer-22 forward, writes the NCHW fp32 network output)
//   correlate B,S -> dW : O[b][c][r][s] = sum_p B[p][b] * S[c][p + (r,s) - 4]
//                      (backward-weight of both layers)
// "S" tensors (3 / 6 channels) live at the module boundary: NCHW fp32 (image batch, network
// output, its gradient).  "B" tensors (32 channels) are split-bf16 NHWC like every activation.
#include "common.cuh"

namespace {

constexpr int K9 = 9, PAD9 = 4, CB = 32;          // kernel size, padding, channels of the big side

// ---------------------------------------------------------------------------------------------
// expand: tile of 8 x 32 output pixels per iteration, 256 threads = 64 pixel quads x 4 channel
// groups; each thread: 4 pixels (along w) x 8 channels = 32 accumulators.
// ---------------------------------------------------------------------------------------------
constexpr int EX_TH = 8, EX_TW = 32, EX_THREADS = 256;
constexpr int EX_SW = EX_TW + 2 * PAD9;            // staged tile width (40 floats: rows stay 16-byte aligned)
constexpr int EX_SH = EX_TH + 2 * PAD9;

template <int CS>
__global__ void __launch_bounds__(EX_THREADS)
conv9_expand_kernel(const float* __restrict__ s0, const float* __restrict__ s1, int sc, const float* __restrict__ w,
                    const float* __restrict__ bias, int flip,
                    bf16* __restrict__ out_hi, bf16* __restrict__ out_lo, int planes,
                    int N, int H, int W, int tiles_h, int tiles_w) {
    pdl_prologue();
    extern __shared__ float sm[];
    float* sw = sm;                                  // [CS*81][32]
    float* st = sm + CS * 81 * CB;                   // [CS][EX_SH][EX_SW]
    // weights: forward  w [32][CS][9][9]  -> sw[(c,r,s)][co] = w[co][c][r][s]
    //          flipped  w [CS][32][9][9]  -> sw[(c,r,s)][ci] = w[c][ci][8-r][8-s]   (backward-data)
    for (int i = threadIdx.x; i < CS * 81 * CB; i += EX_THREADS) {
        const int ch = i % CB, t = i / CB;
        const int c = t / 81, rs = t % 81;
        sw[i] = flip ? w[(c * CB + ch) * 81 + (80 - rs)] : w[(ch * CS + c) * 81 + rs];
    }
    const int g = threadIdx.x & 3, quad = threadIdx.x >> 2;      // channel group, pixel quad
    const int qy = quad / (EX_TW / 4), qx = quad % (EX_TW / 4);
    const size_t HW = static_cast<size_t>(H) * W;
    const int tiles = N * tiles_h * tiles_w;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int q = tile; const int tw = q % tiles_w; q /= tiles_w; const int th = q % tiles_h; const int n = q / tiles_h;
        const int h0 = th * EX_TH, w0 = tw * EX_TW;
        __syncthreads();
        for (int i = threadIdx.x; i < CS * EX_SH * EX_SW; i += EX_THREADS) {
            const int x = i % EX_SW; int r = i / EX_SW; const int y = r % EX_SH; const int c = r / EX_SH;
            const int ih = h0 + y - PAD9, iw = w0 + x - PAD9;
            float v = 0.f;
            if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
                const float* src = c < 3 ? s0 : s1;
                v = __ldg(src + (static_cast<size_t>(n) * sc + (c < 3 ? c : c - 3)) * HW + static_cast<size_t>(ih) * W + iw);
            }
            st[i] = v;
        }
        __syncthreads();
        float acc[4][8];
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[p][j] = bias ? bias[g * 8 + j] : 0.f;
#pragma unroll 1
        for (int c = 0; c < CS; ++c) {
#pragma unroll 1
            for (int r = 0; r < K9; ++r) {
                const float* row = st + (c * EX_SH + qy + r) * EX_SW + qx * 4;
                float in[12];
#pragma unroll
                for (int v = 0; v < 3; ++v) {
                    const float4 t = *reinterpret_cast<const float4*>(row + v * 4);
                    in[v * 4] = t.x; in[v * 4 + 1] = t.y; in[v * 4 + 2] = t.z; in[v * 4 + 3] = t.w;
                }
                const float* wr = sw + ((c * K9 + r) * K9) * CB + g * 8;
#pragma unroll
                for (int s = 0; s < K9; ++s) {
                    const float4 wa = *reinterpret_cast<const float4*>(wr + s * CB);
                    const float4 wb = *reinterpret_cast<const float4*>(wr + s * CB + 4);
                    const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int j = 0; j < 8; ++j) acc[p][j] = fmaf(in[p + s], wv[j], acc[p][j]);
                }
            }
        }
        const int oh = h0 + qy;
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int ow = w0 + qx * 4 + p;
            if (oh < H && ow < W)
                store8p(out_hi, out_lo, planes, ((static_cast<size_t>(n) * H + oh) * W + ow) * CB + g * 8, acc[p]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// reduce: tile of 8 x 32 output pixels, 256 threads = 32 strips of 8 pixels x 8 groups of 4 input
// channels; each thread: 8 pixels x 3 outputs over its 4 channels, then a shuffle reduction over
// the 8 channel groups.  The input tile is staged as fp32 [row][pixel][32].
// ---------------------------------------------------------------------------------------------
constexpr int RD_TH = 8, RD_TW = 32, RD_THREADS = 256;
constexpr int RD_SW = RD_TW + 2 * PAD9, RD_SH = RD_TH + 2 * PAD9;

__global__ void __launch_bounds__(RD_THREADS)
conv9_reduce_kernel(const bf16* __restrict__ x_hi, const bf16* __restrict__ x_lo, int planes,
                    const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ out,
                    int N, int H, int W, int tiles_h, int tiles_w) {
    pdl_prologue();
    extern __shared__ float sm[];
    float* sw = sm;                                  // [81][32][4]  (co padded 3 -> 4)
    float* st = sm + 81 * CB * 4;                    // [RD_SH][RD_SW][32]
    for (int i = threadIdx.x; i < 81 * CB * 4; i += RD_THREADS) {
        const int co = i & 3, ci = (i >> 2) % CB, rs = i / (4 * CB);
        sw[i] = co < 3 ? w[(co * CB + ci) * 81 + rs] : 0.f;      // w [3][32][9][9]
    }
    const int cg = threadIdx.x & 7, strip = threadIdx.x >> 3;    // channel group (4 ch), strip of 8 pixels
    const int sy = strip / (RD_TW / 8), sx = strip % (RD_TW / 8);
    const size_t HW = static_cast<size_t>(H) * W;
    const int tiles = N * tiles_h * tiles_w;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int q = tile; const int tw = q % tiles_w; q /= tiles_w; const int th = q % tiles_h; const int n = q / tiles_h;
        const int h0 = th * RD_TH, w0 = tw * RD_TW;
        __syncthreads();
        for (int i = threadIdx.x; i < RD_SH * RD_SW * (CB / 8); i += RD_THREADS) {
            const int v8 = i % (CB / 8); int r = i / (CB / 8); const int x = r % RD_SW; const int y = r / RD_SW;
            const int ih = h0 + y - PAD9, iw = w0 + x - PAD9;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.f;
            if (ih >= 0 && ih < H && iw >= 0 && iw < W)
                load8p(x_hi, x_lo, planes, ((static_cast<size_t>(n) * H + ih) * W + iw) * CB + v8 * 8, v);
            float4* dst = reinterpret_cast<float4*>(st + (y * RD_SW + x) * CB + v8 * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
        __syncthreads();
        float acc[8][3];
#pragma unroll
        for (int p = 0; p < 8; ++p) { acc[p][0] = 0.f; acc[p][1] = 0.f; acc[p][2] = 0.f; }
#pragma unroll 1
        for (int r = 0; r < K9; ++r) {
            float4 in[16];
            const float* row = st + ((sy + r) * RD_SW + sx * 8) * CB + cg * 4;
#pragma unroll
            for (int x = 0; x < 16; ++x) in[x] = *reinterpret_cast<const float4*>(row + x * CB);
#pragma unroll
            for (int s = 0; s < K9; ++s) {
                const float4* wp = reinterpret_cast<const float4*>(sw + ((r * K9 + s) * CB + cg * 4) * 4);
                const float4 w0v = wp[0], w1v = wp[1], w2v = wp[2], w3v = wp[3];     // per input channel: (co0, co1, co2, 0)
#pragma unroll
                for (int p = 0; p < 8; ++p) {
                    const float4 v = in[p + s];
                    acc[p][0] = fmaf(v.x, w0v.x, fmaf(v.y, w1v.x, fmaf(v.z, w2v.x, fmaf(v.w, w3v.x, acc[p][0]))));
                    acc[p][1] = fmaf(v.x, w0v.y, fmaf(v.y, w1v.y, fmaf(v.z, w2v.y, fmaf(v.w, w3v.y, acc[p][1]))));
                    acc[p][2] = fmaf(v.x, w0v.z, fmaf(v.y, w1v.z, fmaf(v.z, w2v.z, fmaf(v.w, w3v.z, acc[p][2]))));
                }
            }
        }
        // sum over the 8 channel groups (lanes cg = 0..7 of each strip are adjacent lanes)
#pragma unroll
        for (int p = 0; p < 8; ++p)
#pragma unroll
            for (int co = 0; co < 3; ++co) {
                float v = acc[p][co];
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                acc[p][co] = v;
            }
        // lane cg writes pixel p = cg of the strip (3 channels)
        const int oh = h0 + sy, ow = w0 + sx * 8 + cg;
        if (oh < H && ow < W) {
#pragma unroll
            for (int co = 0; co < 3; ++co) {
                float v = 0.f;
#pragma unroll
                for (int p = 0; p < 8; ++p) v = (p == cg) ? acc[p][co] : v;
                out[(static_cast<size_t>(n) * 3 + co) * HW + static_cast<size_t>(oh) * W + ow] = v + (bias ? bias[co] : 0.f);
            }
        }
    }
}

}  // namespace

#include "conv9_wgrad.cuh"

// ---------------------------------------------------------------------------------------------
// Layer-22 forward (32 -> 3, 9x9) on the tensor cores, in two stages:
//   (1) T[p][co*9 + s] = sum_{r, ci} X[p + (r - 4) rows][ci] * W[co][ci][r][s]   -- a VERTICAL 9-tap conv
//       32 -> 27 (padded 32) channels: K = 288 per output instead of K = 2592 for 3 outputs, so the
//       activation tile is read 9 times from shared memory, not 81 (conv_tcx mode 3, fp32 output);
//   (2) out[co][p] = bias[co] + sum_s T[p + (s - 4) columns][co*9 + s]            -- horizontal gather-sum.
// ---------------------------------------------------------------------------------------------
namespace {

// weights [3][32][9][9] fp32 -> split-bf16 [planes][row = co*9 + s (32, 27 used)][tap = r (9)][ci (32)]
// (plane code with the fp16 bit: fp16 planes of ST_W_F16_SCALE * w)
__global__ void conv9_pack_vertical_kernel(const float* __restrict__ w, bf16* __restrict__ hi, int code) {
    pdl_prologue();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = 32 * 9 * CB;
    if (idx >= total) return;
    const int ci = idx % CB; int q = idx / CB; const int r = q % 9; const int row = q / 9;
    float v = 0.f;
    if (row < 27) { const int co = row / 9, s = row % 9; v = w[((co * CB + ci) * 9 + r) * 9 + s]; }
    const int planes = pc_planes(code);
    if (pc_f16(code)) {
        uint32_t h2, l2; split2h(v * ST_W_F16_SCALE, 0.f, h2, l2);
        reinterpret_cast<uint16_t*>(hi)[idx] = static_cast<uint16_t>(h2 & 0xffffu);
        if (planes > 1) reinterpret_cast<uint16_t*>(hi)[static_cast<size_t>(total) + idx] = static_cast<uint16_t>(l2 & 0xffffu);
        return;
    }
#pragma unroll 1
    for (int pl = 0; pl < planes; ++pl) {
        const bf16 h = __float2bfloat16_rn(v);
        hi[static_cast<size_t>(pl) * total + idx] = h;
        v -= __bfloat162float(h);
    }
}

// one thread per output pixel; the T rows of 256 consecutive pixels (+4 on each side) are staged in
// shared memory with coalesced 128-bit loads (row stride 33 words: conflict-free gathers)
constexpr int GA_PIX = 256, GA_ROWS = GA_PIX + 2 * PAD9, GA_LD = 33;
__global__ void __launch_bounds__(GA_PIX)
conv9_gather_kernel(const float* __restrict__ T, const float* __restrict__ bias, float* __restrict__ out, int N, int H, int W) {
    pdl_prologue();
    __shared__ float sT[GA_ROWS * GA_LD];
    const size_t HW = static_cast<size_t>(H) * W;
    const long long total = static_cast<long long>(N) * HW;
    const float b0 = bias ? bias[0] : 0.f, b1 = bias ? bias[1] : 0.f, b2 = bias ? bias[2] : 0.f;
    for (long long base = static_cast<long long>(blockIdx.x) * GA_PIX; base < total; base += static_cast<long long>(gridDim.x) * GA_PIX) {
        __syncthreads();
        for (int i = threadIdx.x; i < GA_ROWS * 8; i += GA_PIX) {
            const int row = i >> 3, q4 = i & 7;
            const long long gp = base - PAD9 + row;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gp >= 0 && gp < total) v = __ldg(reinterpret_cast<const float4*>(T + gp * CB) + q4);
            float* d = sT + row * GA_LD + q4 * 4;
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        }
        __syncthreads();
        const long long pix = base + threadIdx.x;
        if (pix < total) {
            const size_t n = pix / HW; const size_t rem = pix - n * HW;
            const int w = static_cast<int>(rem % W);
            float a0 = b0, a1 = b1, a2 = b2;
#pragma unroll
            for (int s = 0; s < 9; ++s) {
                const int iw = w + s - PAD9;
                if (iw < 0 || iw >= W) continue;
                const float* t = sT + (threadIdx.x + s) * GA_LD;
                a0 += t[s]; a1 += t[9 + s]; a2 += t[18 + s];
            }
            out[(n * 3 + 0) * HW + rem] = a0;
            out[(n * 3 + 1) * HW + rem] = a1;
            out[(n * 3 + 2) * HW + rem] = a2;
        }
    }
}

}  // namespace

int conv_tcx_run(int mode, const bf16*, const bf16*, int, int, int, int, int, const bf16*, const bf16*, int, int, int, int, int,
                 const float*, const bf16*, const bf16*, const bf16*, int, bf16*, bf16*, bf16*, bf16*, int, float*,
                 float, const float*, const float*, cudaStream_t);

size_t conv9_reduce_workspace_bytes(int N, int H, int W) {
    return static_cast<size_t>(N) * H * W * CB * sizeof(float) + 3 * 32 * 9 * CB * sizeof(bf16) + 1024;
}

// workspace: [T fp32 N*H*W*32][packed weights 3 planes]
int conv9_reduce_tc_run(const bf16* x_hi, const bf16* x_lo, int planes, const float* w, const float* bias, float* out,
                        void* workspace, size_t workspace_bytes, int N, int H, int W, cudaStream_t st) {
    ST_CHECK_ARG(workspace && workspace_bytes >= conv9_reduce_workspace_bytes(N, H, W), "conv9_reduce: workspace too small");
    float* T = static_cast<float*>(workspace);
    const size_t t_bytes = (static_cast<size_t>(N) * H * W * CB * sizeof(float) + 1023) / 1024 * 1024;
    bf16* wp = reinterpret_cast<bf16*>(static_cast<uint8_t*>(workspace) + t_bytes);
    // the weights are packed in the format of the activation planes (a tensor-core product takes one format)
    const bool f16 = pc_f16(planes);
    const int np = pc_planes(planes);
    const int wplanes = np >= 3 ? 3 : (np == 2 ? 2 : 1);
    const int wcode = wplanes | (f16 ? ST_FMT_F16 : 0);
    const int total = 32 * 9 * CB;
    st_launch((total + 255) / 256, 256, 0, st)(conv9_pack_vertical_kernel, w, wp, wcode);
    ST_CHECK_LAUNCH("conv9_pack_vertical_kernel");
    int rc = conv_tcx_run(3, x_hi, x_lo, planes, N, H, W, CB, wp, wplanes > 1 ? wp + total : nullptr, wcode, 32, 32, H, W,
                          nullptr, nullptr, nullptr, nullptr, 1, nullptr, nullptr, nullptr, nullptr, 1, T,
                          f16 ? ST_W_F16_INV_SCALE : 1.f, nullptr, nullptr, st);
    if (rc) return rc;
    size_t blocks = (static_cast<size_t>(N) * H * W + GA_PIX - 1) / GA_PIX;
    const size_t cap = static_cast<size_t>(st_num_sms()) * 6;
    if (blocks > cap) blocks = cap;
    st_launch(static_cast<unsigned>(blocks), GA_PIX, 0, st)(conv9_gather_kernel, T, bias, out, N, H, W);
    ST_CHECK_LAUNCH("conv9_gather_kernel");
    return ST_OK;
}
