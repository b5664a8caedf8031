// Part of conv9.cu: backward-weight of the 9x9 layers ("correlate") and the host launchers.
//
//   O[b][c][r][s] = sum over pixels p of  B[p][b] * S[c][p + (r - 4, s - 4)]
//
// Thread = (group of 8 big channels, vertical tap r, small channel c): it walks along an image
// row keeping the 9 horizontal neighbours S[w-4 .. w+4] in a register window (one new value per
// pixel) and 8 x 9 = 72 accumulators; per pixel 2 LDS.128 (the 8 B values, a broadcast) + 1 LDS
// feed 72 FMAs.  A CTA handles ROWS image rows at a time (216 threads = ROWS x 4 x 9 x CS),
// 126 columns per staging step, and writes its partial sums once at the end; a finish kernel adds
// the partials in a fixed order (deterministic) and maps them to the PyTorch weight layout.
#pragma once

namespace {

constexpr int CR_WC = 126, CR_SWS = 135, CR_THREADS = 224, CR_ACTIVE = 216;

template <int CS>
__global__ void __launch_bounds__(CR_THREADS)
conv9_correlate_kernel(const bf16* __restrict__ b_hi, const bf16* __restrict__ b_lo, int planes,
                       const float* __restrict__ s0, const float* __restrict__ s1, int sc,
                       float* __restrict__ partial, int N, int H, int W) {
    pdl_prologue();
    constexpr int ROWS = 6 / CS;                     // 2 rows at a time for 3 channels, 1 for 6
    constexpr int SROWS = ROWS + 2 * PAD9;
    extern __shared__ float sm[];
    float* Bs = sm;                                  // [ROWS][CR_WC][32]
    float* Ss = sm + ROWS * CR_WC * CB;              // [CS][SROWS][CR_SWS]
    const int t = threadIdx.x;
    const bool active = t < CR_ACTIVE;
    const int per_row = 36 * CS;
    const int row_i = active ? t / per_row : 0;
    const int u = active ? t % per_row : 0;
    const int bg = u & 3, rc = u >> 2, r = rc % K9, c = rc / K9;
    const size_t HW = static_cast<size_t>(H) * W;
    float acc[8][K9];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int s = 0; s < K9; ++s) acc[j][s] = 0.f;

    const int row_groups = (H + ROWS - 1) / ROWS;
    const int chunks = (W + CR_WC - 1) / CR_WC;
    const int work = N * row_groups * chunks;
    for (int wk = blockIdx.x; wk < work; wk += gridDim.x) {
        int q = wk; const int ck = q % chunks; q /= chunks; const int hg = q % row_groups; const int n = q / row_groups;
        const int h0 = hg * ROWS, wc0 = ck * CR_WC;
        __syncthreads();
        // stage the big tensor rows as fp32
        for (int i = t; i < ROWS * CR_WC * (CB / 8); i += CR_THREADS) {
            const int v8 = i % (CB / 8); int rr = i / (CB / 8); const int x = rr % CR_WC; const int y = rr / CR_WC;
            const int ih = h0 + y, iw = wc0 + x;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.f;
            if (ih < H && iw < W) load8p(b_hi, b_lo, planes, ((static_cast<size_t>(n) * H + ih) * W + iw) * CB + v8 * 8, v);
            float4* dst = reinterpret_cast<float4*>(Bs + (y * CR_WC + x) * CB + v8 * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
        // stage the small tensor rows h0-4 .. h0+ROWS+3, columns wc0-4 .. wc0+CR_WC+3 (zero outside)
        for (int i = t; i < CS * SROWS * (CR_WC + 2 * PAD9); i += CR_THREADS) {
            const int x = i % (CR_WC + 2 * PAD9); int rr = i / (CR_WC + 2 * PAD9); const int y = rr % SROWS; const int cc = rr / SROWS;
            const int ih = h0 + y - PAD9, iw = wc0 + x - PAD9;
            float v = 0.f;
            if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
                const float* src = cc < 3 ? s0 : s1;
                v = __ldg(src + (static_cast<size_t>(n) * sc + (cc < 3 ? cc : cc - 3)) * HW + static_cast<size_t>(ih) * W + iw);
            }
            Ss[(cc * SROWS + y) * CR_SWS + x] = v;
        }
        __syncthreads();
        if (active) {
            const float* srow = Ss + (c * SROWS + row_i + r) * CR_SWS;      // tap r of output row h0 + row_i
            const float* brow = Bs + row_i * CR_WC * CB + bg * 8;
            float win[K9];
#pragma unroll
            for (int s = 0; s < K9 - 1; ++s) win[s] = srow[s];
#pragma unroll 1
            for (int base = 0; base < CR_WC; base += K9) {
#pragma unroll
                for (int k = 0; k < K9; ++k) {
                    const int x = base + k;
                    win[(k + K9 - 1) % K9] = srow[x + K9 - 1];              // the window's newest element
                    const float4 ba = *reinterpret_cast<const float4*>(brow + x * CB);
                    const float4 bb = *reinterpret_cast<const float4*>(brow + x * CB + 4);
                    const float bv[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
#pragma unroll
                    for (int j = 0; j < 8; ++j)
#pragma unroll
                        for (int s = 0; s < K9; ++s) acc[j][s] = fmaf(bv[j], win[(k + s) % K9], acc[j][s]);
                }
            }
        }
    }
    if (active) {
        float* out = partial + (static_cast<size_t>(blockIdx.x) * ROWS + row_i) * (CB * CS * 81);
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int s = 0; s < K9; ++s)
                out[(((bg * 8 + j) * CS + c) * K9 + r) * K9 + s] = acc[j][s];
    }
}

// mode 0: dw[b][c][r][s] = O[b][c][r][s]                 (layer 0:  B = dy, S = x)
// mode 1: dw[c][b][r][s] = O[b][c][8-r][8-s]             (layer 22: B = x,  S = dy)
// mul_a / mul_b (optional scale slots of the operands): the sum is multiplied by slot[1] of each
__global__ void conv9_wgrad_finish_kernel(const float* __restrict__ partial, int nparts, int CS, int mode,
                                          const float* __restrict__ mul_a, const float* __restrict__ mul_b, float* __restrict__ dw) {
    pdl_prologue();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = CB * CS * 81;
    if (idx >= total) return;
    int src = idx;
    if (mode == 1) {
        const int rs = idx % 81; int q = idx / 81; const int b = q % CB; const int c = q / CB;
        src = (b * CS + c) * 81 + (80 - rs);
    }
    float s = 0.f;
    for (int p = 0; p < nparts; ++p) s += partial[static_cast<size_t>(p) * total + src];
    dw[idx] = s * slot_inv(mul_a) * slot_inv(mul_b);
}

// per-channel sum of an NCHW fp32 tensor with 3 channels (bias gradient of layer 22), two
// deterministic stages: grid (CS_PARTS, 3) partial sums, then one block adds them in order
constexpr int CS_PARTS = 128;
__global__ void __launch_bounds__(256)
nchw3_channel_sum_kernel(const float* __restrict__ g, int sc, int N, size_t HW, float* __restrict__ part) {
    pdl_prologue();
    __shared__ float smr[32];
    const int c = blockIdx.y;
    float s = 0.f;
    const size_t total = static_cast<size_t>(N) * HW;
    for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < total; i += static_cast<size_t>(CS_PARTS) * 256) {
        const size_t n = i / HW;
        s += __ldg(g + (n * sc + c) * HW + (i - n * HW));
    }
    s = block_sum(s, smr);
    if (threadIdx.x == 0) part[c * CS_PARTS + blockIdx.x] = s;
}
__global__ void __launch_bounds__(96)
nchw3_channel_sum_finish_kernel(const float* __restrict__ part, float* __restrict__ out) {
    pdl_prologue();
    const int c = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float s = 0.f;
    for (int i = lane; i < CS_PARTS; i += 32) s += part[c * CS_PARTS + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[c] = s;
}

inline int conv9_grid() { return st_num_sms() * 2; }

}  // namespace

int ew_channel_sum(const bf16* xh, const bf16* xl, int code, const float* slot, int P, int C, int Cout, float* out, float* ws, cudaStream_t st);

int conv9_expand_tc_run(const float* s0, const float* s1, int sc, int CS, const float* w, const float* bias, int flip,
                        bf16* out_hi, bf16* out_lo, int planes, int N, int H, int W, cudaStream_t stream);

int conv9_expand_run(const float* s0, const float* s1, int sc, int CS, const float* w, const float* bias, int flip,
                     bf16* out_hi, bf16* out_lo, int planes, int N, int H, int W, int use_tc, cudaStream_t st) {
    // tensor-core path (conv9_tc.cu) for 2 / 3 output planes; st_set_conv_path(1) selects the CUDA-core kernel
    if (use_tc && pc_planes(planes) >= 2 && out_lo) return conv9_expand_tc_run(s0, s1, sc, CS, w, bias, flip, out_hi, out_lo, planes, N, H, W, st);
    const int th = (H + EX_TH - 1) / EX_TH, tw = (W + EX_TW - 1) / EX_TW;
    const int tiles = N * th * tw;
    const int grid = tiles < conv9_grid() ? tiles : conv9_grid();
    const size_t smem = (static_cast<size_t>(CS) * 81 * CB + CS * EX_SH * EX_SW) * sizeof(float);
    if (CS == 3) {
        static bool set3 = false;
        if (!set3) { cudaFuncSetAttribute(conv9_expand_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); set3 = true; }
        st_launch(grid, EX_THREADS, smem, st)(conv9_expand_kernel<3>, s0, s1, sc, w, bias, flip, out_hi, out_lo, planes, N, H, W, th, tw);
    } else {
        static bool set6 = false;
        if (!set6) { cudaFuncSetAttribute(conv9_expand_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); set6 = true; }
        st_launch(grid, EX_THREADS, smem, st)(conv9_expand_kernel<6>, s0, s1, sc, w, bias, flip, out_hi, out_lo, planes, N, H, W, th, tw);
    }
    ST_CHECK_LAUNCH("conv9_expand_kernel");
    return ST_OK;
}

int conv9_reduce_run(const bf16* x_hi, const bf16* x_lo, int planes, const float* w, const float* bias, float* out,
                     int N, int H, int W, cudaStream_t st) {
    const int th = (H + RD_TH - 1) / RD_TH, tw = (W + RD_TW - 1) / RD_TW;
    const int tiles = N * th * tw;
    const int grid = tiles < conv9_grid() ? tiles : conv9_grid();
    const size_t smem = (static_cast<size_t>(81) * CB * 4 + RD_SH * RD_SW * CB) * sizeof(float);
    static bool set = false;
    if (!set) { cudaFuncSetAttribute(conv9_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); set = true; }
    st_launch(grid, RD_THREADS, smem, st)(conv9_reduce_kernel, x_hi, x_lo, planes, w, bias, out, N, H, W, th, tw);
    ST_CHECK_LAUNCH("conv9_reduce_kernel");
    return ST_OK;
}

size_t conv9_wgrad_workspace_bytes(int CS) {
    const size_t parts = static_cast<size_t>(conv9_grid()) * (6 / CS);
    const size_t a = parts * CB * CS * 81 * sizeof(float), b = static_cast<size_t>(1024) * CB * sizeof(float);
    return a > b ? a : b;
}

// mode 0: layer 0 (B = dy [N,H,W,32], S = input image(s)); mode 1: layer 22 (B = x, S = dy NCHW fp32)
// f16: the big tensor holds fp16 planes (the builder then converts the small tensor to fp16 too, multiplied by
// small_slot[0] when given); the partial sums carry the scales of both operands
int conv9_correlate_tc_run(const bf16* b_hi, bool f16, const float* small_slot, const float* s, int sc, int cs_total, int c_off,
                           float* partial, int N, int H, int W, int* nparts, cudaStream_t stream);

// use_tc: tensor-core correlation (conv9_tc.cu) when the two planes of the big tensor are one plane
// stride apart; otherwise (or with st_set_conv_path(1)) the CUDA-core kernel below.
// planes: plane code of the big tensor; big_slot / small_slot (optional): power-of-two scale slots -- the big
// tensor is STORED multiplied by big_slot[0]; the small (fp32) tensor is multiplied by small_slot[0] when it is
// converted to fp16 operands (tensor-core path with fp16 planes only)
int conv9_wgrad_run(const bf16* b_hi, const bf16* b_lo, int planes, const float* big_slot, const float* small_slot,
                    const float* s0, const float* s1, int sc, int CS, int mode,
                    float* dw, float* db, float* ws, size_t ws_bytes, int N, int H, int W, int use_tc, cudaStream_t st) {
    ST_CHECK_ARG(ws && ws_bytes >= conv9_wgrad_workspace_bytes(CS), "conv9_wgrad: workspace too small");
    const bool f16 = pc_f16(planes);
    if (use_tc && pc_planes(planes) >= 2 && b_lo == b_hi + static_cast<size_t>(N) * H * W * CB) {
        int nparts = 0, rc;
        const float* sslot = f16 ? small_slot : nullptr;
        if ((rc = conv9_correlate_tc_run(b_hi, f16, sslot, s0, sc, CS, 0, ws, N, H, W, &nparts, st))) return rc;
        if (CS == 6 && (rc = conv9_correlate_tc_run(b_hi, f16, sslot, s1, sc, CS, 3, ws, N, H, W, &nparts, st))) return rc;
        const int total = CB * CS * 81;
        st_launch((total + 255) / 256, 256, 0, st)(conv9_wgrad_finish_kernel, ws, nparts, CS, mode, big_slot, sslot, dw);
        ST_CHECK_LAUNCH("conv9_wgrad_finish_kernel");
        if (db) {
            if (mode == 0) return ew_channel_sum(b_hi, b_lo, planes, big_slot, N * H * W, CB, CB, db, ws, st);
            st_launch(dim3(CS_PARTS, 3), 256, 0, st)(nchw3_channel_sum_kernel, s0, sc, N, static_cast<size_t>(H) * W, ws);
            ST_CHECK_LAUNCH("nchw3_channel_sum_kernel");
            st_launch(1, 96, 0, st)(nchw3_channel_sum_finish_kernel, ws, db);
            ST_CHECK_LAUNCH("nchw3_channel_sum_finish_kernel");
        }
        return ST_OK;
    }
    const int rows = 6 / CS;
    const int work = N * ((H + rows - 1) / rows) * ((W + CR_WC - 1) / CR_WC);
    const int grid = work < conv9_grid() ? work : conv9_grid();
    const size_t smem = (static_cast<size_t>(rows) * CR_WC * CB + CS * (rows + 2 * PAD9) * CR_SWS) * sizeof(float);
    static bool set = false;
    if (!set) {
        cudaFuncSetAttribute(conv9_correlate_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
        cudaFuncSetAttribute(conv9_correlate_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
        set = true;
    }
    if (CS == 3) st_launch(grid, CR_THREADS, smem, st)(conv9_correlate_kernel<3>, b_hi, b_lo, planes, s0, s1, sc, ws, N, H, W);
    else st_launch(grid, CR_THREADS, smem, st)(conv9_correlate_kernel<6>, b_hi, b_lo, planes, s0, s1, sc, ws, N, H, W);
    ST_CHECK_LAUNCH("conv9_correlate_kernel");
    const int total = CB * CS * 81;
    st_launch((total + 255) / 256, 256, 0, st)(conv9_wgrad_finish_kernel, ws, grid * rows, CS, mode, big_slot, nullptr, dw);
    ST_CHECK_LAUNCH("conv9_wgrad_finish_kernel");
    if (db) {
        if (mode == 0) return ew_channel_sum(b_hi, b_lo, planes, big_slot, N * H * W, CB, CB, db, ws, st);
        st_launch(dim3(CS_PARTS, 3), 256, 0, st)(nchw3_channel_sum_kernel, s0, sc, N, static_cast<size_t>(H) * W, ws);
        ST_CHECK_LAUNCH("nchw3_channel_sum_kernel");
        st_launch(1, 96, 0, st)(nchw3_channel_sum_finish_kernel, ws, db);
        ST_CHECK_LAUNCH("nchw3_channel_sum_finish_kernel");
    }
    return ST_OK;
}
