// Bandwidth kernels of the VGG loss network: layout conversion, weight packing,
// fused ReLU+maxpool forward/backward, MSE losses and their gradients, the style
// (Gram) MSE and its gradient seed.  All are coalesced 128-bit accesses over the
// split-bf16 NHWC format; reductions are two-stage and deterministic.
// Reference call sites: nn.ReLU/nn.MaxPool2d inside the VGG pieces
// (network.py:257-314), F.mse_loss (network.py:118, :163, :193).
#include "common.cuh"

namespace {

// ---------------------------------------------------------------------------
// NCHW fp32 (one or two sources, channel-concatenated: network.py:984-986) ->
// NHWC split-bf16 with the channel count padded to Cpad (zeros).
// One thread per pixel; reads are coalesced along W per channel plane.
// ---------------------------------------------------------------------------
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                    bf16* __restrict__ hi, bf16* __restrict__ lo,
                                    int N, int C1, int C2, int HW, int Cpad, int planes) {
    pdl_prologue();
    const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<size_t>(N) * HW) return;
    const int n = idx / HW; const int px = idx - static_cast<size_t>(n) * HW;
    for (int c0 = 0; c0 < Cpad; c0 += 8) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int c = c0 + i;
            float x = 0.f;
            if (c < C1) x = a[(static_cast<size_t>(n) * C1 + c) * HW + px];
            else if (c < C1 + C2) x = b[(static_cast<size_t>(n) * C2 + (c - C1)) * HW + px];
            v[i] = x;
        }
        store8p(hi, lo, planes, idx * Cpad + c0, v);
    }
}

// NHWC split-bf16 (Cpad channels stored) -> NCHW fp32 (first C channels)
__global__ void nhwc_to_nchw_kernel(const bf16* __restrict__ hi, const bf16* __restrict__ lo,
                                    float* __restrict__ out, int N, int C, int HW, int Cpad, int planes,
                                    const float* __restrict__ slot) {
    pdl_prologue();
    const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<size_t>(N) * HW) return;
    const int n = idx / HW; const int px = idx - static_cast<size_t>(n) * HW;
    const float mul = slot_inv(slot);
    for (int c0 = 0; c0 < C; c0 += 8) {
        float v[8];
        load8p(hi, lo, planes, idx * Cpad + c0, v);
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (c0 + i < C) out[(static_cast<size_t>(n) * C + c0 + i) * HW + px] = v[i] * mul;
    }
}

// ---------------------------------------------------------------------------
// Weight packing: PyTorch [Cout][Cin][ks][ks] fp32 -> [rows_pad][taps][cols_pad]
// split-bf16, K-major.   mode 0: rows=co, cols=ci, tap=(r,s)          (forward)
//                        mode 1: rows=ci, cols=co, tap=(r,s)          (dgrad, gather form)
//                        mode 2: rows=ci, cols=co, tap=(ks-1-r,ks-1-s) (dgrad as a forward conv)
// ---------------------------------------------------------------------------
__global__ void pack_weight_kernel(const float* __restrict__ w, bf16* __restrict__ hi, bf16* __restrict__ lo,
                                   int Cout, int Cin, int ks, int rows_pad, int cols_pad, int mode, int planes) {
    pdl_prologue();
    const int taps = ks * ks;
    const int total = rows_pad * taps * cols_pad;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    int q = idx; const int col = q % cols_pad; q /= cols_pad; const int tap = q % taps; const int row = q / taps;
    int r = tap / ks, s = tap % ks;
    int co, ci;
    if (mode == 0) { co = row; ci = col; } else { ci = row; co = col; }
    if (mode == 2) { r = ks - 1 - r; s = ks - 1 - s; }
    float v = 0.f;
    if (co < Cout && ci < Cin) v = w[((static_cast<size_t>(co) * Cin + ci) * ks + r) * ks + s];
    if (pc_f16(planes)) {
        // fp16 planes of ST_W_F16_SCALE * w (the consumers multiply their accumulators by ST_W_F16_INV_SCALE)
        uint32_t h2, l2; split2h(v * ST_W_F16_SCALE, 0.f, h2, l2);
        reinterpret_cast<uint16_t*>(hi)[idx] = static_cast<uint16_t>(h2 & 0xffffu);
        if (lo && pc_planes(planes) > 1) reinterpret_cast<uint16_t*>(lo)[idx] = static_cast<uint16_t>(l2 & 0xffffu);
        return;
    }
    bf16 h, l; split_bf16(v, h, l);
    hi[idx] = h;
    if (lo && pc_planes(planes) > 1) lo[idx] = l;
    bf16* p2 = third_plane(hi, lo, planes);
    if (p2) p2[idx] = __float2bfloat16_rn(v - __bfloat162float(h) - __bfloat162float(l));
}

// ---------------------------------------------------------------------------
// ReLU + MaxPool2d(2,2) forward (VGG pieces 2 and 5): y = max over the window of relu(x)
// ---------------------------------------------------------------------------
__global__ void relu_maxpool_fwd_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl,
                                        bf16* __restrict__ yh, bf16* __restrict__ yl,
                                        int N, int H, int W, int C, int planes) {
    pdl_prologue();
    const int Ho = H / 2, Wo = W / 2, Cv = C / 8;
    const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<size_t>(N) * Ho * Wo * Cv) return;
    size_t q = idx; const int cv = q % Cv; q /= Cv; const int ow = q % Wo; q /= Wo; const int oh = q % Ho; const int n = q / Ho;
    float m[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) m[i] = 0.f;                 // relu folded into the max
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        const size_t off = ((static_cast<size_t>(n) * H + oh * 2 + (d >> 1)) * W + ow * 2 + (d & 1)) * C + cv * 8;
        float v[8]; load8p(xh, xl, planes, off, v);
#pragma unroll
        for (int i = 0; i < 8; ++i) m[i] = fmaxf(m[i], v[i]);
    }
    store8p(yh, yl, planes, ((static_cast<size_t>(n) * Ho + oh) * Wo + ow) * C + cv * 8, m);
}

// Backward of the above, fused with the loss gradient of the tap:
//   g_x = (x is the first maximum of relu(x) in its window and x > 0) ? g_y : 0   (+ add)
// One thread per (window, 8 channels); windows are enumerated over ceil(H/2) x ceil(W/2)
// so that odd borders (which max_pool2d drops) still receive `add`.
__global__ void relu_maxpool_bwd_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl,
                                        const bf16* __restrict__ gyh, const bf16* __restrict__ gyl,
                                        const bf16* __restrict__ addh, const bf16* __restrict__ addl,
                                        bf16* __restrict__ gxh, bf16* __restrict__ gxl,
                                        int N, int H, int W, int C, int x_planes) {
    pdl_prologue();
    const int Ho = H / 2, Wo = W / 2, Hc = (H + 1) / 2, Wc = (W + 1) / 2, Cv = C / 8;
    const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<size_t>(N) * Hc * Wc * Cv) return;
    size_t q = idx; const int cv = q % Cv; q /= Cv; const int ow = q % Wc; q /= Wc; const int oh = q % Hc; const int n = q / Hc;
    const bool pooled = oh < Ho && ow < Wo;
    float x[4][8]; bool inb[4];
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        const int h = oh * 2 + (d >> 1), w = ow * 2 + (d & 1);
        inb[d] = h < H && w < W;
        if (inb[d]) load8p(xh, xl, x_planes, ((static_cast<size_t>(n) * H + h) * W + w) * C + cv * 8, x[d]);
    }
    float g[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = 0.f;
    if (pooled) load8(gyh, gyl, ((static_cast<size_t>(n) * Ho + oh) * Wo + ow) * C + cv * 8, g);
    int arg[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float best = -1.f, bx = 0.f; int a = -1;
        if (pooled) {
#pragma unroll
            for (int d = 0; d < 4; ++d) { const float v = fmaxf(x[d][i], 0.f); if (v > best) { best = v; a = d; bx = x[d][i]; } }
            if (!(bx > 0.f)) a = -1;                         // relu'(x) = 0
        }
        arg[i] = a;
    }
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        if (!inb[d]) continue;
        const int h = oh * 2 + (d >> 1), w = ow * 2 + (d & 1);
        const size_t off = ((static_cast<size_t>(n) * H + h) * W + w) * C + cv * 8;
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = (arg[i] == d) ? g[i] : 0.f;
        if (addh) {
            float a[8]; load8(addh, addl, off, a);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] += a[i];
        }
        store8(gxh, gxl, off, o);
    }
}

// ---- the same pair with the routing decision saved by the forward: arg[n][oh][ow][c] (one byte) = position
// 0..3 of the first maximum of relu(x) in the 2x2 window, or 4 when relu'(x) = 0 there.  The backward then reads
// 1 byte per pooled element instead of all planes of the full-resolution tap (100 MB -> 4 MB at 512x512x64).
__global__ void relu_maxpool_fwd_idx_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl,
                                            bf16* __restrict__ yh, bf16* __restrict__ yl, unsigned char* __restrict__ arg,
                                            int N, int H, int W, int C, int planes) {
    pdl_prologue();
    const int Ho = H / 2, Wo = W / 2, Cv = C / 8;
    const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<size_t>(N) * Ho * Wo * Cv) return;
    size_t q = idx; const int cv = q % Cv; q /= Cv; const int ow = q % Wo; q /= Wo; const int oh = q % Ho; const int n = q / Ho;
    float x[4][8];
    {
        Raw8p r[4];                                          // every load of the window in flight before the first conversion
#pragma unroll
        for (int d = 0; d < 4; ++d)
            load8p_raw(xh, xl, planes, ((static_cast<size_t>(n) * H + oh * 2 + (d >> 1)) * W + ow * 2 + (d & 1)) * C + cv * 8, r[d]);
#pragma unroll
        for (int d = 0; d < 4; ++d) decode8p(r[d], planes, x[d]);
    }
    float m[8]; uint32_t a_lo = 0, a_hi = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float best = -1.f, bx = 0.f; uint32_t a = 4;
#pragma unroll
        for (int d = 0; d < 4; ++d) { const float v = fmaxf(x[d][i], 0.f); if (v > best) { best = v; a = d; bx = x[d][i]; } }
        if (!(bx > 0.f)) a = 4;                              // relu'(x) = 0: nothing is routed
        m[i] = best;
        if (i < 4) a_lo |= a << (8 * i); else a_hi |= a << (8 * (i - 4));
    }
    const size_t o = ((static_cast<size_t>(n) * Ho + oh) * Wo + ow) * C + cv * 8;
    store8p(yh, yl, planes, o, m);
    *reinterpret_cast<uint2*>(arg + o) = make_uint2(a_lo, a_hi);
}
__global__ void relu_maxpool_bwd_idx_kernel(const unsigned char* __restrict__ arg,
                                            const bf16* __restrict__ gyh, const bf16* __restrict__ gyl,
                                            const bf16* __restrict__ addh, const bf16* __restrict__ addl,
                                            bf16* __restrict__ gxh, bf16* __restrict__ gxl, int N, int H, int W, int C) {
    pdl_prologue();
    const int Ho = H / 2, Wo = W / 2, Hc = (H + 1) / 2, Wc = (W + 1) / 2, Cv = C / 8;
    const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<size_t>(N) * Hc * Wc * Cv) return;
    size_t q = idx; const int cv = q % Cv; q /= Cv; const int ow = q % Wc; q /= Wc; const int oh = q % Hc; const int n = q / Hc;
    const bool pooled = oh < Ho && ow < Wo;
    // every load of this thread (pooled gradient, routing bytes, the four addends) is issued before the first use
    const size_t o = ((static_cast<size_t>(n) * Ho + (pooled ? oh : 0)) * Wo + (pooled ? ow : 0)) * C + cv * 8;
    Raw8p rg, ra[4];
    load8p_raw(gyh, gyl, 2, o, rg);
    uint2 a = *reinterpret_cast<const uint2*>(arg + o);
    size_t off[4]; bool in[4];
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        const int h = oh * 2 + (d >> 1), w = ow * 2 + (d & 1);
        in[d] = h < H && w < W;
        off[d] = ((static_cast<size_t>(n) * H + (in[d] ? h : oh * 2)) * W + (in[d] ? w : ow * 2)) * C + cv * 8;
        if (addh) load8p_raw(addh, addl, 2, off[d], ra[d]);
    }
    float g[8];
    decode8p(rg, 2, g);
    if (!pooled) {
        a = make_uint2(0x04040404u, 0x04040404u);
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = 0.f;
    }
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        if (!in[d]) continue;
        float ov[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t ai = ((i < 4 ? a.x : a.y) >> (8 * (i & 3))) & 0xffu;
            ov[i] = (ai == static_cast<uint32_t>(d)) ? g[i] : 0.f;
        }
        if (addh) {
            float ad[8]; decode8p(ra[d], 2, ad);
#pragma unroll
            for (int i = 0; i < 8; ++i) ov[i] += ad[i];
        }
        store8(gxh, gxl, off[d], ov);
    }
}

// ---------------------------------------------------------------------------
// MSE forward: partial[b] = sum over this block's slice of (f(x) - f(t))^2,
// f = identity or relu (FeatureReconstructionLoss sits on ReLU_4, network.py:230-232)
// ---------------------------------------------------------------------------
__global__ void mse_partial_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl, int x_code,
                                   const bf16* __restrict__ th, const bf16* __restrict__ tl, int t_code,
                                   size_t nvec, int relu, float* __restrict__ partial) {
    pdl_prologue();
    __shared__ float sm[32];
    float s = 0.f;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        float a[8], b[8];
        Raw8p ra, rb;
        load8p_raw(xh, xl, x_code, i * 8, ra); load8p_raw(th, tl, t_code, i * 8, rb);
        decode8p(ra, x_code, a); decode8p(rb, t_code, b);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float d = relu ? (fmaxf(a[k], 0.f) - fmaxf(b[k], 0.f)) : (a[k] - b[k]);
            s += d * d;
        }
    }
    s = block_sum(s, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

// out[0] = scale * sum(partial[0..n)) ; optionally squared first (feature loss: mse^2 / numel)
// amax_out (optional): also max(partial[n .. 2n)) -- the second half of the partials holds per-block maxima
__global__ void sum_partials_kernel(const float* __restrict__ partial, int n, float scale, int square,
                                    float post_scale, float* __restrict__ out, float* __restrict__ amax_out) {
    pdl_prologue();
    __shared__ float sm[32];
    float s = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
    s = block_sum(s, sm);
    if (threadIdx.x == 0) { s *= scale; if (square) s = s * s * post_scale; out[0] = s; }
    if (amax_out) {
        float m = 0.f;
        for (int i = threadIdx.x; i < n; i += blockDim.x) m = fmaxf(m, partial[n + i]);
        m = block_max(m, sm);
        if (threadIdx.x == 0) amax_out[0] = m;
    }
}

// MSE backward: g = coef * (f(x) - f(t)) * f'(x),  coef = scale * (*gw) * (mul ? *mul : 1)
__global__ void mse_bwd_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl, int x_code,
                               const bf16* __restrict__ th, const bf16* __restrict__ tl, int t_code,
                               size_t nvec, int relu, float scale, const float* __restrict__ gw,
                               const float* __restrict__ mul, int accumulate,
                               bf16* __restrict__ gh, bf16* __restrict__ gl) {
    pdl_prologue();
    const float coef = scale * gw[0] * (mul ? mul[0] : 1.f);
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        float a[8], b[8], o[8];
        Raw8p ra, rb;
        load8p_raw(xh, xl, x_code, i * 8, ra); load8p_raw(th, tl, t_code, i * 8, rb);
        decode8p(ra, x_code, a); decode8p(rb, t_code, b);
        if (accumulate) load8(gh, gl, i * 8, o);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float d = relu ? ((a[k] > 0.f) ? (a[k] - fmaxf(b[k], 0.f)) * coef : 0.f) : (a[k] - b[k]) * coef;
            o[k] = accumulate ? o[k] + d : d;
        }
        store8(gh, gl, i * 8, o);
    }
}

// ---------------------------------------------------------------------------
// Style loss on the (small) Gram matrices: diff = G - T (T broadcast over the batch),
// partial sums of diff^2 per block.   network.py:110-123
// ---------------------------------------------------------------------------
// partial[0 .. grid) = per-block sums of diff^2, partial[grid .. 2 grid) = per-block max |diff|
__global__ void style_diff_kernel(const float* __restrict__ G, const float* __restrict__ T, int CC, int total,
                                  float* __restrict__ diff, float* __restrict__ partial) {
    pdl_prologue();
    __shared__ float sm[32];
    float s = 0.f, m = 0.f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const float d = G[i] - T[i % CC];
        diff[i] = d;
        s += d * d; m = fmaxf(m, fabsf(d));
    }
    s = block_sum(s, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
    m = block_max(m, sm);
    if (threadIdx.x == 0) partial[gridDim.x + blockIdx.x] = m;
}

// D = (*gw) * scale * diff, packed split-bf16 [B*C rows][C] (the per-image "weights" of the
// 1x1 conv that evaluates dF = F * D on the tensor cores)
// f16 (plane code bit): D is stored as fp16 planes of s * D with the power of two s that maps
// max|D| = amax(diff) * |scale * gw| to 2^10; slot = {s, 1/s} for the epilogue of the dF = F * D convolution
__global__ void style_seed_kernel(const float* __restrict__ diff, int total, float scale, const float* __restrict__ gw,
                                  bf16* __restrict__ dh, bf16* __restrict__ dl, int code,
                                  const float* __restrict__ diff_amax, float* __restrict__ slot) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const float coef = scale * gw[0];
    if (pc_f16(code)) {
        const float sc = pow2_scale_for(diff_amax[0] * fabsf(coef), 10.f);
        if (i == 0) { slot[0] = sc; slot[1] = 1.f / sc; }
        if (i >= total) return;
        uint32_t h2, l2; split2h(diff[i] * coef * sc, 0.f, h2, l2);
        reinterpret_cast<uint16_t*>(dh)[i] = static_cast<uint16_t>(h2 & 0xffffu);
        if (dl) reinterpret_cast<uint16_t*>(dl)[i] = static_cast<uint16_t>(l2 & 0xffffu);
        return;
    }
    if (i >= total) return;
    bf16 h, l; split_bf16(diff[i] * coef, h, l);
    dh[i] = h;
    if (dl) dl[i] = l;
}

// the same for every style node of a loss network at once: blockIdx.y = node
__global__ void style_seed_multi_kernel(const StyleSeedTable t) {
    pdl_prologue();
    const StyleSeedEntry& e = t.e[blockIdx.y];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const float coef = e.scale * e.gw[0];
    if (pc_f16(e.code)) {
        const float sc = pow2_scale_for(e.amax[0] * fabsf(coef), 10.f);
        if (i == 0) { e.slot[0] = sc; e.slot[1] = 1.f / sc; }
        if (i >= e.total) return;
        uint32_t h2, l2; split2h(e.diff[i] * coef * sc, 0.f, h2, l2);
        reinterpret_cast<uint16_t*>(e.dh)[i] = static_cast<uint16_t>(h2 & 0xffffu);
        if (e.dl) reinterpret_cast<uint16_t*>(e.dl)[i] = static_cast<uint16_t>(l2 & 0xffffu);
        return;
    }
    if (i >= e.total) return;
    bf16 h, l; split_bf16(e.diff[i] * coef, h, l);
    e.dh[i] = h;
    if (e.dl) e.dl[i] = l;
}

// total = sum_i losses[i] * weights[i] + e0 * s0 + e1 * s1: the weighted sum a training closure returns, in one launch
// (the reference's stack / sum / mul / add chain of the loss getters, network.py:358-364, :704-722)
__global__ void loss_total_kernel(const float* __restrict__ losses, const float* __restrict__ weights, int n,
                                  const float* __restrict__ e0, float s0, const float* __restrict__ e1, float s1,
                                  float* __restrict__ out) {
    pdl_prologue();
    float t = 0.f;
    for (int i = threadIdx.x; i < n; i += 32) t += losses[i] * weights[i];
    t = warp_sum(t);
    if (threadIdx.x == 0) out[0] = t + (e0 ? e0[0] * s0 : 0.f) + (e1 ? e1[0] * s1 : 0.f);
}

// max |x| over a flat fp32 array: per-block partial maxima, then one block -> slot = {s, 1/s} with the power of two s
// that maps the maximum to 2^target_log2 (the scale under which a gradient image is fed to fp16 tensor-core operands)
__global__ void absmax_partial_kernel(const float* __restrict__ x, size_t n, float* __restrict__ partial) {
    pdl_prologue();
    __shared__ float sm[32];
    float m = 0.f;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x)
        m = fmaxf(m, fabsf(x[i]));
    m = block_max(m, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = m;
}
__global__ void absmax_finish_kernel(const float* __restrict__ partial, int n, float target_log2, float* __restrict__ slot) {
    pdl_prologue();
    __shared__ float sm[32];
    float m = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) m = fmaxf(m, partial[i]);
    m = block_max(m, sm);
    if (threadIdx.x == 0) { const float sc = pow2_scale_for(m, target_log2); slot[0] = sc; slot[1] = 1.f / sc; }
}

// per-channel sum over pixels (conv bias gradient): partial[blk][C], then finish
// per-channel sum over pixels (conv bias gradient): 8 channels per thread (128-bit loads),
// 256/Cv pixel lanes per block, grid-stride over pixels; per-block partials, then a
// fixed-order finish (deterministic, no atomics)
__global__ void __launch_bounds__(256)
channel_sum_partial_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl, int code,
                           int P, int C, float* __restrict__ partial) {
    pdl_prologue();
    extern __shared__ float sm[];                 // [lanes][C]
    const int Cv = C / 8, lanes = 256 / Cv;
    const int cv = threadIdx.x % Cv, pl = threadIdx.x / Cv;
    float s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = 0.f;
    if (pl < lanes) {
        for (int p = blockIdx.x * lanes + pl; p < P; p += gridDim.x * lanes) {
            float v[8];
            load8p(xh, xl, code, static_cast<size_t>(p) * C + cv * 8, v);
#pragma unroll
            for (int i = 0; i < 8; ++i) s[i] += v[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) sm[pl * C + cv * 8 + i] = s[i];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += 256) {
        float t = 0.f;
        for (int l = 0; l < lanes; ++l) t += sm[l * C + c];
        partial[static_cast<size_t>(blockIdx.x) * C + c] = t;
    }
}
// out[c] = sum_b partial[b][c]: 32 columns x 8 row lanes per block
__global__ void __launch_bounds__(256)
channel_sum_finish_kernel(const float* __restrict__ partial, int nblk, int C, int Cout, const float* __restrict__ slot,
                          float* __restrict__ out) {
    pdl_prologue();
    __shared__ float sm[8][33];
    const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl;
    float s = 0.f;
    if (c < C) {
        float s1 = 0.f, s2 = 0.f, s3 = 0.f;        // independent accumulators: the loads are L2 round trips
        int b = rl;
        for (; b + 24 < nblk; b += 32) {
            s += partial[static_cast<size_t>(b) * C + c];       s1 += partial[static_cast<size_t>(b + 8) * C + c];
            s2 += partial[static_cast<size_t>(b + 16) * C + c]; s3 += partial[static_cast<size_t>(b + 24) * C + c];
        }
        for (; b < nblk; b += 8) s += partial[static_cast<size_t>(b) * C + c];
        s = (s + s1) + (s2 + s3);
    }
    sm[rl][cl] = s;
    __syncthreads();
    if (rl == 0 && c < Cout) {
        float t = 0.f;
#pragma unroll
        for (int l = 0; l < 8; ++l) t += sm[l][cl];
        out[c] = t * slot_inv(slot);
    }
}

// ---------------------------------------------------------------------------
// Image boundary (reference img_utils.py:33-42 and :105-109), one thread per pixel:
//   u8 HWC [N][H][W][3] -> normalised NCHW fp32: ((x / 255) - mean) / std      (ToTensor + Normalize)
//   NCHW fp32 -> u8 HWC: clamp(x * std + mean, 0, 1) * 255 + 0.5               (imshow's de-normalise + clamp)
// Every operation is a separately rounded IEEE fp32 op (no FMA contraction), so the results are bit-identical to
// the numpy / torch CPU code of the reference.
// ---------------------------------------------------------------------------
struct ImgStats { float mean[3], std[3]; };
__global__ void image_u8_to_nchw_kernel(const unsigned char* __restrict__ in, float* __restrict__ out, size_t pixels, size_t HW, ImgStats st) {
    pdl_prologue();
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    const size_t n = i / HW, px = i - n * HW;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float v = __fdiv_rn(static_cast<float>(in[i * 3 + c]), 255.0f);
        out[(n * 3 + c) * HW + px] = __fdiv_rn(__fsub_rn(v, st.mean[c]), st.std[c]);
    }
}
__global__ void image_nchw_to_u8_kernel(const float* __restrict__ in, unsigned char* __restrict__ out, size_t pixels, size_t HW, ImgStats st) {
    pdl_prologue();
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    const size_t n = i / HW, px = i - n * HW;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float v = __fadd_rn(__fmul_rn(in[(n * 3 + c) * HW + px], st.std[c]), st.mean[c]);
        v = fminf(fmaxf(v, 0.0f), 1.0f);
        out[i * 3 + c] = static_cast<unsigned char>(__fadd_rn(__fmul_rn(v, 255.0f), 0.5f));      // truncation, as torch's .to(uint8)
    }
}

inline int grid_for(size_t n, int block, int cap_mult = 8) {
    size_t g = (n + block - 1) / block;
    size_t cap = static_cast<size_t>(st_num_sms()) * cap_mult;
    return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

int ew_nchw_to_nhwc(const float* a, const float* b, bf16* hi, bf16* lo, int planes, int N, int C1, int C2, int H, int W, int Cpad, cudaStream_t st) {
    const size_t n = static_cast<size_t>(N) * H * W;
    st_launch((n + 255) / 256, 256, 0, st)(nchw_to_nhwc_kernel, a, b, hi, lo, N, C1, C2, H * W, Cpad, planes);
    ST_CHECK_LAUNCH("nchw_to_nhwc_kernel"); return ST_OK;
}
int ew_nhwc_to_nchw(const bf16* hi, const bf16* lo, int planes, const float* slot, float* out, int N, int C, int H, int W, int Cpad, cudaStream_t st) {
    const size_t n = static_cast<size_t>(N) * H * W;
    st_launch((n + 255) / 256, 256, 0, st)(nhwc_to_nchw_kernel, hi, lo, out, N, C, H * W, Cpad, planes, slot);
    ST_CHECK_LAUNCH("nhwc_to_nchw_kernel"); return ST_OK;
}
int ew_pack_weight(const float* w, bf16* hi, bf16* lo, int planes, int Cout, int Cin, int ks, int rows_pad, int cols_pad, int mode, cudaStream_t st) {
    const int total = rows_pad * ks * ks * cols_pad;
    st_launch((total + 255) / 256, 256, 0, st)(pack_weight_kernel, w, hi, lo, Cout, Cin, ks, rows_pad, cols_pad, mode, planes);
    ST_CHECK_LAUNCH("pack_weight_kernel"); return ST_OK;
}
int ew_relu_maxpool_fwd(const bf16* xh, const bf16* xl, bf16* yh, bf16* yl, int planes, int N, int H, int W, int C, cudaStream_t st) {
    const size_t n = static_cast<size_t>(N) * (H / 2) * (W / 2) * (C / 8);
    if (n == 0) return ST_OK;
    st_launch((n + 255) / 256, 256, 0, st)(relu_maxpool_fwd_kernel, xh, xl, yh, yl, N, H, W, C, planes);
    ST_CHECK_LAUNCH("relu_maxpool_fwd_kernel"); return ST_OK;
}
int ew_relu_maxpool_fwd_idx(const bf16* xh, const bf16* xl, bf16* yh, bf16* yl, unsigned char* arg, int planes, int N, int H, int W,
                            int C, cudaStream_t st) {
    const size_t n = static_cast<size_t>(N) * (H / 2) * (W / 2) * (C / 8);
    st_launch((n + 255) / 256, 256, 0, st)(relu_maxpool_fwd_idx_kernel, xh, xl, yh, yl, arg, N, H, W, C, planes);
    ST_CHECK_LAUNCH("relu_maxpool_fwd_idx_kernel"); return ST_OK;
}
int ew_relu_maxpool_bwd_idx(const unsigned char* arg, const bf16* gyh, const bf16* gyl, const bf16* addh, const bf16* addl,
                            bf16* gxh, bf16* gxl, int N, int H, int W, int C, cudaStream_t st) {
    const size_t n = static_cast<size_t>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
    st_launch((n + 255) / 256, 256, 0, st)(relu_maxpool_bwd_idx_kernel, arg, gyh, gyl, addh, addl, gxh, gxl, N, H, W, C);
    ST_CHECK_LAUNCH("relu_maxpool_bwd_idx_kernel"); return ST_OK;
}
int ew_relu_maxpool_bwd(const bf16* xh, const bf16* xl, const bf16* gyh, const bf16* gyl, const bf16* addh, const bf16* addl,
                        bf16* gxh, bf16* gxl, int x_planes, int N, int H, int W, int C, cudaStream_t st) {
    const size_t n = static_cast<size_t>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
    st_launch((n + 255) / 256, 256, 0, st)(relu_maxpool_bwd_kernel, xh, xl, gyh, gyl, addh, addl, gxh, gxl, N, H, W, C, x_planes);
    ST_CHECK_LAUNCH("relu_maxpool_bwd_kernel"); return ST_OK;
}
// workspace: >= 4096 floats
int ew_mse_fwd(const bf16* xh, const bf16* xl, int x_code, const bf16* th, const bf16* tl, int t_code, size_t n, int relu, int feature_form,
               float* loss_out, float* ws, cudaStream_t st) {
    const size_t nvec = n / 8;
    int grid = grid_for(nvec, 256, 4); if (grid > 4096) grid = 4096;
    st_launch(grid, 256, 0, st)(mse_partial_kernel, xh, xl, x_code, th, tl, t_code, nvec, relu, ws);
    ST_CHECK_LAUNCH("mse_partial_kernel");
    // feature_form: (mse)^2 / numel   network.py:193-199
    st_launch(1, 256, 0, st)(sum_partials_kernel, ws, grid, 1.0f / static_cast<float>(n), feature_form, 1.0f / static_cast<float>(n), loss_out, nullptr);
    ST_CHECK_LAUNCH("sum_partials_kernel"); return ST_OK;
}
int ew_mse_bwd(const bf16* xh, const bf16* xl, int x_code, const bf16* th, const bf16* tl, int t_code, size_t n, int relu, float scale,
               const float* gw, const float* mul, int accumulate, bf16* gh, bf16* gl, cudaStream_t st) {
    const size_t nvec = n / 8;
    st_launch(grid_for(nvec, 256), 256, 0, st)(mse_bwd_kernel, xh, xl, x_code, th, tl, t_code, nvec, relu, scale, gw, mul, accumulate, gh, gl);
    ST_CHECK_LAUNCH("mse_bwd_kernel"); return ST_OK;
}
// ws: >= 4096 floats (2 x 2048 per-block partials); diff_amax (optional): receives max |diff|
int ew_style_fwd(const float* G, const float* T, int B, int C, float* diff, float* loss_out, float* diff_amax, float* ws, cudaStream_t st) {
    const int total = B * C * C;
    int grid = grid_for(total, 256, 2); if (grid > 2048) grid = 2048;
    st_launch(grid, 256, 0, st)(style_diff_kernel, G, T, C * C, total, diff, ws);
    ST_CHECK_LAUNCH("style_diff_kernel");
    st_launch(1, 256, 0, st)(sum_partials_kernel, ws, grid, 1.0f / static_cast<float>(total), 0, 1.f, loss_out, diff_amax);
    ST_CHECK_LAUNCH("sum_partials_kernel"); return ST_OK;
}
int ew_style_seed(const float* diff, int B, int C, float scale, const float* gw, bf16* dh, bf16* dl, int code,
                  const float* diff_amax, float* slot, cudaStream_t st) {
    const int total = B * C * C;
    st_launch((total + 255) / 256, 256, 0, st)(style_seed_kernel, diff, total, scale, gw, dh, dl, code, diff_amax, slot);
    ST_CHECK_LAUNCH("style_seed_kernel"); return ST_OK;
}
int ew_style_seed_multi(const StyleSeedTable& t, int max_total, cudaStream_t st) {
    st_launch(dim3((max_total + 255) / 256, t.n), 256, 0, st)(style_seed_multi_kernel, t);
    ST_CHECK_LAUNCH("style_seed_multi_kernel"); return ST_OK;
}
int ew_loss_total(const float* losses, const float* weights, int n, const float* e0, float s0, const float* e1, float s1, float* out, cudaStream_t st) {
    st_launch(1, 32, 0, st)(loss_total_kernel, losses, weights, n, e0, s0, e1, s1, out);
    ST_CHECK_LAUNCH("loss_total_kernel"); return ST_OK;
}
int ew_image_u8_to_nchw(const unsigned char* in, float* out, int N, int H, int W, const float* mean3, const float* std3, cudaStream_t st) {
    ImgStats s; for (int c = 0; c < 3; ++c) { s.mean[c] = mean3[c]; s.std[c] = std3[c]; }
    const size_t px = static_cast<size_t>(N) * H * W;
    st_launch((px + 255) / 256, 256, 0, st)(image_u8_to_nchw_kernel, in, out, px, static_cast<size_t>(H) * W, s);
    ST_CHECK_LAUNCH("image_u8_to_nchw_kernel"); return ST_OK;
}
int ew_image_nchw_to_u8(const float* in, unsigned char* out, int N, int H, int W, const float* mean3, const float* std3, cudaStream_t st) {
    ImgStats s; for (int c = 0; c < 3; ++c) { s.mean[c] = mean3[c]; s.std[c] = std3[c]; }
    const size_t px = static_cast<size_t>(N) * H * W;
    st_launch((px + 255) / 256, 256, 0, st)(image_nchw_to_u8_kernel, in, out, px, static_cast<size_t>(H) * W, s);
    ST_CHECK_LAUNCH("image_nchw_to_u8_kernel"); return ST_OK;
}
// slot = {s, 1/s}: power of two mapping max|x| to 2^target_log2; ws >= 1024 floats
int ew_absmax_slot(const float* x, size_t n, float target_log2, float* slot, float* ws, cudaStream_t st) {
    int grid = grid_for(n, 1024, 2); if (grid > 1024) grid = 1024;
    st_launch(grid, 256, 0, st)(absmax_partial_kernel, x, n, ws);
    ST_CHECK_LAUNCH("absmax_partial_kernel");
    st_launch(1, 256, 0, st)(absmax_finish_kernel, ws, grid, target_log2, slot);
    ST_CHECK_LAUNCH("absmax_finish_kernel"); return ST_OK;
}
// ws: >= nblk*C floats with nblk = min(1024, ceil(P/64))
int ew_channel_sum(const bf16* xh, const bf16* xl, int code, const float* slot, int P, int C, int Cout, float* out, float* ws, cudaStream_t st) {
    const int lanes = 256 / (C / 8);
    int nblk = (P + lanes * 8 - 1) / (lanes * 8);               // >= 8 pixels per lane
    const int cap = st_num_sms() * 4 < 1024 ? st_num_sms() * 4 : 1024;   // ws holds >= 1024 * C floats
    if (nblk > cap) nblk = cap;
    if (nblk < 1) nblk = 1;
    st_launch(nblk, 256, lanes * C * sizeof(float), st)(channel_sum_partial_kernel, xh, xl, code, P, C, ws);
    ST_CHECK_LAUNCH("channel_sum_partial_kernel");
    st_launch((C + 31) / 32, 256, 0, st)(channel_sum_finish_kernel, ws, nblk, C, Cout, slot, out);
    ST_CHECK_LAUNCH("channel_sum_finish_kernel"); return ST_OK;
}
