// InstanceNorm2d(affine) forward/backward with fused ReLU and fused nearest x2
// upsample-on-write (network.py:474, :483, :531-600), the total-variation loss
// (network.py:621-641) and the temporal loss (network.py:885-903).
// HBM-bound kernels: 128-bit accesses, channels-last so a warp reads contiguous
// memory, per-(n,c) statistics reduced in two deterministic stages.
#include "common.cuh"
#include <stdlib.h>

namespace {

constexpr int IN_THREADS = 256;

// stage 1: per (image, pixel split) partial sums of a[c] and b[c] over the split's pixels.
// MODE 0 (forward stats): a = x, b = x*x
// MODE 1 (backward stats): a = dy_eff, b = dy_eff * xhat, with
//        dy_eff = (relu ? dy * (gamma*xhat+beta > 0) : dy), dy summed over the 2x2
//        upsampled positions when up != 0.
// gradient of one pixel: dy itself, or the sum over the 2x2 fan-out of a fused nearest upsample
template <int UP>
__device__ __forceinline__ void in_load_grad(const bf16* __restrict__ gh, const bf16* __restrict__ gl,
                                             int n, int HW, int W, int C, int p, int cv, float (&g)[8]) {
    if (!UP) {
        load8(gh, gl, (static_cast<size_t>(n) * HW + p) * C + cv * 8, g);
    } else {
        const int h = p / W, w = p - h * W, W2 = W * 2;
        const size_t base = ((static_cast<size_t>(n) * HW * 4) + static_cast<size_t>(h * 2) * W2 + w * 2) * C + cv * 8;
        float t0[8], t1[8], t2[8], t3[8];
        load8(gh, gl, base, t0); load8(gh, gl, base + C, t1);
        load8(gh, gl, base + static_cast<size_t>(W2) * C, t2); load8(gh, gl, base + static_cast<size_t>(W2) * C + C, t3);
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = (t0[i] + t1[i]) + (t2[i] + t3[i]);
    }
}

template <int MODE, int UP>
__global__ void __launch_bounds__(IN_THREADS, 2)
in_stats_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl,
                const bf16* __restrict__ gh, const bf16* __restrict__ gl,
                const float* __restrict__ mean, const float* __restrict__ rstd,
                const float* __restrict__ gamma, const float* __restrict__ beta,
                int HW, int W, int C, int splits, int chunk, int relu, int x_planes,
                float* __restrict__ partial /* [N][splits][2][C] */, float* __restrict__ gmax_partial /* [N*splits] or null */) {
    pdl_prologue();
    extern __shared__ float sm[];                 // [lanes][2][C]
    __shared__ float smx[32];
    float gmax = 0.f;                             // MODE 1: max |dy_eff| of this block (bound of the fp16 scale of dx)
    const int n = blockIdx.x / splits, sp = blockIdx.x % splits;
    const int Cv = C / 8, lanes = IN_THREADS / Cv;
    const int cv = threadIdx.x % Cv, pl = threadIdx.x / Cv;
    const int p0 = sp * chunk, p1 = min(HW, p0 + chunk);
    float sa[8], sb[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { sa[i] = 0.f; sb[i] = 0.f; }
    // backward: relu mask  gamma*xhat + beta > 0  <=>  ra*x + rb > 0 ; sums of ge and ge*(x - mu), scaled by rstd at the end
    float mu[8], ra[8], rb[8];
    if (MODE == 1) {
        const float4* m4 = reinterpret_cast<const float4*>(mean + n * C + cv * 8);
        const float4* r4 = reinterpret_cast<const float4*>(rstd + n * C + cv * 8);
        const float4* g4 = reinterpret_cast<const float4*>(gamma + cv * 8);
        const float4* b4 = reinterpret_cast<const float4*>(beta + cv * 8);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float4 m = m4[h], r = r4[h], g = g4[h], b = b4[h];
            const float mm[4] = {m.x, m.y, m.z, m.w}, rr[4] = {r.x, r.y, r.z, r.w};
            const float gg[4] = {g.x, g.y, g.z, g.w}, bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                mu[h * 4 + i] = mm[i];
                ra[h * 4 + i] = relu ? gg[i] * rr[i] : 0.f;
                rb[h * 4 + i] = relu ? bb[i] - gg[i] * rr[i] * mm[i] : 1.f;
            }
        }
    }
    if (pl < lanes) {
        if (MODE == 0) {
            // four pixels per iteration, their (up to) twelve 16-byte loads in flight together; a missing pixel re-reads
            // the first one and contributes nothing
            for (int p = p0 + pl; p < p1; p += 4 * lanes) {
                Raw8p r[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int pj = p + j * lanes < p1 ? p + j * lanes : p;
                    load8p_raw(xh, xl, x_planes, (static_cast<size_t>(n) * HW + pj) * C + cv * 8, r[j]);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float x[8];
                    decode8p(r[j], x_planes, x);
                    if (p + j * lanes < p1) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { sa[i] += x[i]; sb[i] += x[i] * x[i]; }
                    }
                }
            }
        } else {
            for (int p = p0 + pl; p < p1; p += 2 * lanes) {
                const bool two = p + lanes < p1;
                float x0[8], x1[8], g0[8], g1[8];
                if (UP) {
                    load8p(xh, xl, x_planes, (static_cast<size_t>(n) * HW + p) * C + cv * 8, x0);
                    in_load_grad<UP>(gh, gl, n, HW, W, C, p, cv, g0);
                    if (two) {
                        load8p(xh, xl, x_planes, (static_cast<size_t>(n) * HW + p + lanes) * C + cv * 8, x1);
                        in_load_grad<UP>(gh, gl, n, HW, W, C, p + lanes, cv, g1);
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { x1[i] = 0.f; g1[i] = 0.f; }
                    }
                } else {
                    // all ten 16-byte loads of the two pixels are issued before the first conversion (no branch in between);
                    // a missing second pixel re-reads the first one and its gradient is zeroed
                    const size_t o0 = (static_cast<size_t>(n) * HW + p) * C + cv * 8;
                    const size_t o1 = (static_cast<size_t>(n) * HW + (two ? p + lanes : p)) * C + cv * 8;
                    Raw8p rx0, rx1, rg0, rg1;
                    load8p_raw(xh, xl, x_planes, o0, rx0); load8p_raw(gh, gl, 2, o0, rg0);
                    load8p_raw(xh, xl, x_planes, o1, rx1); load8p_raw(gh, gl, 2, o1, rg1);
                    decode8p(rx0, x_planes, x0); decode8p(rg0, 2, g0);
                    decode8p(rx1, x_planes, x1); decode8p(rg1, 2, g1);
                    if (!two) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) { x1[i] = 0.f; g1[i] = 0.f; }
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    // NaN-safe form of the mask: keep the gradient only where the activation is > 0
                    const float e0 = (fmaf(ra[i], x0[i], rb[i]) > 0.f) ? g0[i] : 0.f;
                    const float e1 = (fmaf(ra[i], x1[i], rb[i]) > 0.f) ? g1[i] : 0.f;
                    gmax = fmaxf(gmax, fmaxf(fabsf(e0), fabsf(e1)));
                    sa[i] += e0 + e1;
                    sb[i] = fmaf(e0, x0[i] - mu[i], fmaf(e1, x1[i] - mu[i], sb[i]));
                }
            }
            const float4* r4 = reinterpret_cast<const float4*>(rstd + n * C + cv * 8);
            const float4 r0 = r4[0], r1 = r4[1];
            const float rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) sb[i] *= rr[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            sm[(pl * 2 + 0) * C + cv * 8 + i] = sa[i];
            sm[(pl * 2 + 1) * C + cv * 8 + i] = sb[i];
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < 2 * C; j += IN_THREADS) {
        float s = 0.f;
        for (int l = 0; l < lanes; ++l) s += sm[l * 2 * C + j];
        partial[(static_cast<size_t>(blockIdx.x)) * 2 * C + j] = s;
    }
    if (MODE == 1 && gmax_partial) {
        gmax = block_max(gmax, smx);
        if (threadIdx.x == 0) gmax_partial[blockIdx.x] = gmax;
    }
}

// forward finalize: mean / rstd per (n,c)    (biased variance, eps inside the sqrt)
// block = (32 channels, image n) x 8 split lanes, combined in double in a fixed order
__global__ void __launch_bounds__(256)
in_finalize_fwd_kernel(const float* __restrict__ partial, int C, int splits, int HW, float eps,
                       float* __restrict__ mean, float* __restrict__ rstd) {
    pdl_prologue();
    __shared__ double ss[8][33], sq[8][33];
    const int cl = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl, n = blockIdx.y;
    double s = 0.0, q = 0.0;
    if (c < C) {
        const float* base = partial + (static_cast<size_t>(n) * splits) * 2 * C + c;
        // four independent load pairs per round: the partials come from L2, a dependent chain of splits / 8 round trips
        // made this launch cost as much as a statistics pass at small batch (profiles/r02aj_launches_fast_st_b2_f16x3.csv)
        double s1 = 0.0, q1 = 0.0, s2 = 0.0, q2 = 0.0, s3 = 0.0, q3 = 0.0;
        int sp = sl;
        for (; sp + 24 < splits; sp += 32) {
            const float* b0 = base + static_cast<size_t>(sp) * 2 * C;
            const size_t st = static_cast<size_t>(8) * 2 * C;
            const float x0 = b0[0], y0 = b0[C], x1 = b0[st], y1 = b0[st + C], x2 = b0[2 * st], y2 = b0[2 * st + C], x3 = b0[3 * st], y3 = b0[3 * st + C];
            s += x0; q += y0; s1 += x1; q1 += y1; s2 += x2; q2 += y2; s3 += x3; q3 += y3;
        }
        for (; sp < splits; sp += 8) {
            s += base[static_cast<size_t>(sp) * 2 * C];
            q += base[static_cast<size_t>(sp) * 2 * C + C];
        }
        s = (s + s1) + (s2 + s3); q = (q + q1) + (q2 + q3);
    }
    ss[sl][cl] = s; sq[sl][cl] = q;
    __syncthreads();
    if (sl == 0 && c < C) {
        double S = 0.0, Q = 0.0;
#pragma unroll
        for (int l = 0; l < 8; ++l) { S += ss[l][cl]; Q += sq[l][cl]; }
        const double m = S / HW;
        double var = Q / HW - m * m; if (var < 0.0) var = 0.0;
        mean[n * C + c] = static_cast<float>(m);
        rstd[n * C + c] = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
    }
}

// backward finalize: s1 = sum dy_eff, s2 = sum dy_eff*xhat per (n,c); dgamma/dbeta = sums over n.
// block = 32 channels x 32 lanes; a lane is (image lane, split lane); fixed-order combination (deterministic)
__global__ void __launch_bounds__(1024)
in_finalize_bwd_kernel(const float* __restrict__ partial, int N, int C, int splits,
                       float* __restrict__ s1, float* __restrict__ s2,
                       float* __restrict__ dgamma, float* __restrict__ dbeta,
                       const float* __restrict__ gmax_partial, int ngmax, const float* __restrict__ gamma,
                       const float* __restrict__ rstd, float inv_hw, float* __restrict__ bound_part) {
    pdl_prologue();
    __shared__ float sa[32][33], sb[32][33];
    __shared__ float smx[32];
    // bound of |dx| = |gamma rstd| * |dy_eff - (s1 + xhat s2) / HW| for the fp16 scale of dx (when asked for):
    // max |dy_eff| is measured, |xhat| <= 16 is assumed (the scale leaves another factor 64 of head room)
    float G = 0.f, bmax = 0.f;
    if (bound_part) {
        for (int i = threadIdx.x; i < ngmax; i += 1024) G = fmaxf(G, gmax_partial[i]);
        G = block_max(G, smx);
        if (threadIdx.x == 0) smx[0] = G;
        __syncthreads();
        G = smx[0];
        __syncthreads();
    }
    const int cl = threadIdx.x & 31, l = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl;
    const int NL = N < 32 ? N : 32, SL = 32 / NL;            // image lanes x split lanes (<= 32)
    const int nl = l % NL, sl = l / NL;
    const bool lane_on = sl < SL && c < C;
    float dg = 0.f, db = 0.f;
    for (int n0 = 0; n0 < N; n0 += NL) {
        const int n = n0 + nl;
        float a = 0.f, b = 0.f;
        if (lane_on && n < N) {
            const float* base = partial + (static_cast<size_t>(n) * splits) * 2 * C + c;
            float a1 = 0.f, b1 = 0.f, a2 = 0.f, b2 = 0.f, a3 = 0.f, b3 = 0.f;    // independent accumulators (latency)
            int sp = sl;
            for (; sp + 3 * SL < splits; sp += 4 * SL) {
                const float* q = base + static_cast<size_t>(sp) * 2 * C;
                const size_t st = static_cast<size_t>(SL) * 2 * C;
                a += q[0];       b += q[C];
                a1 += q[st];     b1 += q[st + C];
                a2 += q[2 * st]; b2 += q[2 * st + C];
                a3 += q[3 * st]; b3 += q[3 * st + C];
            }
            for (; sp < splits; sp += SL) {
                a += base[static_cast<size_t>(sp) * 2 * C];
                b += base[static_cast<size_t>(sp) * 2 * C + C];
            }
            a = (a + a1) + (a2 + a3); b = (b + b1) + (b2 + b3);
        }
        sa[l][cl] = a; sb[l][cl] = b;
        __syncthreads();
        if (sl == 0 && c < C && n < N) {
            float A = 0.f, B = 0.f;
            for (int k = 0; k < SL; ++k) { A += sa[k * NL + nl][cl]; B += sb[k * NL + nl][cl]; }
            s1[n * C + c] = A; s2[n * C + c] = B;
            db += A; dg += B;
            if (bound_part) bmax = fmaxf(bmax, fabsf(gamma[c] * rstd[n * C + c]) * (G + (fabsf(A) + 16.f * fabsf(B)) * inv_hw));
        }
        __syncthreads();
    }
    sa[l][cl] = db; sb[l][cl] = dg;
    __syncthreads();
    if (l == 0 && c < C) {
        float g = 0.f, b = 0.f;
        for (int k = 0; k < NL; ++k) { b += sa[k][cl]; g += sb[k][cl]; }      // lanes with sl == 0 are l = nl < NL
        dgamma[c] = g; dbeta[c] = b;
    }
    if (bound_part) {
        bmax = block_max(bmax, smx);
        if (threadIdx.x == 0) bound_part[blockIdx.x] = bmax;
    }
}

// Apply kernels: one block = (image n, pixel chunk); a thread keeps its 8 channels for the whole chunk,
// so the per-channel statistics and affine parameters are loaded into registers once.
__device__ __forceinline__ void load8f(const float* p, float (&v)[8]) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// forward apply: y = gamma * (x - mean) * rstd + beta  (+ReLU) (+ x2 nearest upsample on write)
template <int UP>
__global__ void __launch_bounds__(IN_THREADS)
in_apply_fwd_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl,
                    const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ gamma, const float* __restrict__ beta,
                    bf16* __restrict__ yh, bf16* __restrict__ yl,
                    int HW, int W, int C, int chunk, int relu, int planes) {
    pdl_prologue();
    const int n = blockIdx.y;
    const int Cv = C / 8, lanes = IN_THREADS / Cv;
    const int cv = threadIdx.x % Cv, pl = threadIdx.x / Cv;
    if (pl >= lanes) return;
    const int p0 = blockIdx.x * chunk, p1 = min(HW, p0 + chunk);
    float mu[8], rs[8], ga[8], be[8];
    load8f(mean + n * C + cv * 8, mu); load8f(rstd + n * C + cv * 8, rs);
    load8f(gamma + cv * 8, ga); load8f(beta + cv * 8, be);
    const size_t img = static_cast<size_t>(n) * HW;
    // two pixels per iteration, their loads in flight together (load8p_raw: no branch between the loads)
    for (int p = p0 + pl; p < p1; p += 2 * lanes) {
        const bool two = p + lanes < p1;
        Raw8p r0, r1;
        load8p_raw(xh, xl, planes, (img + p) * C + cv * 8, r0);
        load8p_raw(xh, xl, planes, (img + (two ? p + lanes : p)) * C + cv * 8, r1);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            if (j == 1 && !two) break;
            const int q = p + j * lanes;
            float x[8];
            decode8p(j ? r1 : r0, planes, x);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float y = ga[i] * (x[i] - mu[i]) * rs[i] + be[i];
                x[i] = relu ? fmaxf(y, 0.f) : y;
            }
            if (!UP) {
                store8p(yh, yl, planes, (img + q) * C + cv * 8, x);
            } else {
                const int h = q / W, w = q - h * W;
                const size_t o = ((img * 4) + static_cast<size_t>(h * 2) * (W * 2) + w * 2) * C + cv * 8;
                store8p(yh, yl, planes, o, x); store8p(yh, yl, planes, o + C, x);
                store8p(yh, yl, planes, o + static_cast<size_t>(W) * 2 * C, x); store8p(yh, yl, planes, o + static_cast<size_t>(W) * 2 * C + C, x);
            }
        }
    }
}

// backward apply: dx = gamma*rstd/HW * (HW*dy_eff - s1 - xhat*s2)
template <int UP>
__global__ void __launch_bounds__(IN_THREADS, 2)
in_apply_bwd_kernel(const bf16* __restrict__ xh, const bf16* __restrict__ xl,
                    const bf16* __restrict__ gh, const bf16* __restrict__ gl,
                    const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ s1, const float* __restrict__ s2,
                    bf16* __restrict__ dxh, bf16* __restrict__ dxl, int dx_code,
                    const float* __restrict__ bound_part, int nbound, float* __restrict__ dx_slot,
                    int HW, int W, int C, int chunk, int relu, int x_planes, float* __restrict__ dsum_partial) {
    pdl_prologue();
    extern __shared__ float sm_ds[];                  // [lanes][C], only when dsum_partial != null
    const int n = blockIdx.y;
    // dx_slot: dx is stored multiplied by the power of two that maps its bound to 2^10 (fp16 planes)
    float dx_scale = 1.f;
    if (dx_slot) {
        float b = 0.f;
        for (int i = 0; i < nbound; ++i) b = fmaxf(b, bound_part[i]);
        dx_scale = pow2_scale_for(b, 10.f);
        if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) { dx_slot[0] = dx_scale; dx_slot[1] = 1.f / dx_scale; }
    }
    const int Cv = C / 8, lanes = IN_THREADS / Cv;
    const int cv = threadIdx.x % Cv, pl = threadIdx.x / Cv;
    const int p0 = blockIdx.x * chunk, p1 = min(HW, p0 + chunk);
    const float inv_hw = 1.0f / HW;
    float ds[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) ds[i] = 0.f;
    if (pl < lanes) {
    float mu[8], rs[8], ga[8], be[8], a1[8], a2[8];
    load8f(mean + n * C + cv * 8, mu); load8f(rstd + n * C + cv * 8, rs);
    load8f(gamma + cv * 8, ga); load8f(beta + cv * 8, be);
    load8f(s1 + n * C + cv * 8, a1); load8f(s2 + n * C + cv * 8, a2);
    float gr[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) gr[i] = ga[i] * rs[i];
    const size_t img = static_cast<size_t>(n) * HW;
    auto finish_pixel = [&](float (&x0)[8], const float (&g0)[8], int p) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float xhat = (x0[i] - mu[i]) * rs[i];
            const float ge = (relu && !(ga[i] * xhat + be[i] > 0.f)) ? 0.f : g0[i];
            x0[i] = gr[i] * (ge - inv_hw * (a1[i] + xhat * a2[i]));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) ds[i] += x0[i];
        if (dx_slot) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x0[i] *= dx_scale;
        }
        store8p(dxh, dxl, dx_code, (img + p) * C + cv * 8, x0);
    };
    if (UP) {
        for (int p = p0 + pl; p < p1; p += lanes) {
            float x0[8], g0[8];
            load8p(xh, xl, x_planes, (img + p) * C + cv * 8, x0);
            in_load_grad<UP>(gh, gl, n, HW, W, C, p, cv, g0);
            finish_pixel(x0, g0, p);
        }
    } else {
        // two pixels per iteration, all ten 16-byte loads issued before the first conversion
        for (int p = p0 + pl; p < p1; p += 2 * lanes) {
            const bool two = p + lanes < p1;
            const size_t o0 = (img + p) * C + cv * 8, o1 = (img + (two ? p + lanes : p)) * C + cv * 8;
            Raw8p rx0, rx1, rg0, rg1;
            load8p_raw(xh, xl, x_planes, o0, rx0); load8p_raw(gh, gl, 2, o0, rg0);
            load8p_raw(xh, xl, x_planes, o1, rx1); load8p_raw(gh, gl, 2, o1, rg1);
            float x0[8], g0[8], x1[8], g1[8];
            decode8p(rx0, x_planes, x0); decode8p(rg0, 2, g0);
            decode8p(rx1, x_planes, x1); decode8p(rg1, 2, g1);
            finish_pixel(x0, g0, p);
            if (two) finish_pixel(x1, g1, p + lanes);
        }
    }
    }
    // sum of dx per channel over this block's pixels (= bias gradient of the convolution in front)
    if (dsum_partial) {
        if (pl < lanes) {
#pragma unroll
            for (int i = 0; i < 8; ++i) sm_ds[pl * C + cv * 8 + i] = ds[i];
        }
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += IN_THREADS) {
            float t = 0.f;
            for (int l = 0; l < lanes; ++l) t += sm_ds[l * C + c];
            dsum_partial[(static_cast<size_t>(n) * gridDim.x + blockIdx.x) * C + c] = t;
        }
    }
}

// out[c] = sum over rows of partial[row][c].  8 channels x 128 row lanes per block (C / 8 blocks): with ~1200 rows
// (8 CTAs per SM of the apply kernel) a lane walks ~9 rows in 3 rounds of independent loads instead of 37 in a
// dependent chain on 1..4 blocks; fixed-order tree over the lanes (deterministic)
__global__ void __launch_bounds__(1024)
in_colsum_kernel(const float* __restrict__ partial, int rows, int C, float* __restrict__ out) {
    pdl_prologue();
    __shared__ float sm[128][9];
    const int cl = threadIdx.x & 7, rl = threadIdx.x >> 3;
    const int c = blockIdx.x * 8 + cl;
    float s = 0.f;
    if (c < C) {
        float s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int r = rl;
        for (; r + 384 < rows; r += 512) {
            s += partial[static_cast<size_t>(r) * C + c];        s1 += partial[static_cast<size_t>(r + 128) * C + c];
            s2 += partial[static_cast<size_t>(r + 256) * C + c]; s3 += partial[static_cast<size_t>(r + 384) * C + c];
        }
        for (; r < rows; r += 128) s += partial[static_cast<size_t>(r) * C + c];
        s = (s + s1) + (s2 + s3);
    }
    sm[rl][cl] = s;
    __syncthreads();
#pragma unroll
    for (int half = 64; half >= 1; half >>= 1) {
        if (rl < half) sm[rl][cl] += sm[rl + half][cl];
        __syncthreads();
    }
    if (rl == 0 && c < C) out[c] = sm[0][cl];
}

// ---------------------------------------------------------------------------
// Total variation on an NCHW fp32 image batch: partial sums of |dx| + |dy|
// ---------------------------------------------------------------------------
__global__ void tv_partial_kernel(const float* __restrict__ x, size_t total, int H, int W, float* __restrict__ partial) {
    pdl_prologue();
    __shared__ float sm[32];
    float s = 0.f;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const int w = i % W; const int h = (i / W) % H;
        const float v = x[i];
        if (w + 1 < W) s += fabsf(v - x[i + 1]);
        if (h + 1 < H) s += fabsf(v - x[i + W]);
    }
    s = block_sum(s, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}
__device__ __forceinline__ float sgn(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }
// g = coef * d/dx TV,  coef = factor * (*gw)
__global__ void tv_bwd_kernel(const float* __restrict__ x, size_t total, int H, int W, float factor,
                              const float* __restrict__ gw, float* __restrict__ g) {
    pdl_prologue();
    const float coef = factor * gw[0];
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const int w = i % W; const int h = (i / W) % H;
        const float v = x[i];
        float d = 0.f;
        if (w + 1 < W) d += sgn(v - x[i + 1]);
        if (w > 0) d -= sgn(x[i - 1] - v);
        if (h + 1 < H) d += sgn(v - x[i + W]);
        if (h > 0) d -= sgn(x[i - W] - v);
        g[i] = coef * d;
    }
}
__global__ void finish_scalar_kernel(const float* __restrict__ partial, int n, float scale, float* __restrict__ out) {
    pdl_prologue();
    __shared__ float sm[32];
    float s = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
    s = block_sum(s, sm);
    if (threadIdx.x == 0) out[0] = s * scale;
}

// ---------------------------------------------------------------------------
// Temporal loss: tw * ||cs - os|| / (||cc - oc|| + 1)     (fp32 NCHW tensors)
// ---------------------------------------------------------------------------
__global__ void temporal_partial_kernel(const float* __restrict__ oc, const float* __restrict__ os,
                                        const float* __restrict__ cc, const float* __restrict__ cs,
                                        size_t total, float* __restrict__ partial /* [grid][2] */) {
    pdl_prologue();
    __shared__ float sm[32];
    float a = 0.f, b = 0.f;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const float ds = cs[i] - os[i], dc = cc[i] - oc[i];
        a += ds * ds; b += dc * dc;
    }
    a = block_sum(a, sm);
    b = block_sum(b, sm);
    if (threadIdx.x == 0) { partial[blockIdx.x * 2] = a; partial[blockIdx.x * 2 + 1] = b; }
}
// out[0] = loss, out[1] = ||ds||, out[2] = ||dc||
__global__ void temporal_finish_kernel(const float* __restrict__ partial, int n, float tw, float* __restrict__ out) {
    pdl_prologue();
    __shared__ float sm[32];
    float a = 0.f, b = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { a += partial[i * 2]; b += partial[i * 2 + 1]; }
    a = block_sum(a, sm);
    b = block_sum(b, sm);
    if (threadIdx.x == 0) {
        const float ns = sqrtf(a), nc = sqrtf(b);
        out[0] = tw * ns / (nc + 1.f); out[1] = ns; out[2] = nc;
    }
}
// g = (*gw) * tw / ((nc + 1) * ns) * (cs - os)   (0 when ns == 0, as torch's norm backward)
__global__ void temporal_bwd_kernel(const float* __restrict__ os, const float* __restrict__ cs, size_t total, float tw,
                                    const float* __restrict__ stats, const float* __restrict__ gw, float* __restrict__ g) {
    pdl_prologue();
    const float ns = stats[1], nc = stats[2];
    const float coef = ns > 0.f ? gw[0] * tw / ((nc + 1.f) * ns) : 0.f;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<size_t>(gridDim.x) * blockDim.x)
        g[i] = coef * (cs[i] - os[i]);
}

// ---------------------------------------------------------------------------
// EXTENSION (no reference code: the reference's temporal loss has no optical flow; BASELINE.json
// words it as "temporal (optical-flow-warped) loss").  The previous frame is warped by a dense flow field before
// the comparison:  tw * ||cs - warp(os)|| / (||cc - warp(oc)|| + 1),  warp(img)(x, y) = bilinear(img, x + dx, y + dy)
// with zeros outside the frame (torch grid_sample, bilinear, padding_mode='zeros', align_corners=True in pixel units).
// flow: [N][2][H][W] fp32 (dx, dy) in pixels.  The warped frames are never materialised: both kernels gather on the fly.
// ---------------------------------------------------------------------------
struct WarpTaps { int i00, i01, i10, i11; float w00, w01, w10, w11; };
__device__ __forceinline__ WarpTaps warp_taps(float fx, float fy, int H, int W) {
    WarpTaps t;
    const float x0f = floorf(fx), y0f = floorf(fy);
    const int x0 = (int)x0f, y0 = (int)y0f, x1 = x0 + 1, y1 = y0 + 1;
    const float ax = fx - x0f, ay = fy - y0f;
    const bool vx0 = x0 >= 0 && x0 < W, vx1 = x1 >= 0 && x1 < W, vy0 = y0 >= 0 && y0 < H, vy1 = y1 >= 0 && y1 < H;
    t.w00 = (vx0 && vy0) ? (1.f - ax) * (1.f - ay) : 0.f; t.w01 = (vx1 && vy0) ? ax * (1.f - ay) : 0.f;
    t.w10 = (vx0 && vy1) ? (1.f - ax) * ay : 0.f;         t.w11 = (vx1 && vy1) ? ax * ay : 0.f;
    const int cx0 = min(max(x0, 0), W - 1), cx1 = min(max(x1, 0), W - 1), cy0 = min(max(y0, 0), H - 1), cy1 = min(max(y1, 0), H - 1);
    t.i00 = cy0 * W + cx0; t.i01 = cy0 * W + cx1; t.i10 = cy1 * W + cx0; t.i11 = cy1 * W + cx1;
    return t;
}
__device__ __forceinline__ float warp_gather(const float* __restrict__ plane, const WarpTaps& t) {
    return t.w00 * plane[t.i00] + t.w01 * plane[t.i01] + t.w10 * plane[t.i10] + t.w11 * plane[t.i11];
}
__global__ void temporal_warp_partial_kernel(const float* __restrict__ oc, const float* __restrict__ os,
                                             const float* __restrict__ cc, const float* __restrict__ cs,
                                             const float* __restrict__ flow, int N, int C, int H, int W,
                                             float* __restrict__ partial /* [grid][2] */) {
    pdl_prologue();
    __shared__ float sm[32];
    const size_t HW = (size_t)H * W, pixels = (size_t)N * HW;
    float a = 0.f, b = 0.f;
    for (size_t p = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; p < pixels;
         p += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const int n = (int)(p / HW); const size_t r = p - (size_t)n * HW;
        const int y = (int)(r / W), x = (int)(r - (size_t)y * W);
        const float* f = flow + (size_t)n * 2 * HW;
        const WarpTaps t = warp_taps((float)x + f[r], (float)y + f[HW + r], H, W);
        for (int c = 0; c < C; ++c) {
            const size_t base = ((size_t)n * C + c) * HW;
            const float ds = cs[base + r] - warp_gather(os + base, t), dc = cc[base + r] - warp_gather(oc + base, t);
            a += ds * ds; b += dc * dc;
        }
    }
    a = block_sum(a, sm);
    b = block_sum(b, sm);
    if (threadIdx.x == 0) { partial[blockIdx.x * 2] = a; partial[blockIdx.x * 2 + 1] = b; }
}
__global__ void temporal_warp_bwd_kernel(const float* __restrict__ os, const float* __restrict__ cs, const float* __restrict__ flow,
                                         int N, int C, int H, int W, float tw, const float* __restrict__ stats,
                                         const float* __restrict__ gw, float* __restrict__ g) {
    pdl_prologue();
    const float ns = stats[1], nc = stats[2];
    const float coef = ns > 0.f ? gw[0] * tw / ((nc + 1.f) * ns) : 0.f;
    const size_t HW = (size_t)H * W, pixels = (size_t)N * HW;
    for (size_t p = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; p < pixels;
         p += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const int n = (int)(p / HW); const size_t r = p - (size_t)n * HW;
        const int y = (int)(r / W), x = (int)(r - (size_t)y * W);
        const float* f = flow + (size_t)n * 2 * HW;
        const WarpTaps t = warp_taps((float)x + f[r], (float)y + f[HW + r], H, W);
        for (int c = 0; c < C; ++c) {
            const size_t base = ((size_t)n * C + c) * HW;
            g[base + r] = coef * (cs[base + r] - warp_gather(os + base, t));
        }
    }
}

inline int in_splits(int N, int HW) {
    // STRANSFER_IN_SPLITS_INVARIANT=1: a split that depends on the image size only, so an image's statistics (and with them
    // every ReLU decision downstream) do not depend on how many images share the launch -- used by tools/check_dp.py to
    // compare a data-parallel step with the single-GPU step bit for bit in the forward; 4 % slower at batch 16
    static const bool inv = []() { const char* e = getenv("STRANSFER_IN_SPLITS_INVARIANT"); return e && e[0] == '1'; }();
    int splits = inv ? 128 : (2 * st_num_sms()) / N;              // N * splits = one full wave at 2 CTAs per SM
    // at least 64 pixels per split: a 256-pixel floor left 16 images of 64 x 64 with 256 blocks = 0.43 waves, 21 % of the
    // warp slots busy and 2.7 TB/s (profiles/r02ac_in_kernels_fast_st.txt)
    const int maxs = (HW + 63) / 64;
    if (splits > maxs) splits = maxs;
    if (splits > 128) splits = 128;                  // = IN_MAX_SPLITS (workspace sizing below)
    if (splits < 1) splits = 1;
    return splits;
}
// blocks per image of the apply kernels: about 8 CTAs per SM over the batch, at least 64 pixels each
inline int in_apply_blocks(int N, int HW) {
    int b = (8 * st_num_sms() + N - 1) / N;
    const int maxb = (HW + 63) / 64;
    if (b > maxb) b = maxb;
    return b < 1 ? 1 : b;
}
inline int ew_grid(size_t n, int block) {
    size_t g = (n + block - 1) / block; size_t cap = static_cast<size_t>(st_num_sms()) * 8;
    return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

constexpr int IN_MAX_SPLITS = 128;
size_t in_workspace_floats(int N, int C) {
    // stats partials [N][splits][2][C] + (api adds s1, s2 [N][C]) + per-block sums of dx [8 * SMs + N][C]
    // + per-block max |dy_eff| [N][splits] + per-channel-block bounds of |dx| [16]
    return static_cast<size_t>(N) * IN_MAX_SPLITS * 2 * C + static_cast<size_t>(8 * st_num_sms() + N) * C +
           static_cast<size_t>(N) * IN_MAX_SPLITS + 16;
}

int in_forward(const bf16* xh, const bf16* xl, const float* gamma, const float* beta, float eps,
               bf16* yh, bf16* yl, int planes, float* mean, float* rstd, int N, int H, int W, int C, int relu, int up,
               float* ws, cudaStream_t st) {
    const int HW = H * W, splits = in_splits(N, HW);
    const int chunk = (HW + splits - 1) / splits;
    const int lanes = IN_THREADS / (C / 8);
    st_launch(N * splits, IN_THREADS, lanes * 2 * C * sizeof(float), st)(in_stats_kernel<0, 0>, 
        xh, xl, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, HW, W, C, splits, chunk, 0, planes, ws, nullptr);
    ST_CHECK_LAUNCH("in_stats_kernel<0>");
    st_launch(dim3((C + 31) / 32, N), 256, 0, st)(in_finalize_fwd_kernel, ws, C, splits, HW, eps, mean, rstd);
    ST_CHECK_LAUNCH("in_finalize_fwd_kernel");
    const int ablocks = in_apply_blocks(N, HW), achunk = (HW + ablocks - 1) / ablocks;
    if (up) st_launch(dim3(ablocks, N), IN_THREADS, 0, st)(in_apply_fwd_kernel<1>, xh, xl, mean, rstd, gamma, beta, yh, yl, HW, W, C, achunk, relu, planes);
    else st_launch(dim3(ablocks, N), IN_THREADS, 0, st)(in_apply_fwd_kernel<0>, xh, xl, mean, rstd, gamma, beta, yh, yl, HW, W, C, achunk, relu, planes);
    ST_CHECK_LAUNCH("in_apply_fwd_kernel");
    return ST_OK;
}

// dx_code: plane code of dx; dx_slot (required when dx has fp16 planes, else null): receives {s, 1/s}, dx is stored
// multiplied by s (a power of two chosen from a bound of |dx| that the statistics pass measures)
int in_backward(const bf16* xh, const bf16* xl, const bf16* gh, const bf16* gl,
                const float* mean, const float* rstd, const float* gamma, const float* beta,
                bf16* dxh, bf16* dxl, int dx_code, float* dx_slot, float* dgamma, float* dbeta, float* dsum, int x_planes,
                int N, int H, int W, int C, int relu, int up, float* ws, cudaStream_t st) {
    const int HW = H * W, splits = in_splits(N, HW);
    const int chunk = (HW + splits - 1) / splits;
    const int lanes = IN_THREADS / (C / 8);
    float* s1 = ws + static_cast<size_t>(N) * splits * 2 * C;
    float* s2 = s1 + static_cast<size_t>(N) * C;
    float* gmaxp = ws + static_cast<size_t>(N) * IN_MAX_SPLITS * 2 * C + static_cast<size_t>(8 * st_num_sms() + N) * C + 2 * static_cast<size_t>(N) * C;
    float* boundp = gmaxp + static_cast<size_t>(N) * IN_MAX_SPLITS;
    float* gmax_arg = dx_slot ? gmaxp : nullptr;
    if (up)
        st_launch(N * splits, IN_THREADS, lanes * 2 * C * sizeof(float), st)(in_stats_kernel<1, 1>, 
            xh, xl, gh, gl, mean, rstd, gamma, beta, HW, W, C, splits, chunk, relu, x_planes, ws, gmax_arg);
    else
        st_launch(N * splits, IN_THREADS, lanes * 2 * C * sizeof(float), st)(in_stats_kernel<1, 0>, 
            xh, xl, gh, gl, mean, rstd, gamma, beta, HW, W, C, splits, chunk, relu, x_planes, ws, gmax_arg);
    ST_CHECK_LAUNCH("in_stats_kernel<1>");
    const int nbound = (C + 31) / 32;
    st_launch(nbound, 1024, 0, st)(in_finalize_bwd_kernel, ws, N, C, splits, s1, s2, dgamma, dbeta, gmaxp, N * splits, gamma, rstd,
                                                    1.0f / HW, dx_slot ? boundp : nullptr);
    ST_CHECK_LAUNCH("in_finalize_bwd_kernel");
    const int ablocks = in_apply_blocks(N, HW), achunk = (HW + ablocks - 1) / ablocks;
    float* dpart = dsum ? s2 + static_cast<size_t>(N) * C : nullptr;        // [N * ablocks][C]
    const size_t dsm = dsum ? static_cast<size_t>(IN_THREADS / (C / 8)) * C * sizeof(float) : 0;
    if (up) st_launch(dim3(ablocks, N), IN_THREADS, dsm, st)(in_apply_bwd_kernel<1>, xh, xl, gh, gl, mean, rstd, gamma, beta, s1, s2, dxh, dxl, dx_code, boundp, nbound, dx_slot, HW, W, C, achunk, relu, x_planes, dpart);
    else st_launch(dim3(ablocks, N), IN_THREADS, dsm, st)(in_apply_bwd_kernel<0>, xh, xl, gh, gl, mean, rstd, gamma, beta, s1, s2, dxh, dxl, dx_code, boundp, nbound, dx_slot, HW, W, C, achunk, relu, x_planes, dpart);
    ST_CHECK_LAUNCH("in_apply_bwd_kernel");
    if (dsum) {
        st_launch((C + 7) / 8, 1024, 0, st)(in_colsum_kernel, dpart, N * ablocks, C, dsum);
        ST_CHECK_LAUNCH("in_colsum_kernel");
    }
    return ST_OK;
}

int tv_forward(const float* x, size_t total, int H, int W, float factor, float* loss, float* ws, cudaStream_t st) {
    int grid = ew_grid(total, 256); if (grid > 2048) grid = 2048;
    st_launch(grid, 256, 0, st)(tv_partial_kernel, x, total, H, W, ws);
    ST_CHECK_LAUNCH("tv_partial_kernel");
    st_launch(1, 256, 0, st)(finish_scalar_kernel, ws, grid, factor, loss);
    ST_CHECK_LAUNCH("finish_scalar_kernel");
    return ST_OK;
}
int tv_backward(const float* x, size_t total, int H, int W, float factor, const float* gw, float* g, cudaStream_t st) {
    st_launch(ew_grid(total, 256), 256, 0, st)(tv_bwd_kernel, x, total, H, W, factor, gw, g);
    ST_CHECK_LAUNCH("tv_bwd_kernel");
    return ST_OK;
}
int temporal_forward(const float* oc, const float* os, const float* cc, const float* cs, size_t total, float tw,
                     float* out3, float* ws, cudaStream_t st) {
    int grid = ew_grid(total, 256); if (grid > 1024) grid = 1024;
    st_launch(grid, 256, 0, st)(temporal_partial_kernel, oc, os, cc, cs, total, ws);
    ST_CHECK_LAUNCH("temporal_partial_kernel");
    st_launch(1, 256, 0, st)(temporal_finish_kernel, ws, grid, tw, out3);
    ST_CHECK_LAUNCH("temporal_finish_kernel");
    return ST_OK;
}
int temporal_warp_forward(const float* oc, const float* os, const float* cc, const float* cs, const float* flow, int N, int C,
                          int H, int W, float tw, float* out3, float* ws, cudaStream_t st) {
    int grid = ew_grid((size_t)N * H * W, 256); if (grid > 1024) grid = 1024;
    st_launch(grid, 256, 0, st)(temporal_warp_partial_kernel, oc, os, cc, cs, flow, N, C, H, W, ws);
    ST_CHECK_LAUNCH("temporal_warp_partial_kernel");
    st_launch(1, 256, 0, st)(temporal_finish_kernel, ws, grid, tw, out3);
    ST_CHECK_LAUNCH("temporal_finish_kernel");
    return ST_OK;
}
int temporal_warp_backward(const float* os, const float* cs, const float* flow, int N, int C, int H, int W, float tw,
                           const float* stats, const float* gw, float* g, cudaStream_t st) {
    st_launch(ew_grid((size_t)N * H * W, 256), 256, 0, st)(temporal_warp_bwd_kernel, os, cs, flow, N, C, H, W, tw, stats, gw, g);
    ST_CHECK_LAUNCH("temporal_warp_bwd_kernel");
    return ST_OK;
}
int temporal_backward(const float* os, const float* cs, size_t total, float tw, const float* stats, const float* gw,
                      float* g, cudaStream_t st) {
    st_launch(ew_grid(total, 256), 256, 0, st)(temporal_bwd_kernel, os, cs, total, tw, stats, gw, g);
    ST_CHECK_LAUNCH("temporal_bwd_kernel");
    return ST_OK;
}
