// Optimizer steps of the two hot loops, on the device.
//
//  * st_adam_step: torch.optim.Adam over ALL parameter tensors of a transform net in one launch
//    (reference network.py:643-649, :765: `optimizer.step(closure)` with optim.Adam defaults).
//    Parameters, gradients and both moment buffers are flat fp32 arrays with one segment per
//    parameter tensor; each segment has its own step counter and an active flag, which is
//    torch's per-parameter state (a frozen tensor -- video_train epoch 0, network.py:927-934 --
//    keeps its moments and its step count untouched).
//
//  * st_lbfgs_iteration: one iteration of torch.optim.LBFGS (reference network.py:435, :456:
//    defaults lr 1, max_iter 20, history 100, no line search) in COEFFICIENT SPACE.  torch walks
//    the history twice per iteration with one dot kernel and one axpy kernel per stored pair
//    (400 launches + host round trips at history 100).  Here every vector of the two-loop
//    recursion is a combination  c_g*g + sum_j cy_j*y_j + sum_j cs_j*s_j, so
//      pass A (lbfgs_dots_kernel):   one sweep over the history gives s_j.g, y_j.g, g.g, g.g_prev, g.d;
//      solve  (lbfgs_solve_kernel):  one CTA keeps SY[i][j] = s_i.y_j and YY[i][j] = y_i.y_j in shared
//                                    memory, appends the new pair's row/column from DIFFERENCES of the
//                                    dot vectors of two consecutive gradients (y_new = g - g_prev,
//                                    s_new = t*d), runs both loops on coefficients in fp64 and applies
//                                    torch's stopping rules to a device-side halt flag;
//      pass B (lbfgs_update_kernel): second sweep: writes the new pair into the ring, forms the
//                                    direction, moves x, keeps g as g_prev, max|t*d| by atomicMax.
//    HBM traffic = 2 sweeps of the 2*k history vectors (the same bytes as torch's two loops) in
//    3 launches and no host synchronisation.  All dots are accumulated in fp64 (the products of
//    fp32 values are exact there), which is what makes the differences of dots safe.
//    oracle/lbfgs_model.py is the CPU restatement of this file, pinned against torch.optim.LBFGS.
#include "common.cuh"
#include "../../include/stransfer_b200.h"

// ============================================================================ Adam
struct AdamChunk { unsigned long long off; unsigned int n; int seg; };
#define ADAM_CHUNK 2048
// Optional per-segment description of the PACKED copies of a convolution weight (st_pack_conv_weight layouts) that
// the Adam kernel rewrites together with the fp32 master value, so that no separate packing launch is needed after
// an optimizer step.  88 bytes per segment; ndst = 0 for tensors without packed copies.
struct AdamPackDst { void* hi; void* lo; int rows_pad, cols_pad, mode, code; };
struct AdamPackSeg { long long seg_off; int cout, cin, ks, ndst; AdamPackDst dst[2]; };
static_assert(sizeof(AdamPackSeg) == 88, "AdamPackSeg layout is part of the ABI (optim.py builds it)");

__device__ __forceinline__ void adam_store_packed(const AdamPackSeg& ps, long long idx, float value) {
    const int taps = ps.ks * ps.ks;
    long long q = idx;
    const int s = (int)(q % ps.ks); q /= ps.ks;
    const int r = (int)(q % ps.ks); q /= ps.ks;
    const int ci = (int)(q % ps.cin); const int co = (int)(q / ps.cin);
    for (int d = 0; d < ps.ndst; ++d) {
        const AdamPackDst& D = ps.dst[d];
        int row, col, tap;
        if (D.mode == 0) { row = co; col = ci; tap = r * ps.ks + s; }
        else { row = ci; col = co; tap = (D.mode == 2) ? (ps.ks - 1 - r) * ps.ks + (ps.ks - 1 - s) : r * ps.ks + s; }
        const size_t o = (static_cast<size_t>(row) * taps + tap) * D.cols_pad + col;
        uint16_t* hi = static_cast<uint16_t*>(D.hi);
        uint16_t* lo = static_cast<uint16_t*>(D.lo);
        if (pc_f16(D.code)) {
            uint32_t h2, l2; split2h(value * ST_W_F16_SCALE, 0.f, h2, l2);
            hi[o] = static_cast<uint16_t>(h2 & 0xffffu);
            if (lo && pc_planes(D.code) > 1) lo[o] = static_cast<uint16_t>(l2 & 0xffffu);
        } else {
            bf16 h, l; split_bf16(value, h, l);
            hi[o] = __bfloat16_as_ushort(h);
            if (lo && pc_planes(D.code) > 1) lo[o] = __bfloat16_as_ushort(l);
            if (lo && pc_planes(D.code) > 2)
                (lo + (lo - hi))[o] = __bfloat16_as_ushort(__float2bfloat16_rn(value - __bfloat162float(h) - __bfloat162float(l)));
        }
    }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, const AdamChunk* __restrict__ chunks,
                                                   const int* __restrict__ steps, const int* __restrict__ active,
                                                   double lr, double b1d, double b2d, float omb1, float omb2, float eps, float wd,
                                                   float gscale, const AdamPackSeg* __restrict__ packs) {
    pdl_prologue();
    const AdamChunk c = chunks[blockIdx.x];
    if (!active[c.seg]) return;
    const bool repack = packs && packs[c.seg].ndst > 0;
    __shared__ float s_step_size, s_rsqrt_bc2;
    if (threadIdx.x == 0) {
        const double t = (double)(steps[c.seg] + 1);
        const double bc1 = 1.0 - pow(b1d, t), bc2 = 1.0 - pow(b2d, t);
        s_step_size = (float)(lr / bc1);
        s_rsqrt_bc2 = (float)sqrt(bc2);
    }
    __syncthreads();
    const float step_size = s_step_size, bc2_sqrt = s_rsqrt_bc2, b2 = (float)b2d;
    for (unsigned int i = threadIdx.x; i < c.n; i += 256) {
        const size_t e = c.off + i;
        float gr = g[e] * gscale;
        const float pe = p[e];
        if (wd != 0.f) gr = fmaf(wd, pe, gr);
        // exp_avg.lerp_(grad, 1-b1); exp_avg_sq.mul_(b2).addcmul_(grad, grad, 1-b2)
        const float me = m[e] + (gr - m[e]) * omb1;
        const float ve = fmaf(gr * omb2, gr, v[e] * b2);
        const float denom = __fdiv_rn(__fsqrt_rn(ve), bc2_sqrt) + eps;
        m[e] = me; v[e] = ve;
        const float pn = pe - step_size * __fdiv_rn(me, denom);
        p[e] = pn;
        if (repack) adam_store_packed(packs[c.seg], (long long)e - packs[c.seg].seg_off, pn);
    }
}
__global__ void adam_tick_kernel(int* steps, const int* active, int nseg) {
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nseg && active[i]) steps[i] += 1;
}

extern "C" size_t st_adam_chunk_count(const long long* seg_numel, int nseg) {
    size_t c = 0;
    for (int i = 0; i < nseg; ++i) c += (size_t)((seg_numel[i] + ADAM_CHUNK - 1) / ADAM_CHUNK);
    return c;
}
extern "C" int st_adam_fill_chunks(const long long* seg_numel, int nseg, void* host_chunks) {
    ST_CHECK_ARG(seg_numel && host_chunks && nseg > 0, "st_adam_fill_chunks: bad arguments");
    AdamChunk* out = (AdamChunk*)host_chunks;
    unsigned long long off = 0; size_t c = 0;
    for (int i = 0; i < nseg; ++i) {
        for (long long s = 0; s < seg_numel[i]; s += ADAM_CHUNK) {
            const long long n = seg_numel[i] - s < ADAM_CHUNK ? seg_numel[i] - s : ADAM_CHUNK;
            out[c].off = off + (unsigned long long)s; out[c].n = (unsigned int)n; out[c].seg = i; ++c;
        }
        off += (unsigned long long)seg_numel[i];
    }
    return ST_OK;
}
extern "C" int st_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, const void* chunks,
                            size_t nchunks, int* steps, const int* active, int nseg, double lr, double beta1, double beta2,
                            double eps, double weight_decay, double grad_scale, const void* packs, void* stream) {
    ST_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && chunks && steps && active && nchunks > 0 && nseg > 0,
                 "st_adam_step: null pointer or empty table");
    cudaStream_t st = (cudaStream_t)stream;
    st_launch((unsigned)nchunks, 256, 0, st)(adam_kernel, param, grad, exp_avg, exp_avg_sq, (const AdamChunk*)chunks, steps, active,
                                                    lr, beta1, beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps,
                                                    (float)weight_decay, (float)grad_scale, (const AdamPackSeg*)packs);
    ST_CHECK_LAUNCH("adam_kernel");
    st_launch((nseg + 127) / 128, 128, 0, st)(adam_tick_kernel, steps, active, nseg);
    ST_CHECK_LAUNCH("adam_tick_kernel");
    return ST_OK;
}

// ============================================================================ L-BFGS
enum { H_K = 0, H_HEAD, H_NITER, H_T, H_HDIAG, H_DGPREV, H_GGPREV, H_HALT, H_SKIP, H_LOSSPREV, H_PUSHED, H_NEW, H_CG,
       H_TOLD, H_RUNB, H_LR, H_TOLG, H_TOLC, H_M, H_GTD, H_GMAX, H_DMAX, H_LOSS, H_DMAXBITS = 31, H_SIZE = 32 };
#define LB_MAX_M 100
#define LB_EPB 2048                      // elements per block of the dots pass (256 threads x 2 float4)
#define LB_SL(m) (2 * (m) + 8)

struct LbfgsState {                      // offsets in doubles
    int m;
    __host__ __device__ size_t ro() const { return H_SIZE; }
    __host__ __device__ size_t cy() const { return ro() + m; }
    __host__ __device__ size_t cs() const { return cy() + m; }
    __host__ __device__ size_t sgp() const { return cs() + m; }
    __host__ __device__ size_t ygp() const { return sgp() + m; }
    __host__ __device__ size_t SY() const { return ygp() + m; }
    __host__ __device__ size_t YY() const { return SY() + (size_t)m * m; }
    __host__ __device__ size_t partials() const { return YY() + (size_t)m * m; }
};
static inline size_t lb_chunks(size_t n) { return (n + LB_EPB - 1) / LB_EPB; }
// chunks per CTA so that the dots pass runs at most 3 CTAs per SM (all resident at 80 registers), evenly loaded
static int g_lbfgs_variant = 0;          // 0: 3 CTAs per SM, loads issued per slot pair; 1: 2 CTAs per SM, loads software-pipelined
extern "C" int st_set_lbfgs_variant(int v) { const int old = g_lbfgs_variant; g_lbfgs_variant = v ? 1 : 0; return old; }
static inline size_t lb_cpb(size_t n) { const size_t lim = 3 * (size_t)st_num_sms(); return (lb_chunks(n) + lim - 1) / lim; }
static inline size_t lb_blocks(size_t n) { const size_t c = lb_cpb(n); return (lb_chunks(n) + c - 1) / c; }

__device__ __forceinline__ double shfl_xor_d(double v, int o) {
    int lo = __double2loint(v), hi = __double2hiint(v);
    lo = __shfl_xor_sync(0xffffffffu, lo, o); hi = __shfl_xor_sync(0xffffffffu, hi, o);
    return __hiloint2double(hi, lo);
}
__device__ __forceinline__ double shfl_idx_d(double v, int src) {
    int lo = __double2loint(v), hi = __double2hiint(v);
    lo = __shfl_sync(0xffffffffu, lo, src); hi = __shfl_sync(0xffffffffu, hi, src);
    return __hiloint2double(hi, lo);
}
// sums of 4 values over the warp with 6 fp64 shuffles; lanes 0/8/16/24 end up holding slots 0/1/2/3
__device__ __forceinline__ double warp_sum4(const double (&a)[4], int lane) {
    const bool h4 = lane & 16, h3 = lane & 8;
    const double r0 = shfl_xor_d(h4 ? a[0] : a[2], 16), r1 = shfl_xor_d(h4 ? a[1] : a[3], 16);
    const double b0 = (h4 ? a[2] : a[0]) + r0, b1 = (h4 ? a[3] : a[1]) + r1;
    double c = (h3 ? b1 : b0) + shfl_xor_d(h3 ? b0 : b1, 8);
    c += shfl_xor_d(c, 4); c += shfl_xor_d(c, 2); c += shfl_xor_d(c, 1);
    return c;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += shfl_xor_d(v, o);
    return v;
}
// 4 consecutive floats at element e (zeros beyond n); vector load when whole and aligned (ld % 4 == 0)
__device__ __forceinline__ float4 ld4g(const float* __restrict__ p, size_t e, size_t n) {
    if (e + 3 < n) return *reinterpret_cast<const float4*>(p + e);
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (e < n) r.x = p[e];
    if (e + 1 < n) r.y = p[e + 1];
    if (e + 2 < n) r.z = p[e + 2];
    return r;
}
__device__ __forceinline__ void st4g(float* __restrict__ p, size_t e, size_t n, float4 v) {
    if (e + 3 < n) { *reinterpret_cast<float4*>(p + e) = v; return; }
    if (e < n) p[e] = v.x;
    if (e + 1 < n) p[e + 1] = v.y;
    if (e + 2 < n) p[e + 2] = v.z;
}
__device__ __forceinline__ double dot4d(const double (&g)[4], float4 v) {
    return g[0] * (double)v.x + g[1] * (double)v.y + g[2] * (double)v.z + g[3] * (double)v.w;
}

// ---- pass A: partial dots of the gradient with the history, the previous gradient and the direction.
// Each CTA walks `chunks_per_block` chunks of LB_EPB elements and keeps its per-warp sums in shared memory,
// so the solve kernel has gridDim.x (<= 3 x SMs) rows to add up.
template <bool PIPE>
__global__ void __launch_bounds__(256, PIPE ? 2 : 3) lbfgs_dots_kernel(const float* __restrict__ g, const float* __restrict__ gp,
                                                         const float* __restrict__ d, const float* __restrict__ S,
                                                         const float* __restrict__ Y, size_t n, size_t ld, int m,
                                                         const double* __restrict__ st, int first, int chunks_per_block,
                                                         double* __restrict__ partials) {
    pdl_prologue();
    if (!first && st[H_HALT] != 0.0) return;
    const int k = (int)st[H_K];
    const int SL = LB_SL(m);
    __shared__ double red[8][LB_SL(LB_MAX_M)];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int c = 0; c < chunks_per_block; ++c) {
        const size_t e0 = ((size_t)blockIdx.x * chunks_per_block + c) * LB_EPB + (size_t)tid * 4, e1 = e0 + 1024;
        const bool more = c > 0;
        const float4 ga = ld4g(g, e0, n), gb = ld4g(g, e1, n);
        const double g0[4] = {ga.x, ga.y, ga.z, ga.w}, g1[4] = {gb.x, gb.y, gb.z, gb.w};
        // software pipeline: the 8 loads of the next slot pair are in flight while this pair is reduced
        float4 nx[8];
        auto fetch = [&](int j) {
            const int j1 = j + 1 < k ? j + 1 : j;
            const float* s0 = S + (size_t)j * ld; const float* s1 = S + (size_t)j1 * ld;
            const float* y0 = Y + (size_t)j * ld; const float* y1 = Y + (size_t)j1 * ld;
            nx[0] = ld4g(s0, e0, n); nx[1] = ld4g(s0, e1, n); nx[2] = ld4g(s1, e0, n); nx[3] = ld4g(s1, e1, n);
            nx[4] = ld4g(y0, e0, n); nx[5] = ld4g(y0, e1, n); nx[6] = ld4g(y1, e0, n); nx[7] = ld4g(y1, e1, n);
        };
        if (PIPE && k > 0) fetch(0);
        for (int j = 0; j < k; j += 2) {
            double a[4];
            if constexpr (PIPE) {
                float4 cu[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) cu[i] = nx[i];
                if (j + 2 < k) fetch(j + 2);
                a[0] = dot4d(g0, cu[0]) + dot4d(g1, cu[1]); a[1] = dot4d(g0, cu[2]) + dot4d(g1, cu[3]);
                a[2] = dot4d(g0, cu[4]) + dot4d(g1, cu[5]); a[3] = dot4d(g0, cu[6]) + dot4d(g1, cu[7]);
            } else {
                const int j1 = j + 1 < k ? j + 1 : j;
                const float* s0 = S + (size_t)j * ld; const float* s1 = S + (size_t)j1 * ld;
                const float* y0 = Y + (size_t)j * ld; const float* y1 = Y + (size_t)j1 * ld;
                const float4 a0 = ld4g(s0, e0, n), a1 = ld4g(s0, e1, n), b0 = ld4g(s1, e0, n), b1 = ld4g(s1, e1, n);
                const float4 c0 = ld4g(y0, e0, n), c1 = ld4g(y0, e1, n), d0 = ld4g(y1, e0, n), d1 = ld4g(y1, e1, n);
                a[0] = dot4d(g0, a0) + dot4d(g1, a1); a[1] = dot4d(g0, b0) + dot4d(g1, b1);
                a[2] = dot4d(g0, c0) + dot4d(g1, c1); a[3] = dot4d(g0, d0) + dot4d(g1, d1);
            }
            const double r = warp_sum4(a, lane);
            if ((lane & 7) == 0) {
                const int q = lane >> 3;                       // 0: s_j  1: s_j1  2: y_j  3: y_j1
                if (!(q & 1) || j + 1 < k) {
                    double* dst = &red[w][(q >> 1) * m + j + (q & 1)];
                    *dst = more ? *dst + r : r;
                }
            }
        }
        const float4 p0 = ld4g(gp, e0, n), p1 = ld4g(gp, e1, n), q0 = ld4g(d, e0, n), q1 = ld4g(d, e1, n);
        double sabs = 0.0; float mx = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) { sabs += fabs(g0[i]) + fabs(g1[i]); mx = fmaxf(mx, fmaxf(fabsf((float)g0[i]), fabsf((float)g1[i]))); }
        const double a[4] = {dot4d(g0, ga) + dot4d(g1, gb), dot4d(g0, p0) + dot4d(g1, p1), dot4d(g0, q0) + dot4d(g1, q1), sabs};
        const double r = warp_sum4(a, lane);
        if ((lane & 7) == 0) { double* dst = &red[w][2 * m + (lane >> 3)]; *dst = more ? *dst + r : r; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) { double* dst = &red[w][2 * m + 4]; *dst = more ? fmax(*dst, (double)mx) : (double)mx; }
    }
    __syncthreads();
    for (int s = tid; s < SL; s += 256) {
        const bool valid = (s < m) ? s < k : (s < 2 * m ? s - m < k : s <= 2 * m + 4);
        double acc = 0.0;
        if (valid) {
            if (s == 2 * m + 4) { for (int i = 0; i < 8; ++i) acc = fmax(acc, red[i][s]); }
            else { for (int i = 0; i < 8; ++i) acc += red[i][s]; }
        }
        partials[(size_t)blockIdx.x * SL + s] = acc;
    }
}

// ---- solve: history bookkeeping + two-loop recursion on coefficients, one CTA
__global__ void __launch_bounds__(1024) lbfgs_solve_kernel(double* __restrict__ st, int m, size_t nblocks, const float* __restrict__ loss_ptr,
                                                           int first) {
    pdl_prologue();
    extern __shared__ double sm[];
    const LbfgsState L{m};
    const int tid = threadIdx.x, lane = tid & 31;
    const int SL = LB_SL(m);
    if (!first && st[H_HALT] != 0.0) { if (tid == 0) st[H_RUNB] = 0.0; return; }
    const int LDM = m | 1;                     // odd row stride (in doubles): column walks of SY stay off a single bank
    double* SY = sm; double* YY = SY + m * LDM;
    double* dots = YY + m * LDM;               // SL
    double* grp = dots + SL;                   // 4 * SL
    double* qy = grp + 4 * SL; double* rs = qy + m; double* al = rs + m; double* ro = al + m;
    double* cy = ro + m; double* cs = cy + m; double* sgp = cs + m; double* ygp = sgp + m;
    double* row_sy = ygp + m; double* sc = row_sy + m;     // sc: 16 scalars
    __shared__ int s_k, s_head, s_pushed, s_new, s_halt, s_kold;
    const int k = (int)st[H_K], head = (int)st[H_HEAD];
    const double* partials = st + L.partials();
    // phase 1: reduce the per-block partials in a fixed order (4 groups of 256 threads, then the groups)
    {
        const int grpi = tid >> 8, s = tid & 255;
        if (s < SL) {
            const bool is_max = s == 2 * m + 4;
            double acc = 0.0;
            const size_t per = (nblocks + 3) / 4, b0 = grpi * per, b1 = b0 + per < nblocks ? b0 + per : nblocks;
            size_t b = b0;
            for (; b + 8 <= b1; b += 8) {
                double v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = partials[(b + u) * SL + s];
                if (is_max) acc = fmax(acc, fmax(fmax(fmax(v[0], v[1]), fmax(v[2], v[3])), fmax(fmax(v[4], v[5]), fmax(v[6], v[7]))));
                else acc += ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
            }
            for (; b < b1; ++b) { const double v = partials[b * SL + s]; if (is_max) acc = fmax(acc, v); else acc += v; }
            grp[grpi * SL + s] = acc;
        }
    }
    // phase 2: history matrices and vectors into shared memory
    for (int idx = tid; idx < k * k; idx += 1024) {
        const int i = idx / k, j = idx - i * k;
        SY[i * LDM + j] = st[L.SY() + (size_t)i * m + j];
        YY[i * LDM + j] = st[L.YY() + (size_t)i * m + j];
    }
    if (tid < m) {
        ro[tid] = st[L.ro() + tid]; cy[tid] = st[L.cy() + tid]; cs[tid] = st[L.cs() + tid];
        sgp[tid] = st[L.sgp() + tid]; ygp[tid] = st[L.ygp() + tid];
        qy[tid] = 0.0; rs[tid] = 0.0; al[tid] = 0.0;
    }
    __syncthreads();
    if (tid < SL) {
        const double a = grp[tid], b = grp[SL + tid], c = grp[2 * SL + tid], d = grp[3 * SL + tid];
        dots[tid] = (tid == 2 * m + 4) ? fmax(fmax(a, b), fmax(c, d)) : (a + b) + (c + d);
    }
    __syncthreads();
    double* sg = dots; double* yg = dots + m;
    const double gg = dots[2 * m], ggp = dots[2 * m + 1], dg = dots[2 * m + 2], gabs = dots[2 * m + 3], gmax = dots[2 * m + 4];
    const double tolg = st[H_TOLG], tolc = st[H_TOLC], lr = st[H_LR], t_old = st[H_T];
    const double loss = (double)*loss_ptr;
    if (tid == 0) {
        int halt = gmax <= tolg;
        if (!first) {
            const double dmax = (double)__uint_as_float(*reinterpret_cast<const unsigned int*>(st + H_DMAXBITS));
            if (dmax <= tolc || fabs(loss - st[H_LOSSPREV]) < tolc) halt = 1;
        }
        s_halt = halt; s_pushed = 0; s_new = -1; s_k = k; s_head = head; s_kold = k;
        st[H_GMAX] = gmax; st[H_LOSS] = loss;
        if (halt) { st[H_HALT] = 1.0; st[H_RUNB] = 0.0; }
        else {
            const double niter = st[H_NITER] + 1.0;
            st[H_NITER] = niter;
            sc[0] = niter;
            if (niter > 1.0) {
                const double ys = t_old * (dg - st[H_DGPREV]);
                const double yy = gg - 2.0 * ggp + st[H_GGPREV];
                sc[1] = ys; sc[2] = yy;
                if (ys > 1e-10) {
                    s_pushed = 1;
                    if (k == m) { s_new = head; s_head = (head + 1) % m; }
                    else { s_new = (head + k) % m; s_k = k + 1; }
                }
            }
        }
    }
    __syncthreads();
    if (s_halt) return;
    const double niter = sc[0];
    const int pushed = s_pushed, nw = s_new, kn = s_k, headn = s_head;
    double H = st[H_HDIAG];
    if (pushed) {
        const double cg_old = st[H_CG];
        // s_new . y_j = t_old * d . y_j, d in coefficient form (rows of a pair about to be dropped are still valid)
        if (tid < k) {
            const int j = tid;
            double acc = cg_old * ygp[j];
            for (int i = 0; i < k; ++i) acc += cy[i] * YY[i * LDM + j] + cs[i] * SY[i * LDM + j];
            row_sy[j] = t_old * acc;
        }
        __syncthreads();
        if (tid < k && tid != nw) {
            const int j = tid;
            const double r = row_sy[j], c = sg[j] - sgp[j], y = yg[j] - ygp[j];
            SY[nw * LDM + j] = r; SY[j * LDM + nw] = c; YY[nw * LDM + j] = y; YY[j * LDM + nw] = y;
            st[L.SY() + (size_t)nw * m + j] = r; st[L.SY() + (size_t)j * m + nw] = c;
            st[L.YY() + (size_t)nw * m + j] = y; st[L.YY() + (size_t)j * m + nw] = y;
        }
        __syncthreads();
        if (tid == 0) {
            const double ys = sc[1], yy = sc[2];
            SY[nw * LDM + nw] = ys; YY[nw * LDM + nw] = yy;
            st[L.SY() + (size_t)nw * m + nw] = ys; st[L.YY() + (size_t)nw * m + nw] = yy;
            ro[nw] = 1.0 / ys; st[L.ro() + nw] = 1.0 / ys;
            sg[nw] = t_old * dg; yg[nw] = gg - ggp;
            st[H_HDIAG] = ys / yy;
        }
        __syncthreads();
        H = sc[1] / sc[2];
    }
    double cg;
    if (niter == 1.0) {
        cg = -1.0;
        if (tid < m) { cy[tid] = 0.0; cs[tid] = 0.0; }
        if (tid == 0) st[H_HDIAG] = 1.0;
    } else {
        const double qg = -1.0;
        // Two-loop recursion without reductions: u[j] = s_j.q and w[j] = y_j.q are kept for ALL j and updated by
        // one axpy per step (q -= al_i*y_i  =>  u[j] -= al_i*SY[j][i], w[j] -= al_i*YY[j][i]); the second loop
        // keeps v[j] = y_j.r the same way (r += c*s_i  =>  v[j] += c*SY[i][j]).  One warp, k/32 entries per lane.
        static_assert(LB_MAX_M <= 128, "four register entries per lane");
        if (tid < 32) {
            // entries j = lane + 32*t of u / w / v live in registers; the pivot entry is fetched with one shuffle
            double u[4], w[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int j = lane + 32 * t;
                u[t] = j < kn ? qg * sg[j] : 0.0; w[t] = j < kn ? qg * yg[j] : 0.0;
            }
            int i = (headn + kn - 1) % m;
            for (int idx = kn - 1; idx >= 0; --idx) {
                const int t = i >> 5;
                const double ui = shfl_idx_d(t == 0 ? u[0] : t == 1 ? u[1] : t == 2 ? u[2] : u[3], i & 31);
                const double a = ro[i] * ui;
#pragma unroll
                for (int t2 = 0; t2 < 4; ++t2) {
                    const int j = lane + 32 * t2;
                    if (j < kn) { u[t2] -= a * SY[j * LDM + i]; w[t2] -= a * YY[i * LDM + j]; }
                }
                if (lane == 0) qy[i] = -a;
                i = i == 0 ? m - 1 : i - 1;
            }
            __syncwarp();
#pragma unroll
            for (int t = 0; t < 4; ++t) w[t] *= H;                       // v[j] = y_j . (H q)
            i = headn;
            for (int idx = 0; idx < kn; ++idx) {
                const int t = i >> 5;
                const double vi = shfl_idx_d(t == 0 ? w[0] : t == 1 ? w[1] : t == 2 ? w[2] : w[3], i & 31);
                const double c = -qy[i] - ro[i] * vi;
#pragma unroll
                for (int t2 = 0; t2 < 4; ++t2) {
                    const int j = lane + 32 * t2;
                    if (j < kn) w[t2] += c * SY[i * LDM + j];
                }
                if (lane == 0) cs[i] = c;
                i = i + 1 == m ? 0 : i + 1;
            }
        }
        __syncthreads();
        cg = H * qg;
        if (tid < m) { cy[tid] = H * qy[tid]; if (tid >= kn) cs[tid] = 0.0; }
    }
    __syncthreads();
    if (tid < m) {
        st[L.cy() + tid] = cy[tid]; st[L.cs() + tid] = cs[tid];
        st[L.sgp() + tid] = tid < kn ? sg[tid] : 0.0; st[L.ygp() + tid] = tid < kn ? yg[tid] : 0.0;
    }
    if (tid < 32) {
        double p = 0.0;
        for (int j = lane; j < kn; j += 32) p += cy[j] * yg[j] + cs[j] * sg[j];
        p = warp_sum_d(p);
        if (lane == 0) {
            const double gtd = cg * gg + p;
            const double t = niter == 1.0 ? fmin(1.0, 1.0 / gabs) * lr : lr;
            const int skip = gtd > -tolc;
            st[H_K] = (double)kn; st[H_HEAD] = (double)headn; st[H_TOLD] = t_old; st[H_T] = t; st[H_DGPREV] = gtd; st[H_GTD] = gtd;
            st[H_GGPREV] = gg; st[H_SKIP] = (double)skip; st[H_LOSSPREV] = loss; st[H_PUSHED] = (double)pushed;
            st[H_NEW] = (double)nw; st[H_CG] = cg; st[H_RUNB] = 1.0; st[H_HALT] = skip ? 1.0 : 0.0;
            *reinterpret_cast<unsigned int*>(st + H_DMAXBITS) = 0u;
        }
    }
}

// ---- pass B: new pair into the ring, direction, x += t*d, g_prev = g
__global__ void __launch_bounds__(256, 3) lbfgs_update_kernel(float* __restrict__ x, const float* __restrict__ g, float* __restrict__ gp,
                                                           float* __restrict__ d, float* __restrict__ S, float* __restrict__ Y,
                                                           size_t n, size_t ld, int m, double* __restrict__ st) {
    pdl_prologue();
    if (st[H_RUNB] == 0.0) return;
    const LbfgsState L{m};
    __shared__ double cy[LB_MAX_M], cs[LB_MAX_M];
    __shared__ float red[8];
    const int tid = threadIdx.x;
    const int k = (int)st[H_K], pushed = (int)st[H_PUSHED], nw = pushed ? (int)st[H_NEW] : -1, skip = (int)st[H_SKIP];
    const double cg = st[H_CG], t = st[H_T];
    const float t_old = (float)st[H_TOLD];
    if (tid < m) { cy[tid] = st[L.cy() + tid]; cs[tid] = st[L.cs() + tid]; }
    __syncthreads();
    const size_t e = ((size_t)blockIdx.x * 256 + tid) * 4;
    float mx = 0.f;
    if (e < n) {
        const float4 g4 = ld4g(g, e, n), p4 = ld4g(gp, e, n), d4 = ld4g(d, e, n);
        const float4 yn = make_float4(g4.x - p4.x, g4.y - p4.y, g4.z - p4.z, g4.w - p4.w);
        const float4 sn = make_float4(d4.x * t_old, d4.y * t_old, d4.z * t_old, d4.w * t_old);
        double acc[4] = {cg * (double)g4.x, cg * (double)g4.y, cg * (double)g4.z, cg * (double)g4.w};
        auto add = [&](double a, float4 v) {
            acc[0] += a * (double)v.x; acc[1] += a * (double)v.y; acc[2] += a * (double)v.z; acc[3] += a * (double)v.w;
        };
        int j = 0;
        for (; j + 4 <= k; j += 4) {                   // 8 independent 16-byte loads in flight per thread
            float4 yv[4], sv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                yv[u] = (j + u == nw) ? yn : ld4g(Y + (size_t)(j + u) * ld, e, n);
                sv[u] = (j + u == nw) ? sn : ld4g(S + (size_t)(j + u) * ld, e, n);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) { add(cy[j + u], yv[u]); add(cs[j + u], sv[u]); }
        }
        for (; j < k; ++j) {
            float4 y0 = (j == nw) ? yn : ld4g(Y + (size_t)j * ld, e, n), s0 = (j == nw) ? sn : ld4g(S + (size_t)j * ld, e, n);
            add(cy[j], y0); add(cs[j], s0);
        }
        if (pushed) { st4g(Y + (size_t)nw * ld, e, n, yn); st4g(S + (size_t)nw * ld, e, n, sn); }
        const float4 dn = make_float4((float)acc[0], (float)acc[1], (float)acc[2], (float)acc[3]);
        st4g(d, e, n, dn);
        st4g(gp, e, n, g4);
        if (!skip) {
            const float tf = (float)t;
            const float4 x4 = ld4g(x, e, n);
            st4g(x, e, n, make_float4(fmaf(tf, dn.x, x4.x), fmaf(tf, dn.y, x4.y), fmaf(tf, dn.z, x4.z), fmaf(tf, dn.w, x4.w)));
            const float a0 = fabsf(tf * dn.x), a1 = e + 1 < n ? fabsf(tf * dn.y) : 0.f, a2 = e + 2 < n ? fabsf(tf * dn.z) : 0.f,
                        a3 = e + 3 < n ? fabsf(tf * dn.w) : 0.f;
            mx = fmaxf(fmaxf(a0, a1), fmaxf(a2, a3));
        }
    }
    if (!skip) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((tid & 31) == 0) red[tid >> 5] = mx;
        __syncthreads();
        if (tid == 0) {
            for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
            atomicMax(reinterpret_cast<unsigned int*>(st + H_DMAXBITS), __float_as_uint(mx));   // non-negative floats order as uints
        }
    }
}

__global__ void lbfgs_init_kernel(double* st, int m, double lr, double tolg, double tolc) {
    pdl_prologue();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        st[H_LR] = lr; st[H_TOLG] = tolg; st[H_TOLC] = tolc; st[H_M] = (double)m; st[H_HDIAG] = 1.0;
    }
}

static size_t lb_solve_smem(int m) {
    return sizeof(double) * ((size_t)2 * m * (m | 1) + 5 * (size_t)LB_SL(m) + 9 * (size_t)m + 16);
}

extern "C" size_t st_lbfgs_state_bytes(int history, size_t n) {
    if (history <= 0 || history > LB_MAX_M) return 0;
    const LbfgsState L{history};
    return sizeof(double) * (L.partials() + lb_blocks(n) * (size_t)LB_SL(history));
}
extern "C" int st_lbfgs_init(void* state, int history, size_t n, double lr, double tolerance_grad, double tolerance_change, void* stream) {
    ST_CHECK_ARG(state && n > 0, "st_lbfgs_init: null state or empty vector");
    ST_CHECK_ARG(history > 0 && history <= LB_MAX_M, "st_lbfgs_init: history_size %d not in 1..%d", history, LB_MAX_M);
    cudaStream_t s = (cudaStream_t)stream;
    if (cudaMemsetAsync(state, 0, st_lbfgs_state_bytes(history, n), s) != cudaSuccess) {
        st_set_error("st_lbfgs_init: cudaMemsetAsync failed"); cudaGetLastError(); return ST_ERR_CUDA;
    }
    st_launch(1, 32, 0, s)(lbfgs_init_kernel, (double*)state, history, lr, tolerance_grad, tolerance_change);
    ST_CHECK_LAUNCH("lbfgs_init_kernel");
    return ST_OK;
}
extern "C" int st_lbfgs_iteration(float* x, const float* grad, const float* loss, float* g_prev, float* dir, float* S, float* Y,
                                  size_t n, size_t ld, int history, void* state, int first_of_step, void* stream) {
    ST_CHECK_ARG(x && grad && loss && g_prev && dir && S && Y && state && n > 0, "st_lbfgs_iteration: null pointer or empty vector");
    ST_CHECK_ARG(history > 0 && history <= LB_MAX_M, "st_lbfgs_iteration: history_size %d not in 1..%d", history, LB_MAX_M);
    ST_CHECK_ARG(ld >= n && ld % 4 == 0, "st_lbfgs_iteration: history row stride must be >= n and a multiple of 4");
    ST_CHECK_ARG(((uintptr_t)x | (uintptr_t)grad | (uintptr_t)g_prev | (uintptr_t)dir | (uintptr_t)S | (uintptr_t)Y) % 16 == 0,
                 "st_lbfgs_iteration: vectors must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    double* st = (double*)state;
    const LbfgsState L{history};
    const size_t nb = lb_blocks(n);
    static bool attr = false;
    if (!attr) {
        if (cudaFuncSetAttribute(lbfgs_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lb_solve_smem(LB_MAX_M)) != cudaSuccess) {
            st_set_error("st_lbfgs_iteration: cannot reserve %zu bytes of shared memory", lb_solve_smem(LB_MAX_M)); cudaGetLastError();
            return ST_ERR_CUDA;
        }
        attr = true;
    }
    if (g_lbfgs_variant)
        st_launch((unsigned)nb, 256, 0, s)(lbfgs_dots_kernel<true>, grad, g_prev, dir, S, Y, n, ld, history, st, first_of_step, (int)lb_cpb(n),
                                                             st + L.partials());
    else
        st_launch((unsigned)nb, 256, 0, s)(lbfgs_dots_kernel<false>, grad, g_prev, dir, S, Y, n, ld, history, st, first_of_step, (int)lb_cpb(n),
                                                              st + L.partials());
    ST_CHECK_LAUNCH("lbfgs_dots_kernel");
    st_launch(1, 1024, lb_solve_smem(history), s)(lbfgs_solve_kernel, st, history, nb, loss, first_of_step);
    ST_CHECK_LAUNCH("lbfgs_solve_kernel");
    st_launch((unsigned)((n + 1023) / 1024), 256, 0, s)(lbfgs_update_kernel, x, grad, g_prev, dir, S, Y, n, ld, history, st);
    ST_CHECK_LAUNCH("lbfgs_update_kernel");
    return ST_OK;
}
