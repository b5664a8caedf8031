// Gather-based implicit-GEMM convolution (mma.sync bf16, fp32 accumulate) for the
// layer shapes the tcgen05 kernel does not take: tiny channel counts (image
// input padded 3/6 -> 8 channels, 3-channel output), 9x9 kernels, stride 2, the
// nearest-upsample-fused loader, and backward-data of strided convs.
// Reference call sites: network.py:525-551 (9x9 / stride-2 convs), :580-609
// (upsample + conv, final 9x9), VGG conv1_1 (:246-275) and their autograd.
//
// Same split-bf16 NHWC data format and the same fused epilogue as conv_tc.cu.
#include "common.cuh"

namespace {

constexpr int BM = 64, BK = 32, LDS = 40;   // LDS: padded smem row (bf16) -> conflict-free ldmatrix

struct GenParams {
    const bf16* x_hi; const bf16* x_lo;
    int x_planes, w_planes, out_planes;   // plane counts of the MMA operands (1..3 bf16 planes after conversion) / plane CODE of the outputs
    int x_f16, w_f16, add_code;           // fp16 source planes are converted to bf16 planes on load; plane code of the addend
    float acc_mul; const float* acc_mul_slot; const float* add_mul_slot;   // as ConvTcParams
    int N, H, W, Cin;          // stored input tensor
    int up;                    // 1: input is read through a nearest x2 upsample (K6)
    const bf16* w_hi; const bf16* w_lo;
    int Ktot;                  // taps * Cin (row length of the packed weight matrix)
    int ks, stride, pad, transposed;
    int Ho, Wo, Cout_pad;      // output size, weight rows (multiple of 8)
    int Cstore;                // channel count of the NHWC output / mask / add tensors
    int Cbias;
    const float* bias; const bf16* mask; const bf16* add_hi; const bf16* add_lo;
    bf16* out_hi; bf16* out_lo; bf16* relu_hi; bf16* relu_lo;
    float* out_nchw; int Cnchw;
    int M;                     // N*Ho*Wo
};

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t addr, uint32_t& a, uint32_t& b) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int BN>
__global__ void __launch_bounds__(128) conv_gen_kernel(const GenParams p) {
    pdl_prologue();
    __shared__ __align__(16) bf16 As[3][BM][LDS];
    __shared__ __align__(16) bf16 Bs[3][BN][LDS];
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int ap = p.x_planes, wp = p.w_planes, maxord = max(ap, wp) - 1;
    const bf16* x_p2 = p.x_f16 ? nullptr : third_plane(p.x_hi, p.x_lo, ap);
    const bf16* w_p2 = p.w_f16 ? nullptr : third_plane(p.w_hi, p.w_lo, wp);
    const float acc_mul = p.acc_mul * slot_inv(p.acc_mul_slot), add_mul = slot_inv(p.add_mul_slot);
    const int Hl = p.H << p.up, Wl = p.W << p.up;

    // the two A rows this thread gathers
    int pn[2], poh[2], pow_[2]; bool pv[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        int m = m0 + (t >> 2) + i * 32;
        pv[i] = m < p.M;
        int mm = pv[i] ? m : 0;
        pow_[i] = mm % p.Wo; mm /= p.Wo; poh[i] = mm % p.Ho; pn[i] = mm / p.Ho;
    }
    const int kv = (t & 3) * 8;

    float acc[BN / 8][4];
#pragma unroll
    for (int j = 0; j < BN / 8; ++j) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f; }

    for (int k0 = 0; k0 < p.Ktot; k0 += BK) {
        // ---- gather A ----
        const int k = k0 + kv;
        int tap = 0, ci = 0, r = 0, s = 0;
        const bool kvalid = k < p.Ktot;
        if (kvalid) { tap = k / p.Cin; ci = k - tap * p.Cin; r = tap / p.ks; s = tap - r * p.ks; }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            uint4 vh = make_uint4(0, 0, 0, 0), vl = make_uint4(0, 0, 0, 0), v2 = make_uint4(0, 0, 0, 0);
            if (kvalid && pv[i]) {
                int ih, iw; bool ok;
                if (!p.transposed) {
                    ih = poh[i] * p.stride + r - p.pad; iw = pow_[i] * p.stride + s - p.pad;
                    ok = ih >= 0 && ih < Hl && iw >= 0 && iw < Wl;
                    ih >>= p.up; iw >>= p.up;
                } else {
                    int th = poh[i] + p.pad - r, tw = pow_[i] + p.pad - s;
                    ok = th >= 0 && tw >= 0 && (th % p.stride) == 0 && (tw % p.stride) == 0;
                    ih = th / p.stride; iw = tw / p.stride;
                    ok = ok && ih < p.H && iw < p.W;
                }
                if (ok) {
                    const size_t off = ((static_cast<size_t>(pn[i]) * p.H + ih) * p.W + iw) * p.Cin + ci;
                    vh = *reinterpret_cast<const uint4*>(p.x_hi + off);
                    if (p.x_lo) vl = *reinterpret_cast<const uint4*>(p.x_lo + off);
                    if (x_p2) v2 = *reinterpret_cast<const uint4*>(x_p2 + off);
                    if (p.x_f16) f16x8_to_bf16_planes(vh, vl, v2);
                }
            }
            const int row = (t >> 2) + i * 32;
            *reinterpret_cast<uint4*>(&As[0][row][kv]) = vh;
            *reinterpret_cast<uint4*>(&As[1][row][kv]) = vl;
            *reinterpret_cast<uint4*>(&As[2][row][kv]) = v2;
        }
        // ---- load B ----
        for (int idx = t; idx < BN * 4; idx += 128) {
            const int row = idx >> 2, kb = (idx & 3) * 8;
            uint4 vh = make_uint4(0, 0, 0, 0), vl = make_uint4(0, 0, 0, 0), v2 = make_uint4(0, 0, 0, 0);
            if (n0 + row < p.Cout_pad && k0 + kb < p.Ktot) {
                const size_t off = static_cast<size_t>(n0 + row) * p.Ktot + k0 + kb;
                vh = *reinterpret_cast<const uint4*>(p.w_hi + off);
                if (p.w_lo) vl = *reinterpret_cast<const uint4*>(p.w_lo + off);
                if (w_p2) v2 = *reinterpret_cast<const uint4*>(w_p2 + off);
                if (p.w_f16) f16x8_to_bf16_planes(vh, vl, v2);
            }
            *reinterpret_cast<uint4*>(&Bs[0][row][kb]) = vh;
            *reinterpret_cast<uint4*>(&Bs[1][row][kb]) = vl;
            *reinterpret_cast<uint4*>(&Bs[2][row][kb]) = v2;
        }
        __syncthreads();
        // ---- MMA ----
#pragma unroll
        for (int kk = 0; kk < BK; kk += 16) {
            uint32_t af[3][4];
            const int arow = warp * 16 + (lane & 15), acol = kk + (lane >> 4) * 8;
#pragma unroll
            for (int i = 0; i < 3; ++i)
                if (i < ap) ldmatrix_x4(static_cast<uint32_t>(__cvta_generic_to_shared(&As[i][arow][acol])), af[i][0], af[i][1], af[i][2], af[i][3]);
#pragma unroll
            for (int j = 0; j < BN / 8; ++j) {
                const int brow = j * 8 + (lane & 7), bcol = kk + ((lane >> 3) & 1) * 8;
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    if (q >= wp) break;
                    uint32_t b0, b1;
                    ldmatrix_x2(static_cast<uint32_t>(__cvta_generic_to_shared(&Bs[q][brow][bcol])), b0, b1);
                    // plane pairs (i, q) with i + q <= maxord: 1, 3 or 6 passes per product
#pragma unroll
                    for (int i = 0; i < 3; ++i)
                        if (i < ap && i + q <= maxord) mma16816(acc[j], af[i], b0, b1);
                }
            }
        }
        __syncthreads();
    }

    // ---- epilogue ----
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int m = m0 + warp * 16 + (lane >> 2) + half * 8;
        if (m >= p.M) continue;
        int mm = m; const int ow = mm % p.Wo; mm /= p.Wo; const int oh = mm % p.Ho; const int n = mm / p.Ho;
#pragma unroll
        for (int j = 0; j < BN / 8; ++j) {
            const int ch = n0 + j * 8 + (lane & 3) * 2;
            float v0 = acc[j][half * 2] * acc_mul, v1 = acc[j][half * 2 + 1] * acc_mul;
            if (p.bias) { if (ch < p.Cbias) v0 += p.bias[ch]; if (ch + 1 < p.Cbias) v1 += p.bias[ch + 1]; }
            if (ch < p.Cstore) {
                const size_t o = static_cast<size_t>(m) * p.Cstore + ch;
                if (p.mask) {
                    const uint32_t mk = *reinterpret_cast<const uint32_t*>(p.mask + o);
                    if (!pos16_lo(mk)) v0 = 0.f;
                    if (!pos16_hi(mk)) v1 = 0.f;
                }
                if (p.add_hi) {
                    const bool af = pc_f16(p.add_code);
                    float2 a = unpack16x2(*reinterpret_cast<const uint32_t*>(p.add_hi + o), af);
                    if (p.add_lo && pc_planes(p.add_code) > 1) {
                        const float2 b = unpack16x2(*reinterpret_cast<const uint32_t*>(p.add_lo + o), af);
                        a.x += b.x; a.y += b.y;
                        const bf16* a2 = third_plane(p.add_hi, p.add_lo, p.add_code);
                        if (a2) { const uint32_t c = *reinterpret_cast<const uint32_t*>(a2 + o); a.x += bf16lo_to_f(c); a.y += bf16hi_to_f(c); }
                    }
                    v0 = fmaf(a.x, add_mul, v0); v1 = fmaf(a.y, add_mul, v1);
                }
                if (p.out_hi) {
                    uint32_t h, m, l; split2c(v0, v1, p.out_planes, h, m, l);
                    *reinterpret_cast<uint32_t*>(p.out_hi + o) = h;
                    if (p.out_lo && pc_planes(p.out_planes) > 1) *reinterpret_cast<uint32_t*>(p.out_lo + o) = m;
                    bf16* o2 = third_plane(p.out_hi, p.out_lo, p.out_planes);
                    if (o2) *reinterpret_cast<uint32_t*>(o2 + o) = l;
                }
                if (p.relu_hi) {
                    uint32_t h, m, l; split2c(fmaxf(v0, 0.f), fmaxf(v1, 0.f), p.out_planes, h, m, l);
                    *reinterpret_cast<uint32_t*>(p.relu_hi + o) = h;
                    if (p.relu_lo && pc_planes(p.out_planes) > 1) *reinterpret_cast<uint32_t*>(p.relu_lo + o) = m;
                    bf16* o2 = third_plane(p.relu_hi, p.relu_lo, p.out_planes);
                    if (o2) *reinterpret_cast<uint32_t*>(o2 + o) = l;
                }
            }
            if (p.out_nchw) {
                const size_t plane = static_cast<size_t>(p.Ho) * p.Wo;
                const size_t base = (static_cast<size_t>(n) * p.Cnchw) * plane + static_cast<size_t>(oh) * p.Wo + ow;
                if (ch < p.Cnchw) p.out_nchw[base + ch * plane] = v0;
                if (ch + 1 < p.Cnchw) p.out_nchw[base + (ch + 1) * plane] = v1;
            }
        }
    }
}

}  // namespace

// x_planes / w_planes / out_planes / add_code: plane CODES; fp16 source planes are re-split into three bf16 planes on load
int conv_gen_run(const bf16* x_hi, const bf16* x_lo, int x_planes, int N, int H, int W, int Cin, int up,
                 const bf16* w_hi, const bf16* w_lo, int w_planes, int ks, int stride, int pad, int transposed,
                 int Ho, int Wo, int Cout_pad, int Cstore, int Cbias,
                 const float* bias, const bf16* mask, const bf16* add_hi, const bf16* add_lo, int add_code,
                 bf16* out_hi, bf16* out_lo, bf16* relu_hi, bf16* relu_lo, int out_planes,
                 float* out_nchw, int Cnchw, float acc_mul, const float* acc_mul_slot, const float* add_mul_slot, cudaStream_t stream) {
    GenParams p;
    p.x_f16 = pc_f16(x_planes); p.w_f16 = pc_f16(w_planes);
    if (!x_lo || pc_planes(x_planes) < 2) x_lo = nullptr;
    if (!w_lo || pc_planes(w_planes) < 2) w_lo = nullptr;
    p.x_planes = p.x_f16 ? 3 : (x_lo ? (pc_planes(x_planes) >= 3 ? 3 : 2) : 1);
    p.w_planes = p.w_f16 ? 3 : (w_lo ? (pc_planes(w_planes) >= 3 ? 3 : 2) : 1);
    p.out_planes = out_planes; p.add_code = add_code;
    p.acc_mul = acc_mul; p.acc_mul_slot = acc_mul_slot; p.add_mul_slot = add_mul_slot;
    p.x_hi = x_hi; p.x_lo = x_lo; p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.up = up;
    p.w_hi = w_hi; p.w_lo = w_lo; p.Ktot = ks * ks * Cin;
    p.ks = ks; p.stride = stride; p.pad = pad; p.transposed = transposed;
    p.Ho = Ho; p.Wo = Wo; p.Cout_pad = Cout_pad; p.Cstore = Cstore; p.Cbias = Cbias;
    p.bias = bias; p.mask = mask; p.add_hi = add_hi; p.add_lo = add_lo;
    p.out_hi = out_hi; p.out_lo = out_lo; p.relu_hi = relu_hi; p.relu_lo = relu_lo;
    p.out_nchw = out_nchw; p.Cnchw = Cnchw;
    p.M = N * Ho * Wo;
    const int mblocks = (p.M + BM - 1) / BM;
    if (Cout_pad <= 8) {
        st_launch(dim3(mblocks, 1), 128, 0, stream)(conv_gen_kernel<8>, p);
    } else if (Cout_pad <= 16) {
        st_launch(dim3(mblocks, 1), 128, 0, stream)(conv_gen_kernel<16>, p);
    } else if (Cout_pad <= 32) {
        st_launch(dim3(mblocks, 1), 128, 0, stream)(conv_gen_kernel<32>, p);
    } else {
        st_launch(dim3(mblocks, (Cout_pad + 63) / 64), 128, 0, stream)(conv_gen_kernel<64>, p);
    }
    ST_CHECK_LAUNCH("conv_gen_kernel");
    return ST_OK;
}
