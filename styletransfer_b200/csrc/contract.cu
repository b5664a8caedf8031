// Pixel-contraction kernel (mma.sync bf16, fp32 accumulate):
//     P[m][kidx] = sum over pixels of  dy[pixel][m] * gather(x)[pixel][kidx]
// with kidx = tap*Cin + ci.  Two users:
//   * conv backward-weight of the transform net (autograd of the nn.Conv2d call
//     sites network.py:468-481, :525-609): dy = grad of the conv output, x = its
//     input, summed over the whole batch;
//   * the Gram matrix F^T F of StyleLoss.gram_matrix (network.py:92-108): a 1x1
//     "wgrad" with dy = x = F, one matrix per image.
// Both operands are pixel-major in memory (NHWC), i.e. the contraction index is the
// slow one, so fragments are fetched with ldmatrix.trans.  The pixel range is split
// across CTAs (split-K); partial tiles go to a workspace and a finish kernel sums
// them in a fixed order (deterministic, no atomics).
#include "common.cuh"

namespace {

constexpr int TM = 64, TN = 64, TK = 32, LD = 72;

struct ContractParams {
    const bf16* dy_hi; const bf16* dy_lo; int Cdy;          // [N,Ho,Wo,Cdy]
    const bf16* x_hi; const bf16* x_lo; int H, W, Cin, up;   // [N,H,W,Cin]
    int N, Ho, Wo, ks, stride, pad;
    int Mrows;        // rows of P (= Cdy rounded up to TM by the grid)
    int Ktot;         // taps*Cin
    int groups;       // 1: sum over the whole batch; N: one result per image
    int splits;       // pixel splits per group
    int chunk;        // pixels per split (multiple of TK)
    float* partial;   // [groups*splits][Mrows_pad][Ktot_pad]
    int Mpad, Kpad;
    int dy_f16, x_f16; // fp16 source planes are re-split into bf16 hi / lo on load
};

__device__ __forceinline__ void ldm_x4_t(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(128) contract_kernel(const ContractParams p) {
    pdl_prologue();
    __shared__ __align__(16) bf16 Ds[2][TK][LD];
    __shared__ __align__(16) bf16 Xs[2][TK][LD];
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int m0 = blockIdx.x * TM, kn0 = blockIdx.y * TN;
    const int z = blockIdx.z, group = z / p.splits, split = z % p.splits;
    const int Pg = (p.N / p.groups) * p.Ho * p.Wo;          // pixels per group
    const int pbeg = split * p.chunk, pend = min(Pg, pbeg + p.chunk);
    const bool d_lo = p.dy_lo != nullptr, x_lo = p.x_lo != nullptr;
    const int Hl = p.H << p.up, Wl = p.W << p.up;
    const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;

    float acc[2][4][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc[i][j][0] = acc[i][j][1] = acc[i][j][2] = acc[i][j][3] = 0.f; }

    // per-thread load assignment: 2 vectors per plane per tile: rows (t>>3) and (t>>3)+16, vector t&7
    const int vcol = (t & 7) * 8;
    // x gather: decode this thread's kidx once (it does not change over the pixel loop)
    const int kidx = kn0 + vcol;
    const bool kvalid = kidx < p.Ktot;
    int ci = 0, r = 0, s = 0;
    if (kvalid) { int tap = kidx / p.Cin; ci = kidx - tap * p.Cin; r = tap / p.ks; s = tap - r * p.ks; }
    const bool mvalid = (m0 + vcol) < p.Cdy;

    for (int pb = pbeg; pb < pend; pb += TK) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int row = (t >> 3) + i * 16;
            const int pl = pb + row;                       // pixel within the group
            uint4 dh = make_uint4(0, 0, 0, 0), dl = dh, xh = dh, xl = dh;
            if (pl < pend) {
                const int g = group * Pg + pl;             // global output pixel
                if (mvalid) {
                    const size_t off = static_cast<size_t>(g) * p.Cdy + m0 + vcol;
                    dh = *reinterpret_cast<const uint4*>(p.dy_hi + off);
                    if (d_lo) dl = *reinterpret_cast<const uint4*>(p.dy_lo + off);
                    if (p.dy_f16) { uint4 d2; f16x8_to_bf16_planes(dh, dl, d2); }
                }
                if (kvalid) {
                    int q = g; const int ow = q % p.Wo; q /= p.Wo; const int oh = q % p.Ho; const int n = q / p.Ho;
                    int ih = oh * p.stride + r - p.pad, iw = ow * p.stride + s - p.pad;
                    if (ih >= 0 && ih < Hl && iw >= 0 && iw < Wl) {
                        ih >>= p.up; iw >>= p.up;
                        const size_t off = ((static_cast<size_t>(n) * p.H + ih) * p.W + iw) * p.Cin + ci;
                        xh = *reinterpret_cast<const uint4*>(p.x_hi + off);
                        if (x_lo) xl = *reinterpret_cast<const uint4*>(p.x_lo + off);
                        if (p.x_f16) { uint4 x2; f16x8_to_bf16_planes(xh, xl, x2); }
                    }
                }
            }
            *reinterpret_cast<uint4*>(&Ds[0][row][vcol]) = dh;
            *reinterpret_cast<uint4*>(&Ds[1][row][vcol]) = dl;
            *reinterpret_cast<uint4*>(&Xs[0][row][vcol]) = xh;
            *reinterpret_cast<uint4*>(&Xs[1][row][vcol]) = xl;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < TK; kk += 16) {
            uint32_t ah[2][4], al[2][4];
            const int arow = kk + (lane & 7) + ((lane >> 4) & 1) * 8;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int acol = wm + i * 16 + ((lane >> 3) & 1) * 8;
                ldm_x4_t(static_cast<uint32_t>(__cvta_generic_to_shared(&Ds[0][arow][acol])), ah[i][0], ah[i][1], ah[i][2], ah[i][3]);
                if (d_lo) ldm_x4_t(static_cast<uint32_t>(__cvta_generic_to_shared(&Ds[1][arow][acol])), al[i][0], al[i][1], al[i][2], al[i][3]);
            }
            const int brow = kk + (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                const int bcol = wn + jj * 16 + ((lane >> 4) & 1) * 8;
                uint32_t bh[4], bl[4];
                ldm_x4_t(static_cast<uint32_t>(__cvta_generic_to_shared(&Xs[0][brow][bcol])), bh[0], bh[1], bh[2], bh[3]);
                if (x_lo) ldm_x4_t(static_cast<uint32_t>(__cvta_generic_to_shared(&Xs[1][brow][bcol])), bl[0], bl[1], bl[2], bl[3]);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        float (&c)[4] = acc[i][jj * 2 + h];
                        mma16816(c, ah[i], bh[h * 2], bh[h * 2 + 1]);
                        if (d_lo) mma16816(c, al[i], bh[h * 2], bh[h * 2 + 1]);
                        if (x_lo) mma16816(c, ah[i], bl[h * 2], bl[h * 2 + 1]);
                    }
                }
            }
        }
        __syncthreads();
    }

    float* out = p.partial + static_cast<size_t>(z) * p.Mpad * p.Kpad;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int m = m0 + wm + i * 16 + (lane >> 2) + half * 8;
                const int k = kn0 + wn + j * 8 + (lane & 3) * 2;
                *reinterpret_cast<float2*>(out + static_cast<size_t>(m) * p.Kpad + k) =
                    make_float2(acc[i][j][half * 2], acc[i][j][half * 2 + 1]);
            }
}

// dW[co][ci][r][s] (PyTorch layout, fp32) = sum over splits of partial[split][co][tap*CinPad + ci]
__global__ void wgrad_finish_kernel(const float* partial, int splits, int Mpad, int Kpad,
                                    int Cout, int Cin, int CinPad, int taps, const float* dy_slot, float* dw) {
    pdl_prologue();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = Cout * Cin * taps;
    if (idx >= total) return;
    int q = idx; const int tap = q % taps; q /= taps; const int ci = q % Cin; const int co = q / Cin;
    const size_t off = static_cast<size_t>(co) * Kpad + tap * CinPad + ci;
    float s = 0.f;
    for (int sp = 0; sp < splits; ++sp) s += partial[static_cast<size_t>(sp) * Mpad * Kpad + off];
    dw[idx] = s * slot_inv(dy_slot);
}

// G[b][i][j] = scale * sum over splits (one result per image): 32 outputs x 8 split lanes per
// block, fixed-order combination (deterministic)
__global__ void __launch_bounds__(256)
gram_finish_kernel(const float* __restrict__ partial, int splits, int Mpad, int Kpad, int C, float scale,
                   float* __restrict__ G, int total) {
    pdl_prologue();
    __shared__ float sm[8][33];
    const int ol = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int idx = blockIdx.x * 32 + ol;
    float s = 0.f;
    if (idx < total) {
        int q = idx; const int j = q % C; q /= C; const int i = q % C; const int b = q / C;
        const float* base = partial + (static_cast<size_t>(b) * splits) * Mpad * Kpad + static_cast<size_t>(i) * Kpad + j;
        for (int sp = sl; sp < splits; sp += 8) s += base[static_cast<size_t>(sp) * Mpad * Kpad];
    }
    sm[sl][ol] = s;
    __syncthreads();
    if (sl == 0 && idx < total) {
        float t = 0.f;
#pragma unroll
        for (int l = 0; l < 8; ++l) t += sm[l][ol];
        G[idx] = t * scale;
    }
}

int pick_splits(int tiles, int pixels_per_group, int groups) {
    const int target = 8 * st_num_sms();       // ~8 resident CTAs per SM hide the gather latency
    int splits = (target + tiles * groups - 1) / (tiles * groups);
    const int max_splits = (pixels_per_group + 4 * TK - 1) / (4 * TK);     // at least 128 pixels per split
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    return splits;
}

}  // namespace

// workspace bytes needed by contract_run for these shapes
size_t contract_workspace_bytes(int N, int Ho, int Wo, int Cdy, int Ktot, int groups) {
    const int mt = (Cdy + TM - 1) / TM, kt = (Ktot + TN - 1) / TN;
    const int Pg = (N / groups) * Ho * Wo;
    const int splits = pick_splits(mt * kt, Pg, groups);
    return static_cast<size_t>(groups) * splits * (mt * TM) * (kt * TN) * sizeof(float);
}

// mode 0: wgrad -> dw [Cout][Cin][ks][ks];  mode 1: gram -> G [N][C][C] * scale
// dy_f16 / x_f16: the planes of that tensor are fp16; dy_slot (optional): power-of-two scale slot of dy (mode 0)
int contract_run(const bf16* dy_hi, const bf16* dy_lo, int Cdy, bool dy_f16, const float* dy_slot,
                 const bf16* x_hi, const bf16* x_lo, bool x_f16, int N, int H, int W, int CinPad, int up,
                 int Ho, int Wo, int ks, int stride, int pad,
                 int mode, int Cout, int Cin, float scale, float* out,
                 void* workspace, size_t workspace_bytes, cudaStream_t stream) {
    const int groups = mode == 1 ? N : 1;
    ContractParams p;
    p.dy_f16 = dy_f16; p.x_f16 = x_f16;
    p.dy_hi = dy_hi; p.dy_lo = dy_lo; p.Cdy = Cdy;
    p.x_hi = x_hi; p.x_lo = x_lo; p.H = H; p.W = W; p.Cin = CinPad; p.up = up;
    p.N = N; p.Ho = Ho; p.Wo = Wo; p.ks = ks; p.stride = stride; p.pad = pad;
    p.Ktot = ks * ks * CinPad;
    const int mt = (Cdy + TM - 1) / TM, kt = (p.Ktot + TN - 1) / TN;
    p.Mpad = mt * TM; p.Kpad = kt * TN; p.Mrows = Cdy;
    p.groups = groups;
    const int Pg = (N / groups) * Ho * Wo;
    p.splits = pick_splits(mt * kt, Pg, groups);
    p.chunk = ((Pg + p.splits - 1) / p.splits + TK - 1) / TK * TK;
    const size_t need = static_cast<size_t>(groups) * p.splits * p.Mpad * p.Kpad * sizeof(float);
    ST_CHECK_ARG(workspace && workspace_bytes >= need, "contract: workspace too small (%zu < %zu)", workspace_bytes, need);
    p.partial = static_cast<float*>(workspace);
    st_launch(dim3(mt, kt, groups * p.splits), 128, 0, stream)(contract_kernel, p);
    ST_CHECK_LAUNCH("contract_kernel");
    if (mode == 0) {
        const int total = Cout * Cin * ks * ks;
        st_launch((total + 255) / 256, 256, 0, stream)(wgrad_finish_kernel, p.partial, p.splits, p.Mpad, p.Kpad, Cout, Cin, CinPad, ks * ks, dy_slot, out);
        ST_CHECK_LAUNCH("wgrad_finish_kernel");
    } else {
        const int total = N * Cdy * Cdy;
        st_launch((total + 31) / 32, 256, 0, stream)(gram_finish_kernel, p.partial, p.splits, p.Mpad, p.Kpad, Cdy, scale, out, total);
        ST_CHECK_LAUNCH("gram_finish_kernel");
    }
    return ST_OK;
}
