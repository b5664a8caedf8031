// Shared device helpers for the stransfer sm_100a kernels.
//
// Data format used by every kernel ("split-bf16 NHWC"): an activation tensor is
// stored channels-last as TWO bf16 planes, hi = bf16(x) and lo = bf16(x - hi).
// hi + lo carries ~16 mantissa bits, and a product a*b is evaluated on the bf16
// tensor cores as a_hi*b_hi + a_lo*b_hi + a_hi*b_lo with fp32 accumulation
// (the "bf16x3" scheme).  A NULL lo plane means plain bf16 (one MMA per
// product).  Why: DESIGN.md section "Precision" (single-pass bf16 gives a 3e-2 error
// on the Gatys input gradient; the north-star tolerance is 1e-3).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda.h>
#include <stdint.h>

typedef __nv_bfloat16 bf16;

// ----- plane codes -----------------------------------------------------------------
// Every "planes" argument of the library is a PLANE CODE: bits 0..3 = number of 16-bit planes (1..3),
// bit 4 = element format of the planes (0 = bf16, 1 = fp16; fp16 tensors have at most 2 planes).  Plane
// pointers stay typed bf16* (a 16-bit slot); the code says how the bits are read.
//   bf16 x 3 planes: 24 mantissa bits, fp32 range, 6 tensor-core passes per product ("bf16x6")
//   fp16 x 2 planes: 22 mantissa bits, fp16 range (values beyond +-65504 saturate, values below 2^-3 keep an
//                    absolute precision of 2^-25), 3 passes per product ("f16x3")
// A tensor whose values may leave the fp16 range (gradients) is stored multiplied by a power of two that lives
// in a 2-float device slot {scale, 1/scale} next to it; consumers multiply their accumulators by slot[1].
#define ST_FMT_F16 16
__host__ __device__ __forceinline__ int pc_planes(int code) { return code & 15; }
__host__ __device__ __forceinline__ bool pc_f16(int code) { return (code & ST_FMT_F16) != 0; }
// fixed power-of-two factor of fp16-packed convolution weights (|w| ~ 1e-2 would leave the lo plane subnormal)
#define ST_W_F16_SCALE 256.0f
#define ST_W_F16_INV_SCALE (1.0f / 256.0f)

// ----- the seeds of several style nodes in one launch (elementwise.cu, api.cu) ----
#define ST_MAX_STYLE_NODES 16
struct StyleSeedEntry {
    const float* diff; int total; float scale; const float* gw;
    __nv_bfloat16* dh; __nv_bfloat16* dl; int code; const float* amax; float* slot;
};
struct StyleSeedTable { int n; StyleSeedEntry e[ST_MAX_STYLE_NODES]; };

// ----- error plumbing (host) -------------------------------------------------
void st_set_error(const char* fmt, ...);
#define ST_OK 0
#define ST_ERR_INVALID (-1)
#define ST_ERR_CUDA (-2)
#define ST_ERR_UNSUPPORTED (-3)
#define ST_CHECK_ARG(cond, ...)                                   \
    do { if (!(cond)) { st_set_error(__VA_ARGS__); return ST_ERR_INVALID; } } while (0)
extern long long g_st_launches;      // kernels launched by this library (st_launch_count)
#define ST_CHECK_LAUNCH(name)                                                       \
    do { ++g_st_launches; cudaError_t e_ = cudaGetLastError();                                       \
         if (e_ != cudaSuccess) { st_set_error("%s: %s", name, cudaGetErrorString(e_)); \
                                  return ST_ERR_CUDA; } } while (0)

// ----- programmatic dependent launch ---------------------------------------------------------------------------------
// Every kernel of the library can be launched with the programmatic-stream-serialization attribute (st_launch) and starts
// with pdl_prologue(): it lets the NEXT kernel of the stream be scheduled as soon as every CTA of this one has started
// (its CTAs become resident as this kernel's CTAs retire, SM by SM, instead of after the whole grid has drained and the
// dependency has been resolved by the front end), and then waits until the PREVIOUS kernel has completed and its memory
// is visible before touching global memory.  Inside a captured CUDA graph the launches become programmatic edges.
// With the attribute off (the default; STRANSFER_PDL=1 turns it on, see api.cu for the measurement) both instructions are no-ops.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_prologue() { pdl_trigger(); pdl_wait(); }
extern int g_st_pdl;
struct StLaunch {
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr;
    StLaunch(dim3 grid, dim3 block, size_t smem, cudaStream_t st) {
        cfg = cudaLaunchConfig_t{};
        cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
        attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr.val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = &attr; cfg.numAttrs = g_st_pdl ? 1 : 0;
    }
    StLaunch(const StLaunch&) = delete;
    template <class... KArgs, class... Args>
    void operator()(void (*kernel)(KArgs...), Args&&... args) {
        cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    }
};
#define st_launch StLaunch

// ----- split-bf16 -------------------------------------------------------------
__device__ __forceinline__ void split_bf16(float v, bf16& hi, bf16& lo) {
    hi = __float2bfloat16_rn(v);
    lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}
__device__ __forceinline__ float join_bf16(bf16 hi, bf16 lo) {
    return __bfloat162float(hi) + __bfloat162float(lo);
}
// pack two floats into a bf16x2 word (a = low half)
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float bf16lo_to_f(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi_to_f(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
// split 2 floats -> (hi word, lo word)
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    hi = pack_bf16x2(a, b);
    lo = pack_bf16x2(a - bf16lo_to_f(hi), b - bf16hi_to_f(hi));
}

// ----- fp16 planes ---------------------------------------------------------------
// two floats -> f16x2 word (a = low half), round to nearest, saturating to +-65504 instead of inf
__device__ __forceinline__ uint32_t pack_f16x2(float a, float b) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}
__device__ __forceinline__ float2 unpack_f16x2(uint32_t w) {
    return __half22float2(*reinterpret_cast<__half2*>(&w));
}
__device__ __forceinline__ void split2h(float a, float b, uint32_t& hi, uint32_t& lo) {
    hi = pack_f16x2(a, b);
    const float2 h = unpack_f16x2(hi);
    lo = pack_f16x2(a - h.x, b - h.y);
}
// sign test of a 16-bit float pair (bf16 or fp16 alike): value > 0  <=>  the bits are a positive non-zero integer
__device__ __forceinline__ bool pos16_lo(uint32_t w) { return static_cast<int16_t>(w & 0xffffu) > 0; }
__device__ __forceinline__ bool pos16_hi(uint32_t w) { return static_cast<int32_t>(w) > 0xffff; }     // one compare, no mask
// one 16-bit-pair word of plane format f16 / bf16 -> two floats
__device__ __forceinline__ float2 unpack16x2(uint32_t w, bool f16) {
    return f16 ? unpack_f16x2(w) : make_float2(bf16lo_to_f(w), bf16hi_to_f(w));
}

// 8 channels (one 16-byte vector per plane) <-> 8 floats
__device__ __forceinline__ void load8(const bf16* hi, const bf16* lo, size_t off, float (&v)[8]) {
    uint4 h = *reinterpret_cast<const uint4*>(hi + off);
    uint32_t hw[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { v[2 * i] = bf16lo_to_f(hw[i]); v[2 * i + 1] = bf16hi_to_f(hw[i]); }
    if (lo) {
        uint4 l = *reinterpret_cast<const uint4*>(lo + off);
        uint32_t lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { v[2 * i] += bf16lo_to_f(lw[i]); v[2 * i + 1] += bf16hi_to_f(lw[i]); }
    }
}
__device__ __forceinline__ void store8(bf16* hi, bf16* lo, size_t off, const float (&v)[8]) {
    uint32_t hw[4], lw[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) split2(v[2 * i], v[2 * i + 1], hw[i], lw[i]);
    *reinterpret_cast<uint4*>(hi + off) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
    if (lo) *reinterpret_cast<uint4*>(lo + off) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
}

// ----- three-plane variants ("bf16x6"): hi + mid + lo carries ~24 mantissa bits.  The
// third plane of a tensor always sits one plane stride after the second:
//     p2 = lo + (lo - hi).
__device__ __forceinline__ const bf16* third_plane(const bf16* hi, const bf16* lo, int code) {
    return (pc_planes(code) >= 3 && !pc_f16(code) && lo) ? lo + (lo - hi) : nullptr;
}
__device__ __forceinline__ bf16* third_plane(bf16* hi, bf16* lo, int code) {
    return (pc_planes(code) >= 3 && !pc_f16(code) && lo) ? lo + (lo - hi) : nullptr;
}
__device__ __forceinline__ void load8h(const bf16* hi, const bf16* lo, size_t off, float (&v)[8]) {
    uint4 h = *reinterpret_cast<const uint4*>(hi + off);
    uint32_t hw[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = unpack_f16x2(hw[i]); v[2 * i] = t.x; v[2 * i + 1] = t.y; }
    if (lo) {
        uint4 l = *reinterpret_cast<const uint4*>(lo + off);
        uint32_t lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 t = unpack_f16x2(lw[i]); v[2 * i] += t.x; v[2 * i + 1] += t.y; }
    }
}
__device__ __forceinline__ void store8h(bf16* hi, bf16* lo, size_t off, const float (&v)[8]) {
    uint32_t hw[4], lw[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) split2h(v[2 * i], v[2 * i + 1], hw[i], lw[i]);
    *reinterpret_cast<uint4*>(hi + off) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
    if (lo) *reinterpret_cast<uint4*>(lo + off) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
}
// `code`: plane code of the tensor (see the top of this file)
__device__ __forceinline__ void load8p(const bf16* hi, const bf16* lo, int code, size_t off, float (&v)[8]) {
    const int planes = pc_planes(code);
    if (pc_f16(code)) { load8h(hi, planes >= 2 ? lo : nullptr, off, v); return; }
    load8(hi, planes >= 2 ? lo : nullptr, off, v);
    const bf16* p2 = third_plane(hi, lo, planes);
    if (p2) {
        uint4 l = *reinterpret_cast<const uint4*>(p2 + off);
        uint32_t lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { v[2 * i] += bf16lo_to_f(lw[i]); v[2 * i + 1] += bf16hi_to_f(lw[i]); }
    }
}
// The same in two steps -- raw 16-byte loads of every plane first, arithmetic afterwards -- for loops that want ALL the
// loads of an iteration in flight before the first use: load8p branches on the plane code between its loads, which
// makes every load -> convert pair a separate exposed round trip (profiles/r02ac_in_kernels_fast_st.txt).  Absent
// planes are read from the hi plane (an L1 hit) and replaced by zeros, so the instruction stream has no branches.
struct Raw8p { uint4 h, l, t; };
__device__ __forceinline__ void load8p_raw(const bf16* hi, const bf16* lo, int code, size_t off, Raw8p& r) {
    const int planes = pc_planes(code);
    const bf16* lp = (planes >= 2 && lo) ? lo : hi;
    const bf16* tp = third_plane(hi, lo, code);
    const bf16* tq = tp ? tp : hi;
    r.h = __ldg(reinterpret_cast<const uint4*>(hi + off));
    r.l = __ldg(reinterpret_cast<const uint4*>(lp + off));
    r.t = __ldg(reinterpret_cast<const uint4*>(tq + off));
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    if (lp == hi) r.l = z;
    if (!tp) r.t = z;
}
__device__ __forceinline__ void decode8p(const Raw8p& r, int code, float (&v)[8]) {
    const uint32_t hw[4] = {r.h.x, r.h.y, r.h.z, r.h.w}, lw[4] = {r.l.x, r.l.y, r.l.z, r.l.w}, tw[4] = {r.t.x, r.t.y, r.t.z, r.t.w};
    if (pc_f16(code)) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 a = unpack_f16x2(hw[i]), b = unpack_f16x2(lw[i]);
            v[2 * i] = a.x + b.x; v[2 * i + 1] = a.y + b.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // same association as load8p: (hi + lo) + third
            v[2 * i] = (bf16lo_to_f(hw[i]) + bf16lo_to_f(lw[i])) + bf16lo_to_f(tw[i]);
            v[2 * i + 1] = (bf16hi_to_f(hw[i]) + bf16hi_to_f(lw[i])) + bf16hi_to_f(tw[i]);
        }
    }
}
__device__ __forceinline__ void store8p(bf16* hi, bf16* lo, int code, size_t off, const float (&v)[8]) {
    const int planes = pc_planes(code);
    if (pc_f16(code)) { store8h(hi, planes >= 2 ? lo : nullptr, off, v); return; }
    uint32_t hw[4], mw[4], lw[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float a = v[2 * i], b = v[2 * i + 1];
        hw[i] = pack_bf16x2(a, b);
        const float ra = a - bf16lo_to_f(hw[i]), rb = b - bf16hi_to_f(hw[i]);
        mw[i] = pack_bf16x2(ra, rb);
        lw[i] = pack_bf16x2(ra - bf16lo_to_f(mw[i]), rb - bf16hi_to_f(mw[i]));
    }
    *reinterpret_cast<uint4*>(hi + off) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
    if (planes >= 2 && lo) *reinterpret_cast<uint4*>(lo + off) = make_uint4(mw[0], mw[1], mw[2], mw[3]);
    bf16* p2 = third_plane(hi, lo, planes);
    if (p2) *reinterpret_cast<uint4*>(p2 + off) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
}
// 256-bit global accesses (sm_100): one full 32-byte sector per lane
__device__ __forceinline__ void stg256(void* p, const uint32_t (&w)[8]) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]),
                 "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]) : "memory");
}
__device__ __forceinline__ void ldg256(const void* p, uint32_t (&w)[8]) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]),
                 "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
}
// 16 consecutive channels -> up to three bf16 planes, 32 bytes each (off: multiple of 16 elements)
__device__ __forceinline__ void store16p(bf16* hi, bf16* lo, int code, size_t off, const float (&v)[16]) {
    const int planes = pc_planes(code);
    uint32_t hw[8], mw[8], lw[8];
    if (pc_f16(code)) {
#pragma unroll
        for (int i = 0; i < 8; ++i) split2h(v[2 * i], v[2 * i + 1], hw[i], mw[i]);
        stg256(hi + off, hw);
        if (planes >= 2 && lo) stg256(lo + off, mw);
        return;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float a = v[2 * i], b = v[2 * i + 1];
        hw[i] = pack_bf16x2(a, b);
        const float ra = a - bf16lo_to_f(hw[i]), rb = b - bf16hi_to_f(hw[i]);
        mw[i] = pack_bf16x2(ra, rb);
        lw[i] = pack_bf16x2(ra - bf16lo_to_f(mw[i]), rb - bf16hi_to_f(mw[i]));
    }
    stg256(hi + off, hw);
    if (planes >= 2 && lo) stg256(lo + off, mw);
    bf16* p2 = third_plane(hi, lo, planes);
    if (p2) stg256(p2 + off, lw);
}
__device__ __forceinline__ void unpack16(const uint32_t (&w)[8], float (&v)[16]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[2 * i] = bf16lo_to_f(w[i]); v[2 * i + 1] = bf16hi_to_f(w[i]); }
}
// 2 values -> the 16-bit-pair words of the planes of a tensor with plane code `code` (l unused for fp16)
__device__ __forceinline__ void split2c(float a, float b, int code, uint32_t& h, uint32_t& m, uint32_t& l) {
    if (pc_f16(code)) { split2h(a, b, h, m); l = 0; return; }
    h = pack_bf16x2(a, b);
    const float ra = a - bf16lo_to_f(h), rb = b - bf16hi_to_f(h);
    m = pack_bf16x2(ra, rb);
    l = pack_bf16x2(ra - bf16lo_to_f(m), rb - bf16hi_to_f(m));
}
// 2 values -> up to three bf16x2 words
__device__ __forceinline__ void split2p(float a, float b, uint32_t& h, uint32_t& m, uint32_t& l) {
    h = pack_bf16x2(a, b);
    const float ra = a - bf16lo_to_f(h), rb = b - bf16hi_to_f(h);
    m = pack_bf16x2(ra, rb);
    l = pack_bf16x2(ra - bf16lo_to_f(m), rb - bf16hi_to_f(m));
}

// The legacy mma.sync kernels multiply bf16 planes only: eight channels of an fp16-plane tensor (hi, lo vectors)
// are re-split into bf16 hi / lo / third planes on the way into shared memory (their value hi + lo is exact in fp32)
__device__ __forceinline__ void f16x8_to_bf16_planes(uint4& vh, uint4& vl, uint4& v2) {
    uint32_t hw[4] = {vh.x, vh.y, vh.z, vh.w}, lw[4] = {vl.x, vl.y, vl.z, vl.w}, o0[4], o1[4], o2[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 a = unpack_f16x2(hw[i]), b = unpack_f16x2(lw[i]);
        split2p(a.x + b.x, a.y + b.y, o0[i], o1[i], o2[i]);
    }
    vh = make_uint4(o0[0], o0[1], o0[2], o0[3]); vl = make_uint4(o1[0], o1[1], o1[2], o1[3]); v2 = make_uint4(o2[0], o2[1], o2[2], o2[3]);
}

// ----- power-of-two scale slots ----------------------------------------------------
// slot = {scale, 1/scale}.  The producer of a tensor bounded by `bound` picks the power of two that maps the
// bound to 2^target_log2 (fp16 holds 65504 ~ 2^16: target 10 leaves a factor 64 of head room).
__device__ __forceinline__ float pow2_scale_for(float bound, float target_log2) {
    if (!(bound > 0.f) || !(bound < 3.0e38f)) return 1.f;
    float e = target_log2 - ceilf(log2f(bound));
    e = fminf(fmaxf(e, -100.f), 100.f);
    return exp2f(e);
}
__device__ __forceinline__ float slot_inv(const float* slot) { return slot ? slot[1] : 1.f; }

// ----- reductions ---------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// block-wide sum; result valid in thread 0.  `sm` must hold >= 32 floats.
__device__ __forceinline__ float block_sum(float v, float* sm) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sm[w] = v;
    __syncthreads();
    if (w == 0) {
        int nw = (blockDim.x + 31) >> 5;
        v = lane < nw ? sm[lane] : 0.f;
        v = warp_sum(v);
    }
    return v;
}

// block-wide max; result valid in thread 0.  `sm` must hold >= 32 floats.
__device__ __forceinline__ float block_max(float v, float* sm) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if (lane == 0) sm[w] = v;
    __syncthreads();
    if (w == 0) {
        int nw = (blockDim.x + 31) >> 5;
        v = lane < nw ? sm[lane] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    }
    return v;
}

static inline int st_num_sms() {
    static int n = 0;
    if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
    return n;
}
