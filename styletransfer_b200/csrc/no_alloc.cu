// st_instnorm_fwd_static: InstanceNorm2d forward for callers that may not allocate (code under CUDA-graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  Same kernels as st_instnorm_fwd (norm_losses.cu); the
// per-split statistics partials live in a file-scope device array instead of a caller-provided workspace.
//
// The array is part of the module image: every GPU that loads the library holds ST_INSTNORM_STATIC_SCRATCH_BYTES of it
// until the process exits.  All calls share it (it keeps the last call's contents), so two calls that overlap on the
// GPU (different streams, no dependency) race; calls on one stream are ordered and safe.  Nothing in it has to start
// at zero: the statistics kernel stores every partial the finalize kernel reads.
#include "common.cuh"
#include "../../include/stransfer_b200.h"
#include <mutex>

int in_forward(const bf16*, const bf16*, const float*, const float*, float, bf16*, bf16*, int, float*, float*, int, int, int, int, int, int, float*, cudaStream_t);

namespace {
constexpr int IN_MAX_SPLITS = 128;                                   // norm_losses.cu: in_splits() never exceeds it
constexpr size_t SCRATCH_FLOATS = static_cast<size_t>(ST_INSTNORM_STATIC_MAX_NC) * IN_MAX_SPLITS * 2;   // [N][splits][2][C]
static_assert(SCRATCH_FLOATS * sizeof(float) == ST_INSTNORM_STATIC_SCRATCH_BYTES, "header and kernel file disagree");
}  // namespace
static __device__ __align__(256) float g_in_static_scratch[SCRATCH_FLOATS];

namespace {
constexpr int MAX_DEVICES = 64;
float* g_scratch_ptr[MAX_DEVICES] = {};
std::mutex g_scratch_mutex;

// In host code g_in_static_scratch is only a handle of the symbol: the device address comes from the runtime, once per
// device, with that device current.
int scratch_for_current_device(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        st_set_error("st_instnorm_fwd_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return ST_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[dev]) {
        float* ptr = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_in_static_scratch);
        if (e != cudaSuccess || !ptr) {
            st_set_error("st_instnorm_fwd_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return ST_ERR_CUDA;
        }
        g_scratch_ptr[dev] = ptr;
    }
    *out = g_scratch_ptr[dev];
    return ST_OK;
}
}  // namespace

extern "C" int st_instnorm_fwd_static(const void* x_hi, const void* x_lo, const float* gamma, const float* beta, float eps,
                                      void* y_hi, void* y_lo, int planes, float* mean, float* rstd,
                                      int N, int H, int W, int C, int relu, int up, void* stream) {
    ST_CHECK_ARG(x_hi && gamma && beta && y_hi && mean && rstd, "st_instnorm_fwd_static: null pointer");
    ST_CHECK_ARG(N > 0 && H > 0 && W > 0, "st_instnorm_fwd_static: empty shape");
    ST_CHECK_ARG(C >= 8 && C <= 256 && (C & (C - 1)) == 0, "st_instnorm_fwd_static: C must be a power of two in [8,256] (got %d)", C);
    const int n = pc_planes(planes);
    ST_CHECK_ARG(n >= 1 && n <= (pc_f16(planes) ? 2 : 3) && (planes & ~31) == 0, "st_instnorm_fwd_static: bad plane code %d", planes);
    ST_CHECK_ARG(static_cast<long long>(N) * C <= ST_INSTNORM_STATIC_MAX_NC,
                 "st_instnorm_fwd_static: N*C = %lld exceeds the built-in scratch (N*C <= %d); use st_instnorm_fwd with a workspace",
                 static_cast<long long>(N) * C, ST_INSTNORM_STATIC_MAX_NC);
    float* scratch = nullptr;
    if (int rc = scratch_for_current_device(&scratch)) return rc;
    const int code = x_lo ? planes : ((planes & ST_FMT_F16) | 1);
    return in_forward(static_cast<const bf16*>(x_hi), static_cast<const bf16*>(x_lo), gamma, beta, eps,
                      static_cast<bf16*>(y_hi), static_cast<bf16*>(y_lo), code, mean, rstd, N, H, W, C, relu, up,
                      scratch, static_cast<cudaStream_t>(stream));
}
