// Shared helpers for the pio_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>
#include <algorithm>

#include "../../include/pio_b200.h"

namespace pio {

// thread-local last error text, readable through pio_last_error()
void set_error(const char* fmt, ...);

#define PIO_CHECK_ARG(cond, ...)                    \
    do {                                            \
        if (!(cond)) {                              \
            pio::set_error(__VA_ARGS__);            \
            return PIO_ERR_ARG;                     \
        }                                           \
    } while (0)

#define PIO_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t _e = (call);                                                         \
        if (_e != cudaSuccess) {                                                         \
            pio::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e),      \
                           __FILE__, __LINE__);                                          \
            return PIO_ERR_CUDA;                                                         \
        }                                                                                \
    } while (0)

// every kernel launch is followed by this: counts the launch (pio_launch_count) and surfaces
// launch-configuration errors
void count_launch();
#define PIO_LAUNCH_CHECK()           \
    do {                             \
        pio::count_launch();         \
        PIO_CUDA(cudaGetLastError()); \
    } while (0)

#define PIO_TRY(expr)                  \
    do {                               \
        int _s = (expr);               \
        if (_s != PIO_OK) return _s;   \
    } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int num_sms();

// out1[i] (+)= sum_s part[s*stride+i] (i<n1); out2[i] (+)= sum_s part[s*stride+n1+i] (i<n2)
int reduce_partials(const float* part, int splits, int64_t stride, float* out1, int64_t n1, float* out2,
                    int64_t n2, int accumulate, cudaStream_t st);

// bump allocator over a caller-provided workspace
struct Workspace {
    char* base;
    size_t size;
    size_t off;
    Workspace(void* p, size_t n) : base((char*)p), size(n), off(0) {}
    void* take(size_t bytes) {
        size_t o = align_up(off, 256);
        if (o + bytes > size || base == nullptr) return nullptr;
        off = o + bytes;
        return base + o;
    }
    float* takef(size_t n) { return (float*)take(n * sizeof(float)); }
};

// Programmatic dependent launch (PDL).  A kernel launched through launch_pdl() may start while its
// predecessor in the stream is still running; it must call pdl_wait() before it touches global memory
// (the wait returns once every prerequisite grid has completed and its writes are visible).  Kernels call
// pdl_trigger() first thing so their own successor can be scheduled early and get its prologue (barrier
// init, TMEM allocation, descriptor prefetch, launch latency) out of the way.  Both are no-ops for a
// kernel launched the ordinary way.  Measured on the cfg2 step: eager launches 9.57 -> 9.02 ms; inside a
// captured graph the early-scheduled dependents only take SM resources from the side-stream branch
// (7.58 -> 8.01 ms), so launches made while the stream is capturing do not use it.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

bool pdl_enabled(cudaStream_t st);

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled(st) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Counter-based dropout: element `idx` of dropout site `seed` is kept iff hash >= threshold, so the
// forward and the backward kernels regenerate the same mask from (seed, index) without storing it.
// Sites are told apart by small seed offsets (seed + 1 = the Residual site, seed + 4 s + {0, 2} for latent layer
// s ...), so the site key is first run through a full 64-bit mixer on the host (make_drop) and only then
// combined with the element index: masks of neighbouring sites are independent bit patterns, not shifted
// copies of one another (tests/test_gpu_dropout.py::test_masks_of_neighbouring_sites_are_uncorrelated).
struct DropCfg {
    uint64_t seed;       // mixed site key (make_drop), not the caller's raw seed
    uint32_t threshold;  // p * 2^32 ; 0 = dropout off
    float inv_keep;      // 1 / (1 - p)
};
static inline uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline DropCfg make_drop(float p, uint64_t seed) {
    DropCfg d;
    d.seed = mix64(mix64(seed) ^ 0xA0761D6478BD642Full);
    if (!(p > 0.f)) {
        d.threshold = 0;
        d.inv_keep = 1.f;
    } else {
        const double t = (double)p * 4294967296.0;
        d.threshold = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
        d.inv_keep = 1.0f / (1.0f - p);
    }
    return d;
}
__device__ __forceinline__ bool drop_keep(const DropCfg& d, uint64_t idx) {
    uint64_t z = (idx * 0x9E3779B97F4A7C15ull) ^ d.seed;
    z ^= z >> 32;
    z *= 0xD6E8FEB86659FD93ull;
    z ^= z >> 32;
    z *= 0xD6E8FEB86659FD93ull;
    z ^= z >> 32;
    return (uint32_t)z >= d.threshold;
}

// out = dropout-mask(in): gradient of a dropout site w.r.t. its input (same mask as the forward)
int dropout_mask(const float* in, float* out, int64_t n, const DropCfg& d, cudaStream_t st);

__device__ __forceinline__ float gelu_erf(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float gelu_erf_grad(float x) {
    // d/dx [x Phi(x)] = Phi(x) + x phi(x)
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
    const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
    return cdf + x * pdf;
}

// GELU and its derivative for the tensor-core epilogues, where four warps per CTA do all the element-wise math:
// erf by Abramowitz-Stegun 7.1.26 (|error| < 1.5e-7, i.e. at fp32 rounding level for the cdf), whose exp(-u^2) with
// u = x / sqrt(2) is the very exp(-x^2 / 2) the derivative's density term needs — one ex2 and one reciprocal per
// element instead of erff + expf.
__device__ __forceinline__ void gelu_erf_both(float x, float& g, float& dg) {
    const float a = fabsf(x) * 0.70710678118654752440f;
    float t;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, a, 1.0f)));
    float poly = fmaf(t, 1.061405429f, -1.453152027f);
    poly = fmaf(t, poly, 1.421413741f);
    poly = fmaf(t, poly, -0.284496736f);
    poly = fmaf(t, poly, 0.254829592f);
    poly *= t;
    float e;  // exp(-x^2 / 2)
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-a * a * 1.44269504088896340736f));
    const float erf_abs = fmaf(-poly, e, 1.0f);
    const float cdf = 0.5f * (1.0f + copysignf(erf_abs, x));
    g = x * cdf;
    dg = fmaf(x * 0.39894228040143267794f, e, cdf);
}

}  // namespace pio
