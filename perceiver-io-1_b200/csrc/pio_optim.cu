// AdamW over a whole parameter set in one launch (include/pio_b200.h: pio_adamw_step / pio_adamw_tick).
//
// Bandwidth kernel: per element it reads p, g, m, v and writes p, m, v (28 bytes) plus, for matrix weights,
// the bf16 shadow the tensor-core GEMMs read (2 bytes) — so the optimizer step is also the re-cast of the
// shadows and no cast pass follows it.  One CTA per chunk of at most PIO_ADAMW_CHUNK elements of one tensor
// (tensor and chunk tables live in device memory; the host rewrites the tensor table only when a pointer moves), float4 accesses where the chunk
// allows.  Hyper-parameters and the step counter are read from device memory, so the launch can sit in a
// captured CUDA graph and still follow a learning-rate schedule or a replay count.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "pio_common.cuh"

namespace pio {

struct AdamHyper {
    float lr, b1, b2, eps, wd, c1, c2s;  // c1 = lr / (1 - b1^t), c2s = 1 / sqrt(1 - b2^t)
};

__device__ __forceinline__ float adamw_one(float p, float g, float& m, float& v, const AdamHyper& h) {
    p *= 1.0f - h.lr * h.wd;
    m = h.b1 * m + (1.0f - h.b1) * g;
    v = h.b2 * v + (1.0f - h.b2) * g * g;
    const float denom = sqrtf(v) * h.c2s + h.eps;
    return p - h.c1 * (m / denom);
}

struct AdamChunk {  // a chunk joined with its tensor entry
    float* p; const float* g; float* m; float* v; void* shadow16;
    int64_t offset;
    int count, cols, ld16;
};

__global__ void __launch_bounds__(256) adamw_kernel(const pio_adamw_tensor* __restrict__ tensors,
                                                    const pio_adamw_chunk* __restrict__ chunks, const float* __restrict__ hyper) {
    pdl_trigger();
    pdl_wait();
    AdamChunk ck;
    {
        const pio_adamw_chunk c = chunks[blockIdx.x];
        const pio_adamw_tensor t = tensors[c.tensor];
        ck.p = t.p; ck.g = t.g; ck.m = t.m; ck.v = t.v; ck.shadow16 = t.shadow16;
        ck.offset = c.offset; ck.count = c.count; ck.cols = t.cols; ck.ld16 = t.ld16;
    }
    AdamHyper h;
    h.lr = hyper[0]; h.b1 = hyper[1]; h.b2 = hyper[2]; h.eps = hyper[3]; h.wd = hyper[4];
    const float t = hyper[5] + 1.0f;
    h.c1 = h.lr / (1.0f - powf(h.b1, t));
    h.c2s = rsqrtf(1.0f - powf(h.b2, t));
    float* __restrict__ p = ck.p + ck.offset;
    const float* __restrict__ g = ck.g + ck.offset;
    float* __restrict__ m = ck.m + ck.offset;
    float* __restrict__ v = ck.v + ck.offset;
    __nv_bfloat16* sh = (__nv_bfloat16*)ck.shadow16;
    const bool vec = ((ck.offset & 3) == 0) && ((((uintptr_t)ck.p | (uintptr_t)ck.g | (uintptr_t)ck.m | (uintptr_t)ck.v) & 15) == 0);
    const bool dense16 = sh != nullptr && ck.cols == ck.ld16 && ((((uintptr_t)sh) & 7) == 0);  // shadow has the parameter's own layout
    const int n4 = vec ? (ck.count >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 pp = reinterpret_cast<float4*>(p)[i];
        const float4 gg = reinterpret_cast<const float4*>(g)[i];
        float4 mm = reinterpret_cast<float4*>(m)[i];
        float4 vv = reinterpret_cast<float4*>(v)[i];
        pp.x = adamw_one(pp.x, gg.x, mm.x, vv.x, h);
        pp.y = adamw_one(pp.y, gg.y, mm.y, vv.y, h);
        pp.z = adamw_one(pp.z, gg.z, mm.z, vv.z, h);
        pp.w = adamw_one(pp.w, gg.w, mm.w, vv.w, h);
        reinterpret_cast<float4*>(p)[i] = pp;
        reinterpret_cast<float4*>(m)[i] = mm;
        reinterpret_cast<float4*>(v)[i] = vv;
        if (sh != nullptr) {
            const int64_t e = ck.offset + 4 * (int64_t)i;
            if (dense16) {
                __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
                uint2 o;
                o.x = *reinterpret_cast<uint32_t*>(&lo);
                o.y = *reinterpret_cast<uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(sh + e) = o;
            } else {
                const float vals[4] = {pp.x, pp.y, pp.z, pp.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int64_t r = (e + j) / ck.cols;
                    sh[r * ck.ld16 + ((e + j) - r * ck.cols)] = __float2bfloat16(vals[j]);
                }
            }
        }
    }
    for (int i = 4 * n4 + threadIdx.x; i < ck.count; i += blockDim.x) {  // tail (or the whole chunk when unaligned)
        float mm = m[i], vv = v[i];
        const float pn = adamw_one(p[i], g[i], mm, vv, h);
        p[i] = pn;
        m[i] = mm;
        v[i] = vv;
        if (sh != nullptr) {
            const int64_t e = ck.offset + i;
            const int64_t r = e / ck.cols;
            sh[r * ck.ld16 + (e - r * ck.cols)] = __float2bfloat16(pn);
        }
    }
}

__global__ void adamw_tick_kernel(float* hyper) {
    pdl_trigger();
    pdl_wait();
    hyper[5] += 1.0f;
}

// flat <-> tensors, one CTA per chunk (the flatten / unflatten around the gradient all-reduce)
__global__ void __launch_bounds__(256) flat_copy_kernel(float* __restrict__ flat, const pio_flat_tensor* __restrict__ tensors,
                                                        const pio_adamw_chunk* __restrict__ chunks, int to_flat, float scale) {
    pdl_trigger();
    pdl_wait();
    const pio_adamw_chunk c = chunks[blockIdx.x];
    const pio_flat_tensor t = tensors[c.tensor];
    float* __restrict__ a = t.tensor + c.offset;
    float* __restrict__ f = flat + t.flat_offset + c.offset;
    const bool vec = ((c.offset & 3) == 0) && ((t.flat_offset & 3) == 0) && ((((uintptr_t)t.tensor | (uintptr_t)flat) & 15) == 0);
    const int n4 = vec ? (c.count >> 2) : 0;
    if (to_flat) {
        for (int i = threadIdx.x; i < n4; i += blockDim.x) reinterpret_cast<float4*>(f)[i] = reinterpret_cast<const float4*>(a)[i];
        for (int i = 4 * n4 + threadIdx.x; i < c.count; i += blockDim.x) f[i] = a[i];
    } else {
        for (int i = threadIdx.x; i < n4; i += blockDim.x) {
            float4 v = reinterpret_cast<const float4*>(f)[i];
            v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
            reinterpret_cast<float4*>(a)[i] = v;
        }
        for (int i = 4 * n4 + threadIdx.x; i < c.count; i += blockDim.x) a[i] = f[i] * scale;
    }
}

// a[i] *= *scale over up to three arrays; every CTA leaves at once when the scale is exactly 1 (what
// loss.backward() hands a loss: the gradients formed in the forward are then already final)
__global__ void __launch_bounds__(256) scale3_kernel(float* __restrict__ a, int64_t na, float* __restrict__ b, int64_t nb,
                                                     float* __restrict__ c, int64_t nc, const float* __restrict__ scale) {
    pdl_trigger();
    pdl_wait();
    const float s = __ldg(scale);
    if (s == 1.0f) return;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int64_t i = i0; i < na; i += stride) a[i] *= s;
    for (int64_t i = i0; i < nb; i += stride) b[i] *= s;
    for (int64_t i = i0; i < nc; i += stride) c[i] *= s;
}

}  // namespace pio

using namespace pio;

extern "C" {

int pio_scale3_f32(float* a, int64_t na, float* b, int64_t nb, float* c, int64_t nc, const float* scale, pio_stream_t stream) {
    PIO_CHECK_ARG(scale && na >= 0 && nb >= 0 && nc >= 0 && (a || na == 0) && (b || nb == 0) && (c || nc == 0),
                  "pio_scale3_f32: bad arguments");
    const int64_t most = std::max(na, std::max(nb, nc));
    if (most == 0) return PIO_OK;
    const int blocks = (int)std::min<int64_t>(cdiv(most, 256 * 8), (int64_t)num_sms() * 8);
    PIO_CUDA(launch_pdl(scale3_kernel, dim3(std::max(blocks, 1)), dim3(256), 0, (cudaStream_t)stream, a, na, b, nb, c, nc, scale));
    count_launch();
    return PIO_OK;
}

int pio_adamw_step(const pio_adamw_tensor* tensors, const pio_adamw_chunk* chunks, int n_chunks, const float* hyper,
                   pio_stream_t stream) {
    PIO_CHECK_ARG(n_chunks >= 0 && (n_chunks == 0 || (chunks && tensors)) && hyper, "pio_adamw_step: bad arguments");
    if (n_chunks == 0) return PIO_OK;
    PIO_CUDA(launch_pdl(adamw_kernel, dim3((unsigned)n_chunks), dim3(256), 0, (cudaStream_t)stream, tensors, chunks, hyper));
    count_launch();
    return PIO_OK;
}

int pio_flat_copy_f32(float* flat, const pio_flat_tensor* tensors, const pio_adamw_chunk* chunks, int n_chunks, int to_flat,
                      float scale, pio_stream_t stream) {
    PIO_CHECK_ARG(n_chunks >= 0 && (n_chunks == 0 || (flat && tensors && chunks)), "pio_flat_copy_f32: bad arguments");
    if (n_chunks == 0) return PIO_OK;
    PIO_CUDA(launch_pdl(flat_copy_kernel, dim3((unsigned)n_chunks), dim3(256), 0, (cudaStream_t)stream, flat, tensors, chunks, to_flat,
                        scale));
    count_launch();
    return PIO_OK;
}

int pio_adamw_tick(float* hyper, pio_stream_t stream) {
    PIO_CHECK_ARG(hyper != nullptr, "pio_adamw_tick: null pointer");
    PIO_CUDA(launch_pdl(adamw_tick_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream, hyper));
    count_launch();
    return PIO_OK;
}

}  // extern "C"
