// HBM-bound kernels of the path: image encode (concat with the Fourier table), Fourier table
// generator, text embedding fwd/bwd, MLM masking select/apply, LayerNorm fwd/bwd, partial
// reductions, cross-entropy fwd/bwd.  All coalesced along the channel axis.
#include <stdarg.h>
#include <stdlib.h>
#include <atomic>
#include <mutex>

#include "pio_common.cuh"

namespace pio {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

bool pdl_enabled(cudaStream_t st) {
    static const bool on = getenv("PIO_NO_PDL") == nullptr;
    if (!on) return false;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) return false;
    return cs == cudaStreamCaptureStatusNone;
}

int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// ---------------------------------------------------------------------------------------------
// image encode: one thread per output element, consecutive threads along the channel axis of a
// token row so both the table read and the output write are fully coalesced.
// ---------------------------------------------------------------------------------------------
__global__ void image_encode_kernel(const float* __restrict__ x, const float* __restrict__ pos,
                                    float* __restrict__ out, int64_t total, int64_t m, int c_img, int c_pos) {
    const int c_out = c_img + c_pos;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / c_out;  // b*m + token
        const int c = (int)(i - row * c_out);
        float v;
        if (c < c_img) v = __ldg(x + row * c_img + c);
        else v = __ldg(pos + (row % m) * c_pos + (c - c_img));
        out[i] = v;
    }
}

// Fourier table on the device (adapter.py:53-94): thread per (token, channel).
struct FourierShape {
    int dim[4];
    int ndim;
};
__global__ void fourier_table_kernel(float* __restrict__ out, FourierShape s, int bands, int64_t tokens) {
    const int c_pos = s.ndim * (2 * bands + 1);
    const int64_t total = tokens * c_pos;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = i / c_pos;
        const int c = (int)(i - t * c_pos);
        int idx[4];
        int64_t r = t;
        for (int d = s.ndim - 1; d >= 0; --d) {
            idx[d] = (int)(r % s.dim[d]);
            r /= s.dim[d];
        }
        int d, band = -1, kind;  // kind 0 = position, 1 = sin, 2 = cos
        if (c < s.ndim) {
            d = c;
            kind = 0;
        } else {
            const int cc = c - s.ndim;
            kind = 1 + cc / (s.ndim * bands);
            const int rem = cc % (s.ndim * bands);
            d = rem / bands;
            band = rem % bands;
        }
        const int n = s.dim[d];
        // torch.linspace(-1, 1, n): symmetric evaluation, step = 2/(n-1)
        float p;
        if (n == 1) p = -1.0f;
        else {
            const float step = 2.0f / (float)(n - 1);
            p = (idx[d] < n / 2) ? (-1.0f + step * (float)idx[d]) : (1.0f - step * (float)(n - 1 - idx[d]));
        }
        float v = p;
        if (kind != 0) {
            float f;
            if (bands == 1) f = 1.0f;
            else {
                const float hi = (float)n / 2.0f;
                const float fs = (hi - 1.0f) / (float)(bands - 1);
                f = (band < bands / 2) ? (1.0f + fs * (float)band) : (hi - fs * (float)(bands - 1 - band));
            }
            const float a = 3.14159265358979323846f * (p * f);
            v = (kind == 1) ? sinf(a) : cosf(a);
        }
        out[i] = v;
    }
}

// ---------------------------------------------------------------------------------------------
// text embedding
// ---------------------------------------------------------------------------------------------
__global__ void text_embed_fwd_kernel(const int64_t* __restrict__ ids, const float* __restrict__ emb,
                                      const float* __restrict__ pos, float* __restrict__ out,
                                      int64_t tokens, int64_t l, int c, int64_t vocab, float scale) {
    // one warp per token row
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t t = warp; t < tokens; t += nwarps) {
        int64_t id = ids[t];
        id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
        const float* e = emb + id * c;
        const float* p = pos + (t % l) * c;
        float* o = out + t * c;
        for (int j = lane; j < c; j += 32) o[j] = __ldg(e + j) * scale + __ldg(p + j);
    }
}

// d_embedding[ids[t]] += dout[t] * scale.  MLM batches repeat a few ids thousands of times ([MASK] alone is
// ~12 % of the tokens), and same-address atomics serialise in L2, so every block first folds its tokens into a
// small direct-mapped cache of embedding rows in shared memory (tag = id; a token whose slot is taken by
// another id goes straight to global memory) and flushes the cache with one atomic per element at the end.
constexpr int EMB_SLOTS = 16;
constexpr int EMB_THREADS = 1024;  // 32 warps, one token each at a time: the kernel is a chain of dependent loads per token
                                   // (id -> tag -> row), so it lives on how many rows are in flight (56 us with 8 warps)
__global__ void __launch_bounds__(EMB_THREADS) text_embed_bwd_emb_kernel(const int64_t* __restrict__ ids, const float* __restrict__ dout,
                                                                         float* __restrict__ demb, int64_t tokens, int c, int64_t vocab,
                                                                         float scale, int slots) {
    extern __shared__ float s_cache[];  // [slots][c]
    __shared__ long long s_tag[EMB_SLOTS];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int i = threadIdx.x; i < slots * c; i += blockDim.x) s_cache[i] = 0.f;
    if (threadIdx.x < EMB_SLOTS) s_tag[threadIdx.x] = -1;
    __syncthreads();
    const bool vec4 = (c & 3) == 0 && c <= 512 && ((((uintptr_t)dout) | ((uintptr_t)demb)) & 15) == 0;
    const int64_t per = (tokens + gridDim.x - 1) / gridDim.x;  // a contiguous run of tokens per block
    const int64_t t_end = min(tokens, (int64_t)(blockIdx.x + 1) * per);
    for (int64_t t = (int64_t)blockIdx.x * per + wid; t < t_end; t += nw) {
        int64_t id = ids[t];
        id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
        const float* g = dout + t * c;
        float4 v[4];  // (vec4) the row, read before the slot is known: 16 bytes per lane and 128 channels
        if (vec4) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (lane * 4 + 128 * k < c) v[k] = *reinterpret_cast<const float4*>(g + lane * 4 + 128 * k);
        }
        int slot = -1;
        if (slots > 0) {
            // the lowest ids are the tokenizer's special tokens ([PAD], [UNK], [MASK], ...: reference tokenizer.py puts
            // them first) and by far the most frequent: they get slots of their own, everything else shares the rest
            const int reserved = slots >= 16 ? 8 : 0;
            const int cand = (id < reserved) ? (int)id
                                             : reserved + (int)((((uint64_t)id * 0x9E3779B97F4A7C15ull) >> 40) % (uint64_t)(slots - reserved));
            long long old = 0;
            if (lane == 0) old = atomicCAS((unsigned long long*)&s_tag[cand], (unsigned long long)-1ll, (unsigned long long)id);
            old = __shfl_sync(0xffffffffu, old, 0);
            if (old == -1 || old == id) slot = cand;
        }
        if (vec4) {
            float* d = (slot >= 0 ? s_cache + (size_t)slot * c : demb + id * c) + lane * 4;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (lane * 4 + 128 * k < c) {
                    if (slot >= 0) {
                        atomicAdd(d + 128 * k, v[k].x * scale);
                        atomicAdd(d + 128 * k + 1, v[k].y * scale);
                        atomicAdd(d + 128 * k + 2, v[k].z * scale);
                        atomicAdd(d + 128 * k + 3, v[k].w * scale);
                    } else {  // most tokens of a block are distinct ids and end up here: one 16-byte reduction per lane
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d + 128 * k), "f"(v[k].x * scale), "f"(v[k].y * scale),
                                     "f"(v[k].z * scale), "f"(v[k].w * scale)
                                     : "memory");
                    }
                }
            }
        } else {
            float* d = slot >= 0 ? s_cache + (size_t)slot * c : demb + id * c;
            for (int j = lane; j < c; j += 32) atomicAdd(d + j, g[j] * scale);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < slots * c; i += blockDim.x) {
        const long long id = s_tag[i / c];
        if (id >= 0) atomicAdd(demb + id * c + (i % c), s_cache[i]);
    }
}

__global__ void text_embed_bwd_pos_kernel(const float* __restrict__ dout, float* __restrict__ dpos,
                                          int64_t batch, int64_t l, int c) {
    const int64_t total = l * c;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int64_t b = 0; b < batch; ++b) acc += dout[b * total + i];
        dpos[i] = acc;
    }
}

// ---------------------------------------------------------------------------------------------
// MLM masking (integer / boolean, bit-exact)
// ---------------------------------------------------------------------------------------------
__global__ void text_mask_select_kernel(const int64_t* __restrict__ ids, const uint8_t* __restrict__ pad,
                                        const float* __restrict__ r0, const float* __restrict__ r1,
                                        const float* __restrict__ r2, uint8_t* __restrict__ flags,
                                        int32_t* __restrict__ block_counts, int64_t n, int64_t unk, float mask_p) {
    // block = PIO_MASK_BLOCK elements, 256 threads x 4
    __shared__ int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const float p90 = 0.9f;
    const float p19 = (float)(1.0 / 9.0);
    int local = 0;
    const int64_t base = (int64_t)blockIdx.x * PIO_MASK_BLOCK;
    for (int j = threadIdx.x; j < PIO_MASK_BLOCK; j += blockDim.x) {
        const int64_t i = base + j;
        if (i >= n) break;
        const bool special = (ids[i] == unk) || (pad != nullptr && pad[i] != 0);
        const bool sel = (r0[i] < mask_p) && !special;
        const bool sel1 = sel && (r1[i] < p90);
        const bool sel2 = sel1 && (r2[i] < p19);
        flags[i] = (uint8_t)((sel ? 1 : 0) | (sel1 ? 2 : 0) | (sel2 ? 4 : 0));
        local += sel2 ? 1 : 0;
    }
    local = (int)warp_sum((float)local);  // <= 1024, exact in fp32
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&s_cnt, local);
    __syncthreads();
    if (threadIdx.x == 0) block_counts[blockIdx.x] = s_cnt;
}

__global__ void text_mask_apply_kernel(int64_t* __restrict__ ids, int64_t* __restrict__ labels,
                                       const uint8_t* __restrict__ flags, const int32_t* __restrict__ block_counts,
                                       const int64_t* __restrict__ random_ids, int64_t n_random, int64_t n,
                                       int64_t mask_id) {
    // one block of 1024 threads per PIO_MASK_BLOCK elements; rank of each bit2 element = (sum of
    // previous blocks' counts) + in-block exclusive scan, i.e. row-major order like x[mask] = values.
    __shared__ int s_warp[32];
    __shared__ long long s_base;
    const int tid = threadIdx.x;
    const int lane = tid & 31, wid = tid >> 5;
    // base offset
    long long part = 0;
    for (int j = tid; j < (int)blockIdx.x; j += blockDim.x) part += block_counts[j];
    // block reduce (values are small integers; use integer shuffles)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    __shared__ long long s_part[32];
    if (lane == 0) s_part[wid] = part;
    __syncthreads();
    if (tid == 0) {
        long long t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_part[w];
        s_base = t;
    }
    const int64_t i = (int64_t)blockIdx.x * PIO_MASK_BLOCK + tid;
    const uint8_t f = (i < n) ? flags[i] : 0;
    const int is2 = (f & 4) ? 1 : 0;
    // warp inclusive scan
    int incl = is2;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int v = s_warp[lane];
        int vi = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, vi, o);
            if (lane >= o) vi += t;
        }
        s_warp[lane] = vi - v;  // exclusive
    }
    __syncthreads();
    if (i < n) {
        const int64_t orig = ids[i];
        if (f & 4) {
            const long long rank = s_base + s_warp[wid] + (incl - 1);
            ids[i] = (rank < n_random) ? random_ids[rank] : mask_id;
        } else if (f & 2) {
            ids[i] = mask_id;
        }
        labels[i] = (f & 1) ? orig : (int64_t)-100;
    }
}

// ---------------------------------------------------------------------------------------------
// LayerNorm
// ---------------------------------------------------------------------------------------------
__global__ void layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                     const float* __restrict__ b, float* __restrict__ y,
                                     float* __restrict__ stats, int64_t rows, int c, float eps) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float inv_c = 1.0f / (float)c;
    for (int64_t r = warp; r < rows; r += nwarps) {
        const float* xr = x + r * c;
        float s = 0.f;
        for (int j = lane; j < c; j += 32) s += xr[j];
        const float mean = warp_sum(s) * inv_c;
        float v = 0.f;
        for (int j = lane; j < c; j += 32) {
            const float d = xr[j] - mean;
            v += d * d;
        }
        const float rstd = rsqrtf(warp_sum(v) * inv_c + eps);
        float* yr = y + r * c;
        for (int j = lane; j < c; j += 32) yr[j] = (xr[j] - mean) * rstd * __ldg(w + j) + __ldg(b + j);
        if (lane == 0 && stats != nullptr) {
            stats[2 * r] = mean;
            stats[2 * r + 1] = rstd;
        }
    }
}

// dx per row (warp per row) + per-block sums of dgamma/dbeta, either written as partials
// [gridDim.x][2][c] for a deterministic second pass or added to dgamma/dbeta with one atomic per
// column per block.  The grid gives every warp one row where it can (this kernel is a chain of
// dependent latencies, not a bandwidth problem, at the latent sizes): all three row loads (dy, x and
// the residual gradient) are issued before the first reduction.
template <int NCH>
__global__ void __launch_bounds__(NCH <= 8 ? 512 : 256, NCH <= 16 ? 2 : 1) layernorm_bwd_kernel(
    const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ stats,
    const float* __restrict__ w, const float* __restrict__ dres, float* __restrict__ dx,
    float* __restrict__ partial, int64_t rows, int c, float* __restrict__ atomic_dw = nullptr,
    float* __restrict__ atomic_db = nullptr, const int* __restrict__ dyn_rows = nullptr, __nv_bfloat16* __restrict__ dx16 = nullptr) {
    extern __shared__ float s_acc[];  // [warps][2][c]
    pdl_trigger();
    pdl_wait();
    if (dyn_rows != nullptr) rows = min(rows, (int64_t)*dyn_rows);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    float gw[NCH], gb[NCH];
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        gw[i] = 0.f;
        gb[i] = 0.f;
    }
    const float inv_c = 1.0f / (float)c;
    for (int64_t r = (int64_t)blockIdx.x * nw + wid; r < rows; r += (int64_t)gridDim.x * nw) {
        const float* dyr = dy + r * c;
        const float* xr = x + r * c;
        float d[NCH], xh[NCH], rs[NCH];
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
            const int j = lane + 32 * i;
            const bool in = j < c;
            d[i] = in ? dyr[j] : 0.f;
            xh[i] = in ? xr[j] : 0.f;
            rs[i] = (in && dres != nullptr) ? dres[r * c + j] : 0.f;
        }
        const float mean = stats[2 * r], rstd = stats[2 * r + 1];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
            const int j = lane + 32 * i;
            const bool in = j < c;
            xh[i] = in ? (xh[i] - mean) * rstd : 0.f;
            gw[i] += d[i] * xh[i];
            gb[i] += d[i];
            d[i] *= in ? __ldg(w + j) : 0.f;  // g = dy * gamma
            c1 += d[i];
            c2 += d[i] * xh[i];
        }
        c1 = warp_sum(c1) * inv_c;
        c2 = warp_sum(c2) * inv_c;
        if (dx != nullptr) {
#pragma unroll
            for (int i = 0; i < NCH; ++i) {
                const int j = lane + 32 * i;
                if (j < c) {
                    const float o = rstd * (d[i] - c1 - xh[i] * c2) + rs[i];
                    dx[r * c + j] = o;
                    if (dx16 != nullptr) dx16[r * c + j] = __float2bfloat16(o);
                }
            }
        }
    }
    float* mine = s_acc + (size_t)wid * 2 * c;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        const int j = lane + 32 * i;
        if (j < c) {
            mine[j] = gw[i];
            mine[c + j] = gb[i];
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < 2 * c; j += blockDim.x) {
        float acc = 0.f;
        for (int w2 = 0; w2 < nw; ++w2) acc += s_acc[(size_t)w2 * 2 * c + j];
        if (atomic_dw != nullptr) atomicAdd(j < c ? atomic_dw + j : atomic_db + (j - c), acc);
        else partial[(size_t)blockIdx.x * 2 * c + j] = acc;
    }
}

// The same for rows of NV * 128 channels: one float4 per lane per 128 channels (three 16-byte loads per lane for
// dy, x and the residual gradient instead of twelve 4-byte ones at C = 256), same block-level reduction.
template <int NV>
__global__ void __launch_bounds__(512, 2) layernorm_bwd_vec_kernel(
    const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ stats,
    const float* __restrict__ w, const float* __restrict__ dres, float* __restrict__ dx,
    float* __restrict__ partial, int64_t rows, float* __restrict__ atomic_dw, float* __restrict__ atomic_db,
    const int* __restrict__ dyn_rows, __nv_bfloat16* __restrict__ dx16) {
    extern __shared__ float s_acc[];  // [warps][2][c]
    pdl_trigger();
    pdl_wait();
    if (dyn_rows != nullptr) rows = min(rows, (int64_t)*dyn_rows);
    constexpr int C = NV * 128;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    float4 gw[NV], gb[NV], wv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        gw[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        gb[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        wv[i] = __ldg(reinterpret_cast<const float4*>(w) + i * 32 + lane);
    }
    for (int64_t r = (int64_t)blockIdx.x * nw + wid; r < rows; r += (int64_t)gridDim.x * nw) {
        const float4* dyr = reinterpret_cast<const float4*>(dy + r * C);
        const float4* xr = reinterpret_cast<const float4*>(x + r * C);
        float4 d[NV], xh[NV], rs[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            d[i] = dyr[i * 32 + lane];
            xh[i] = xr[i * 32 + lane];
            rs[i] = dres != nullptr ? reinterpret_cast<const float4*>(dres + r * C)[i * 32 + lane] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const float mean = stats[2 * r], rstd = stats[2 * r + 1];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
#define PIO_LNV(f)                          \
    xh[i].f = (xh[i].f - mean) * rstd;      \
    gw[i].f += d[i].f * xh[i].f;            \
    gb[i].f += d[i].f;                      \
    d[i].f *= wv[i].f;                      \
    c1 += d[i].f;                           \
    c2 += d[i].f * xh[i].f;
            PIO_LNV(x) PIO_LNV(y) PIO_LNV(z) PIO_LNV(w)
#undef PIO_LNV
        }
        c1 = warp_sum(c1) * (1.0f / C);
        c2 = warp_sum(c2) * (1.0f / C);
        if (dx != nullptr) {
            float4* dxr = reinterpret_cast<float4*>(dx + r * C);
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                float4 o;
                o.x = rstd * (d[i].x - c1 - xh[i].x * c2) + rs[i].x;
                o.y = rstd * (d[i].y - c1 - xh[i].y * c2) + rs[i].y;
                o.z = rstd * (d[i].z - c1 - xh[i].z * c2) + rs[i].z;
                o.w = rstd * (d[i].w - c1 - xh[i].w * c2) + rs[i].w;
                dxr[i * 32 + lane] = o;
                if (dx16 != nullptr) {  // the same row as bf16: the next operator's dY operand (saves its cast pass)
                    __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
                    uint2 u;
                    u.x = *reinterpret_cast<uint32_t*>(&lo);
                    u.y = *reinterpret_cast<uint32_t*>(&hi);
                    reinterpret_cast<uint2*>(dx16 + r * C)[i * 32 + lane] = u;
                }
            }
        }
    }
    float* mine = s_acc + (size_t)wid * 2 * C;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        reinterpret_cast<float4*>(mine)[i * 32 + lane] = gw[i];
        reinterpret_cast<float4*>(mine + C)[i * 32 + lane] = gb[i];
    }
    __syncthreads();
    for (int j = threadIdx.x; j < 2 * C; j += blockDim.x) {
        float acc = 0.f;
        for (int w2 = 0; w2 < nw; ++w2) acc += s_acc[(size_t)w2 * 2 * C + j];
        if (atomic_dw != nullptr) atomicAdd(j < C ? atomic_dw + j : atomic_db + (j - C), acc);
        else partial[(size_t)blockIdx.x * 2 * C + j] = acc;
    }
}

// out1[i] (+)= sum_s part[s*stride + i], i < n1 ; out2[i] (+)= sum_s part[s*stride + n1 + i], i < n2
__global__ void reduce_partials_kernel(const float* __restrict__ part, int splits, int64_t stride,
                                       float* __restrict__ out1, int64_t n1, float* __restrict__ out2,
                                       int64_t n2, int accumulate) {
    pdl_trigger();
    pdl_wait();
    const int64_t total = n1 + n2;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int s = 0; s < splits; ++s) acc += part[(int64_t)s * stride + i];
        float* dst = (i < n1) ? (out1 + i) : (out2 + (i - n1));
        if (accumulate) acc += *dst;
        *dst = acc;
    }
}

// The same reduction for few outputs and many partial rows (LayerNorm / bias gradients): a block owns 32
// outputs, its 32 warps stride over the partial rows, shared memory folds the 32 sub-sums in a fixed order.
__global__ void __launch_bounds__(1024) reduce_partials_tall_kernel(const float* __restrict__ part, int splits, int64_t stride,
                                                                    float* __restrict__ out1, int64_t n1,
                                                                    float* __restrict__ out2, int64_t n2, int accumulate) {
    __shared__ float sub[32][33];
    pdl_trigger();
    pdl_wait();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 32 + tx;
    const int64_t total = n1 + n2;
    float acc = 0.f;
    if (i < total)
        for (int s = ty; s < splits; s += 32) acc += part[(int64_t)s * stride + i];
    sub[ty][tx] = acc;
    __syncthreads();
    if (ty == 0 && i < total) {
        float v = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) v += sub[k][tx];
        float* dst = (i < n1) ? (out1 + i) : (out2 + (i - n1));
        if (accumulate) v += *dst;
        *dst = v;
    }
}

int reduce_partials(const float* part, int splits, int64_t stride, float* out1, int64_t n1, float* out2,
                    int64_t n2, int accumulate, cudaStream_t st) {
    const int64_t total = n1 + n2;
    if (total <= 0) return PIO_OK;
    if (splits >= 32 && total <= 32 * (int64_t)num_sms() * 2) {
        PIO_CUDA(launch_pdl(reduce_partials_tall_kernel, dim3((unsigned)cdiv(total, 32)), dim3(1024), 0, st, part, splits, stride, out1,
                            n1, out2, n2, accumulate));
        count_launch();
        return PIO_OK;
    }
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(cdiv(total, threads), (int64_t)num_sms() * 8);
    PIO_CUDA(launch_pdl(reduce_partials_kernel, dim3(blocks), dim3(threads), 0, st, part, splits, stride, out1, n1, out2, n2, accumulate));
    count_launch();
    return PIO_OK;
}

// out[i] = keep(i) ? in[i] / (1-p) : 0 — the gradient of Residual's dropout w.r.t. its branch
__global__ void dropout_mask_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t n, DropCfg d) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = drop_keep(d, (uint64_t)i) ? in[i] * d.inv_keep : 0.f;
}

int dropout_mask(const float* in, float* out, int64_t n, const DropCfg& d, cudaStream_t st) {
    if (n <= 0) return PIO_OK;
    const int blocks = (int)std::min<int64_t>(cdiv(n, 256), (int64_t)num_sms() * 16);
    dropout_mask_kernel<<<blocks, 256, 0, st>>>(in, out, n, d);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

// ---------------------------------------------------------------------------------------------
// cross entropy
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ce_fwd_rows_kernel(const float* __restrict__ logits,
                                                          const int64_t* __restrict__ labels,
                                                          float* __restrict__ row_lse, float* __restrict__ row_loss,
                                                          int64_t rows, int64_t classes, int64_t ld, int64_t ignore) {
    __shared__ float s_m[8], s_s[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
        const float* z = logits + r * ld;
        float m = -INFINITY, s = 0.f;
        for (int64_t j = threadIdx.x; j < classes; j += blockDim.x) {
            const float v = z[j];
            if (v > m) {
                s = s * expf(m - v) + 1.0f;
                m = v;
            } else {
                s += expf(v - m);
            }
        }
        // warp combine
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
            const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
            const float mm = fmaxf(m, m2);
            const float a = (m == -INFINITY) ? 0.f : s * expf(m - mm);
            const float b = (m2 == -INFINITY) ? 0.f : s2 * expf(m2 - mm);
            m = mm;
            s = a + b;
        }
        if (lane == 0) {
            s_m[wid] = m;
            s_s[wid] = s;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float mm = -INFINITY;
            for (int w = 0; w < 8; ++w) mm = fmaxf(mm, s_m[w]);
            float ss = 0.f;
            for (int w = 0; w < 8; ++w)
                if (s_m[w] != -INFINITY) ss += s_s[w] * expf(s_m[w] - mm);
            const float lse = mm + logf(ss);
            row_lse[r] = lse;
            const int64_t y = labels[r];
            row_loss[r] = (y == ignore || y < 0 || y >= classes) ? 0.f : (lse - z[y]);
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(1024) ce_finalize_kernel(const float* __restrict__ row_loss,
                                                           const int64_t* __restrict__ labels, float* __restrict__ out3,
                                                           int64_t rows, int64_t classes, int64_t ignore) {
    __shared__ double s_l[32];
    __shared__ double s_n[32];
    double l = 0.0, n = 0.0;
    for (int64_t r = threadIdx.x; r < rows; r += blockDim.x) {
        const int64_t y = labels[r];
        if (!(y == ignore || y < 0 || y >= classes)) {
            l += (double)row_loss[r];
            n += 1.0;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        l += __shfl_xor_sync(0xffffffffu, l, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
    }
    if ((threadIdx.x & 31) == 0) {
        s_l[threadIdx.x >> 5] = l;
        s_n[threadIdx.x >> 5] = n;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double tl = 0.0, tn = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
            tl += s_l[w];
            tn += s_n[w];
        }
        out3[0] = (float)(tl / tn);  // 0/0 = NaN like torch when nothing is supervised
        out3[1] = (float)tn;
        out3[2] = (float)tl;
    }
}

__global__ void ce_bwd_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels,
                              const float* __restrict__ out3, const float* __restrict__ row_lse,
                              const float* __restrict__ dloss, float* __restrict__ dlogits, int64_t rows,
                              int64_t classes, int64_t ld, int64_t ignore) {
    const float scale = dloss[0] / out3[1];
    for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) {
        const int64_t y = labels[r];
        const bool valid = !(y == ignore || y < 0 || y >= classes);
        const float lse = row_lse[r];
        const float* z = logits + r * ld;
        float* d = dlogits + r * ld;
        if (!valid) {
            for (int64_t j = threadIdx.x; j < classes; j += blockDim.x) d[j] = 0.f;
        } else {
            for (int64_t j = threadIdx.x; j < classes; j += blockDim.x) {
                float p = expf(z[j] - lse);
                if (j == y) p -= 1.0f;
                d[j] = p * scale;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Scratch for pio_layernorm_bwd_noalloc (see that entry point below): the per-block partial sums of
// dgamma / dbeta, held in device memory that is part of the loaded library image instead of a
// caller-provided workspace.  PIO_LN_NOALLOC_BLOCKS x 2 x PIO_LN_NOALLOC_MAX_C floats = 1 MiB,
// resident on every GPU that loads the library until the process exits.
// ---------------------------------------------------------------------------------------------
constexpr int PIO_LN_NOALLOC_BLOCKS = 128;
static __device__ __align__(256) float g_ln_noalloc_scratch[PIO_LN_NOALLOC_BLOCKS * 2 * PIO_LN_NOALLOC_MAX_C];

// On the host the array's name is only a handle for the runtime, not a device address: a kernel
// handed &g_ln_noalloc_scratch[0] from host code would fault (or scribble on host-reachable memory).
// The device address is resolved once per device, with that device current, and cached.
static int ln_noalloc_scratch(float** out) {
    static std::mutex mu;
    static float* per_device[64] = {nullptr};
    int dev = -1;
    PIO_CUDA(cudaGetDevice(&dev));
    PIO_CHECK_ARG(dev >= 0 && dev < 64, "pio_layernorm_bwd_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (per_device[dev] == nullptr) {
        float* ptr = nullptr;
        PIO_CUDA(cudaGetSymbolAddress((void**)&ptr, g_ln_noalloc_scratch));
        PIO_CHECK_ARG(ptr != nullptr, "pio_layernorm_bwd_noalloc: could not resolve the scratch array on device %d", dev);
        per_device[dev] = ptr;
    }
    *out = per_device[dev];
    return PIO_OK;
}

}  // namespace pio

using namespace pio;

extern "C" {

const char* pio_last_error(void) { return pio::g_err; }
int pio_version(void) { return 100; }
long long pio_launch_count(void) { return pio::g_launches.load(std::memory_order_relaxed); }

int pio_image_encode(const float* x, const float* pos_enc, float* out, int64_t batch, int64_t m, int c_img,
                     int c_pos, pio_stream_t stream) {
    PIO_CHECK_ARG(batch >= 0 && m >= 0 && c_img > 0 && c_pos >= 0, "pio_image_encode: bad dims");
    const int64_t total = batch * m * (c_img + c_pos);
    if (total == 0) return PIO_OK;  // empty batch: torch hands out null data pointers
    PIO_CHECK_ARG(x && pos_enc && out, "pio_image_encode: null pointer");
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(cdiv(total, threads * 4), (int64_t)num_sms() * 16);
    image_encode_kernel<<<std::max(blocks, 1), threads, 0, (cudaStream_t)stream>>>(x, pos_enc, out, total, m, c_img, c_pos);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

int pio_fourier_table(float* out, const int* shape, int ndim, int bands, pio_stream_t stream) {
    PIO_CHECK_ARG(out && shape && ndim >= 1 && ndim <= 4 && bands >= 1, "pio_fourier_table: bad arguments");
    FourierShape s;
    int64_t tokens = 1;
    for (int d = 0; d < 4; ++d) s.dim[d] = 1;
    for (int d = 0; d < ndim; ++d) {
        PIO_CHECK_ARG(shape[d] >= 1, "pio_fourier_table: bad shape");
        s.dim[d] = shape[d];
        tokens *= shape[d];
    }
    s.ndim = ndim;
    const int64_t total = tokens * ndim * (2 * bands + 1);
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(cdiv(total, threads), (int64_t)num_sms() * 16);
    fourier_table_kernel<<<std::max(blocks, 1), threads, 0, (cudaStream_t)stream>>>(out, s, bands, tokens);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

int pio_text_embed_fwd(const int64_t* ids, const float* emb, const float* pos, float* out, int64_t batch,
                       int64_t l, int c, int64_t vocab, float scale, pio_stream_t stream) {
    PIO_CHECK_ARG(ids && emb && pos && out && c > 0 && vocab > 0, "pio_text_embed_fwd: bad arguments");
    const int64_t tokens = batch * l;
    if (tokens == 0) return PIO_OK;
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(cdiv(tokens, threads / 32), (int64_t)num_sms() * 16);
    text_embed_fwd_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(ids, emb, pos, out, tokens, l, c, vocab, scale);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

int pio_text_embed_bwd(const int64_t* ids, const float* dout, float* demb, float* dpos, int64_t batch, int64_t l,
                       int c, int64_t vocab, float scale, pio_stream_t stream) {
    PIO_CHECK_ARG(ids && dout && c > 0 && vocab > 0, "pio_text_embed_bwd: bad arguments");
    const int64_t tokens = batch * l;
    if (tokens == 0) return PIO_OK;
    const int threads = 256;
    if (demb) {
        // blocks own >= 64 tokens each so the per-block cache has duplicates to absorb
        const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(cdiv(tokens, 64), (int64_t)num_sms() * 2));
        const int slots = c <= 512 ? EMB_SLOTS : (c <= 2048 ? 4 : 0);
        text_embed_bwd_emb_kernel<<<blocks, EMB_THREADS, (size_t)slots * c * sizeof(float), (cudaStream_t)stream>>>(
            ids, dout, demb, tokens, c, vocab, scale, slots);
        PIO_LAUNCH_CHECK();
    }
    if (dpos) {
        const int blocks = (int)std::min<int64_t>(cdiv(l * c, threads), (int64_t)num_sms() * 16);
        text_embed_bwd_pos_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(dout, dpos, batch, l, c);
        PIO_LAUNCH_CHECK();
    }
    return PIO_OK;
}

int pio_text_mask_select(const int64_t* ids, const uint8_t* pad, const float* r0, const float* r1, const float* r2,
                         uint8_t* flags, int32_t* block_counts, int64_t n, int64_t unk_token_id, float mask_p,
                         pio_stream_t stream) {
    PIO_CHECK_ARG(ids && r0 && r1 && r2 && flags && block_counts && n >= 0, "pio_text_mask_select: bad arguments");
    if (n == 0) return PIO_OK;
    const int blocks = (int)cdiv(n, PIO_MASK_BLOCK);
    text_mask_select_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(ids, pad, r0, r1, r2, flags, block_counts, n,
                                                                      unk_token_id, mask_p);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

int pio_text_mask_apply(int64_t* ids, int64_t* labels, const uint8_t* flags, const int32_t* block_counts,
                        const int64_t* random_ids, int64_t n_random, int64_t n, int64_t mask_token_id,
                        pio_stream_t stream) {
    PIO_CHECK_ARG(ids && labels && flags && block_counts && n >= 0, "pio_text_mask_apply: bad arguments");
    PIO_CHECK_ARG(n_random == 0 || random_ids != nullptr, "pio_text_mask_apply: random_ids missing");
    if (n == 0) return PIO_OK;
    const int blocks = (int)cdiv(n, PIO_MASK_BLOCK);
    text_mask_apply_kernel<<<blocks, PIO_MASK_BLOCK, 0, (cudaStream_t)stream>>>(ids, labels, flags, block_counts,
                                                                                 random_ids, n_random, n, mask_token_id);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

int pio_layernorm_fwd(const float* x, const float* w, const float* b, float* y, float* stats, int64_t rows, int c,
                      float eps, pio_stream_t stream) {
    PIO_CHECK_ARG(x && w && b && y && c > 0 && rows >= 0, "pio_layernorm_fwd: bad arguments");
    if (rows == 0) return PIO_OK;
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(cdiv(rows, threads / 32), (int64_t)num_sms() * 16);
    layernorm_fwd_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(x, w, b, y, stats, rows, c, eps);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

static int ln_bwd_warps(int c) { return c <= 256 ? 16 : 8; }  // matches the kernel's launch bounds per NCH
static int ln_bwd_blocks(int64_t rows, int c) {
    return (int)std::max<int64_t>(1, std::min<int64_t>(cdiv(rows, ln_bwd_warps(c)), (int64_t)pio::num_sms() * 4));
}

size_t pio_layernorm_bwd_ws_bytes(int64_t rows, int c) {
    return align_up((size_t)ln_bwd_blocks(rows, c) * 2 * c * sizeof(float), 256) + 256;
}

}  // extern "C"
namespace pio {
int layernorm_bwd(const float* dy, const float* x, const float* stats, const float* w, const float* dres, float* dx, float* dw,
                  float* db, int accumulate, int64_t rows, int c, void* ws, size_t ws_bytes, cudaStream_t stream,
                  const int* dyn_rows, __nv_bfloat16* dx16);
}
extern "C" {
int pio_layernorm_bwd(const float* dy, const float* x, const float* stats, const float* w, const float* dres,
                      float* dx, float* dw, float* db, int accumulate, int64_t rows, int c, void* ws, size_t ws_bytes,
                      pio_stream_t stream) {
    return pio::layernorm_bwd(dy, x, stats, w, dres, dx, dw, db, accumulate, rows, c, ws, ws_bytes, (cudaStream_t)stream, nullptr,
                              nullptr);
}
}  // extern "C"

namespace pio {
int layernorm_bwd(const float* dy, const float* x, const float* stats, const float* w, const float* dres, float* dx, float* dw,
                  float* db, int accumulate, int64_t rows, int c, void* ws, size_t ws_bytes, cudaStream_t stream,
                  const int* dyn_rows, __nv_bfloat16* dx16) {
    PIO_CHECK_ARG(dy && x && stats && w && dw && db && c > 0 && rows >= 0, "pio_layernorm_bwd: bad arguments");
    PIO_CHECK_ARG(dx16 == nullptr || dx != nullptr, "pio_layernorm_bwd: the bf16 copy needs dx");
    PIO_CHECK_ARG(c <= 1024, "pio_layernorm_bwd: channels %d > 1024 unsupported", c);
    if (rows == 0) return PIO_OK;
    Workspace W(ws, ws_bytes);
    const int blocks = ln_bwd_blocks(rows, c);
    const bool atomic = (accumulate == 2);  // dw / db hold a running sum (or zeros): add with atomics, no reduction pass
    float* partial = atomic ? nullptr : W.takef((size_t)blocks * 2 * c);
    if (!partial && !atomic) {
        set_error("pio_layernorm_bwd: workspace too small (%zu bytes)", ws_bytes);
        return PIO_ERR_WORKSPACE;
    }
    float* adw = atomic ? dw : nullptr;
    float* adb = atomic ? db : nullptr;
    cudaStream_t st = (cudaStream_t)stream;
    const int warps = ln_bwd_warps(c);
    const size_t smem = (size_t)warps * 2 * c * sizeof(float);
    const bool vec_ok = (c == 128 || c == 256) &&
                        ((((uintptr_t)dy | (uintptr_t)x | (uintptr_t)w | (uintptr_t)dres | (uintptr_t)dx) & 15) == 0);
    if (vec_ok) {  // warps == 16 for these widths; smem = 16 * 2 * c * 4 <= 32 KB
        if (c == 128)
            PIO_CUDA(launch_pdl(layernorm_bwd_vec_kernel<1>, dim3(blocks), dim3(32 * warps), smem, st, dy, x, stats, w, dres, dx, partial, rows, adw, adb, dyn_rows, dx16));
        else
            PIO_CUDA(launch_pdl(layernorm_bwd_vec_kernel<2>, dim3(blocks), dim3(32 * warps), smem, st, dy, x, stats, w, dres, dx, partial, rows, adw, adb, dyn_rows, dx16));
        PIO_LAUNCH_CHECK();
        if (atomic) return PIO_OK;
        return reduce_partials(partial, blocks, (int64_t)2 * c, dw, c, db, c, accumulate, st);
    }
#define PIO_LN_BWD(NCH)                                                                                         \
    do {                                                                                                        \
        if (smem > 48 * 1024)                                                                                   \
            PIO_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          (int)smem));                                                          \
        PIO_CUDA(launch_pdl(layernorm_bwd_kernel<NCH>, dim3(blocks), dim3(32 * warps), smem, st, dy, x, stats, w, dres, dx, partial, rows, c, adw, \
                            adb, dyn_rows, dx16));                                                              \
    } while (0)
    if (c <= 128) PIO_LN_BWD(4);
    else if (c <= 256) PIO_LN_BWD(8);
    else if (c <= 512) PIO_LN_BWD(16);
    else PIO_LN_BWD(32);
#undef PIO_LN_BWD
    PIO_LAUNCH_CHECK();
    if (atomic) return PIO_OK;
    // partial rows are [dw(c) | db(c)] : reduce both with one launch
    return reduce_partials(partial, blocks, (int64_t)2 * c, dw, c, db, c, accumulate, st);
}
}  // namespace pio

extern "C" {

// LayerNorm backward for callers that cannot allocate device memory.
//
// Same arithmetic and arguments as pio_layernorm_bwd, minus the workspace: the partial sums go to
// the library's own static device array.  Meant for callers that have nowhere to get scratch from —
// code that runs while a stream is being captured or inside a hook where cudaMalloc (and torch's
// caching allocator, which can fall through to it) must not be called, plugin hosts that pass a
// kernel library only their input and output buffers, and embedders that must not grow device
// memory after start-up.
//
// Contract:
//   * c <= PIO_LN_NOALLOC_MAX_C, otherwise PIO_ERR_ARG with a message saying so (nothing launched);
//   * the first call on each device resolves the array's device address (cudaGetSymbolAddress);
//     make that first call outside any stream capture;
//   * every call on a device uses the same array: two calls whose kernels can overlap on the GPU
//     (e.g. on different streams) race — serialise them;
//   * the array keeps whatever the previous call left; that is harmless, because every partial row
//     the final reduction reads is written in full by the same call first (nothing needs zeroing);
//   * no cudaMalloc / cudaFree / allocating library routine is called, here or underneath.
int pio_layernorm_bwd_noalloc(const float* dy, const float* x, const float* stats, const float* w, const float* dres,
                              float* dx, float* dw, float* db, int accumulate, int64_t rows, int c, pio_stream_t stream) {
    PIO_CHECK_ARG(dy && x && stats && w && dw && db && c > 0 && rows >= 0, "pio_layernorm_bwd_noalloc: bad arguments");
    PIO_CHECK_ARG(c <= PIO_LN_NOALLOC_MAX_C,
                  "pio_layernorm_bwd_noalloc: %d channels exceed the built-in scratch (at most %d); use pio_layernorm_bwd "
                  "with a caller-provided workspace",
                  c, PIO_LN_NOALLOC_MAX_C);
    if (rows == 0) return PIO_OK;
    float* partial = nullptr;
    PIO_TRY(ln_noalloc_scratch(&partial));
    const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(cdiv(rows, 8), PIO_LN_NOALLOC_BLOCKS));
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = (size_t)8 * 2 * c * sizeof(float);
#define PIO_LN_BWD_NA(NCH)                                                                                      \
    do {                                                                                                        \
        if (smem > 48 * 1024)                                                                                   \
            PIO_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          (int)smem));                                                          \
        layernorm_bwd_kernel<NCH><<<blocks, 256, smem, st>>>(dy, x, stats, w, dres, dx, partial, rows, c);       \
    } while (0)
    if (c <= 128) PIO_LN_BWD_NA(4);
    else if (c <= 256) PIO_LN_BWD_NA(8);
    else if (c <= 512) PIO_LN_BWD_NA(16);
    else PIO_LN_BWD_NA(32);
#undef PIO_LN_BWD_NA
    PIO_LAUNCH_CHECK();
    return reduce_partials(partial, blocks, (int64_t)2 * c, dw, c, db, c, accumulate, st);
}

int pio_cross_entropy_fwd(const float* logits, const int64_t* labels, float* out3, float* row_lse, int64_t rows,
                          int64_t classes, int64_t ld, int64_t ignore_index, pio_stream_t stream) {
    PIO_CHECK_ARG(logits && labels && out3 && row_lse && rows >= 0 && classes > 0 && ld >= classes,
                  "pio_cross_entropy_fwd: bad arguments");
    // row losses are staged in row_lse's second half: caller provides row_lse of 2*rows floats
    cudaStream_t st = (cudaStream_t)stream;
    float* row_loss = row_lse + rows;
    if (rows > 0) {
        const int blocks = (int)std::min<int64_t>(rows, (int64_t)num_sms() * 8);
        ce_fwd_rows_kernel<<<blocks, 256, 0, st>>>(logits, labels, row_lse, row_loss, rows, classes, ld, ignore_index);
        PIO_LAUNCH_CHECK();
    }
    ce_finalize_kernel<<<1, 1024, 0, st>>>(row_loss, labels, out3, rows, classes, ignore_index);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

int pio_cross_entropy_bwd(const float* logits, const int64_t* labels, const float* out3, const float* row_lse,
                          const float* dloss, float* dlogits, int64_t rows, int64_t classes, int64_t ld,
                          int64_t ignore_index, pio_stream_t stream) {
    PIO_CHECK_ARG(logits && labels && out3 && row_lse && dloss && dlogits && classes > 0 && ld >= classes,
                  "pio_cross_entropy_bwd: bad arguments");
    if (rows == 0) return PIO_OK;
    const int blocks = (int)std::min<int64_t>(rows, (int64_t)num_sms() * 8);
    ce_bwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(logits, labels, out3, row_lse, dloss, dlogits, rows, classes,
                                                            ld, ignore_index);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

}  // extern "C"
