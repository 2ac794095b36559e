// Bandwidth-bound helpers of the bf16 pipeline: LayerNorm with bf16 output, fp32->bf16 cast fused
// with the column sums that give bias gradients, bf16 column sums.
#include "pio_ops_common.cuh"

namespace pio {

__global__ void layernorm_fwd_bf16_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                          const float* __restrict__ b, __nv_bfloat16* __restrict__ y, int64_t ldy,
                                          float* __restrict__ stats, int64_t rows, int c, float eps,
                                          const int* __restrict__ dyn_rows) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float inv_c = 1.0f / (float)c;
    const int64_t alloc_rows = rows;
    if (dyn_rows != nullptr) {  // device-side row count: rows up to the next multiple of 128 are zero-filled (operand padding)
        rows = min(rows, (int64_t)*dyn_rows);
        const int64_t upto = min(alloc_rows, (rows + 127) / 128 * 128);
        for (int64_t r = rows + warp; r < upto; r += nwarps)
            for (int j = lane; j < (int)ldy; j += 32) y[r * ldy + j] = __float2bfloat16(0.f);
    }
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
        __nv_bfloat16* yr = y + r * ldy;
        for (int j = lane; j < (int)ldy; j += 32)
            yr[j] = __float2bfloat16(j < c ? (xr[j] - mean) * rstd * __ldg(w + j) + __ldg(b + j) : 0.f);
        if (lane == 0 && stats != nullptr) {
            stats[2 * r] = mean;
            stats[2 * r + 1] = rstd;
        }
    }
}

// The same for rows of NV * 128 channels with a dense bf16 output (ldy == c): the row lives in registers
// (one float4 per lane per 128 channels), is read once and written as 8-byte bf16 quads.
template <int NV>
__global__ void __launch_bounds__(256) layernorm_fwd_bf16_vec_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                                     const float* __restrict__ b, __nv_bfloat16* __restrict__ y,
                                                                     float* __restrict__ stats, int64_t rows, float eps,
                                                                     const int* __restrict__ dyn_rows) {
    pdl_trigger();
    pdl_wait();
    constexpr int C = NV * 128;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    if (dyn_rows != nullptr) {
        const int64_t alloc_rows = rows;
        rows = min(rows, (int64_t)*dyn_rows);
        const int64_t upto = min(alloc_rows, (rows + 127) / 128 * 128);
        for (int64_t r = rows + warp; r < upto; r += nwarps)
            for (int i = 0; i < NV; ++i) reinterpret_cast<uint2*>(y + r * C)[i * 32 + lane] = make_uint2(0u, 0u);
    }
    float4 wv[NV], bv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        wv[i] = __ldg(reinterpret_cast<const float4*>(w) + i * 32 + lane);
        bv[i] = __ldg(reinterpret_cast<const float4*>(b) + i * 32 + lane);
    }
    for (int64_t r = warp; r < rows; r += nwarps) {
        const float4* xr = reinterpret_cast<const float4*>(x + r * C);
        float4 v[NV];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            v[i] = xr[i * 32 + lane];
            s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
        }
        const float mean = warp_sum(s) * (1.0f / C);
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
            q += (v[i].x * v[i].x + v[i].y * v[i].y) + (v[i].z * v[i].z + v[i].w * v[i].w);
        }
        const float rstd = rsqrtf(warp_sum(q) * (1.0f / C) + eps);
        uint2* yr = reinterpret_cast<uint2*>(y + r * C);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            __nv_bfloat162 lo = __floats2bfloat162_rn(v[i].x * rstd * wv[i].x + bv[i].x, v[i].y * rstd * wv[i].y + bv[i].y);
            __nv_bfloat162 hi = __floats2bfloat162_rn(v[i].z * rstd * wv[i].z + bv[i].z, v[i].w * rstd * wv[i].w + bv[i].w);
            uint2 o;
            o.x = *reinterpret_cast<uint32_t*>(&lo);
            o.y = *reinterpret_cast<uint32_t*>(&hi);
            yr[i * 32 + lane] = o;
        }
        if (lane == 0 && stats != nullptr) {
            stats[2 * r] = mean;
            stats[2 * r + 1] = rstd;
        }
    }
}

int layernorm_fwd_bf16(const float* x, const float* w, const float* b, __nv_bfloat16* y, int64_t ldy, float* stats,
                       int64_t rows, int c, float eps, cudaStream_t st, const int* dyn_rows) {
    PIO_CHECK_ARG(x && w && b && y && c > 0 && ldy >= c, "layernorm_fwd_bf16: bad arguments");
    if (rows == 0) return PIO_OK;
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(cdiv(rows, threads / 32), (int64_t)num_sms() * 16);
    const bool vec_ok = ldy == c && (c % 128) == 0 && c <= 512 &&
                        ((((uintptr_t)x | (uintptr_t)w | (uintptr_t)b) & 15) == 0) && ((((uintptr_t)y) & 7) == 0);
    if (vec_ok) {
        switch (c / 128) {
            case 1: PIO_CUDA(launch_pdl(layernorm_fwd_bf16_vec_kernel<1>, dim3(blocks), dim3(threads), 0, st, x, w, b, y, stats, rows, eps, dyn_rows)); break;
            case 2: PIO_CUDA(launch_pdl(layernorm_fwd_bf16_vec_kernel<2>, dim3(blocks), dim3(threads), 0, st, x, w, b, y, stats, rows, eps, dyn_rows)); break;
            case 3: PIO_CUDA(launch_pdl(layernorm_fwd_bf16_vec_kernel<3>, dim3(blocks), dim3(threads), 0, st, x, w, b, y, stats, rows, eps, dyn_rows)); break;
            default: PIO_CUDA(launch_pdl(layernorm_fwd_bf16_vec_kernel<4>, dim3(blocks), dim3(threads), 0, st, x, w, b, y, stats, rows, eps, dyn_rows)); break;
        }
        count_launch();
        return PIO_OK;
    }
    PIO_CUDA(launch_pdl(layernorm_fwd_bf16_kernel, dim3(blocks), dim3(threads), 0, st, x, w, b, y, ldy, stats, rows, c, eps, dyn_rows));
    count_launch();
    return PIO_OK;
}

// grid (ceil(c/128), chunks): block covers 128 columns x rows [chunk*rpc, (chunk+1)*rpc)
template <typename TIn>
__global__ void __launch_bounds__(256) colsum_partial_kernel(const TIn* __restrict__ src, int64_t ld_src,
                                                             __nv_bfloat16* __restrict__ dst16, int64_t ld_dst,
                                                             float* __restrict__ partial, int64_t rows, int c,
                                                             int64_t rows_per_chunk, int pad_cols, float* __restrict__ atomic_out,
                                                             const int* __restrict__ dyn_rows) {
    __shared__ float s_part[256];
    pdl_trigger();
    pdl_wait();
    const int col = blockIdx.x * 128 + (threadIdx.x & 127);
    const int sub = threadIdx.x >> 7;
    if (dyn_rows != nullptr) {
        // device-side row count: the live rows are re-divided among the chunks (with the host's split the few live rows of
        // a mostly empty buffer all fell to the first chunks), and the bf16 copy is zero-filled from the count up to the
        // next multiple of 128 rows (operand padding) by the first chunk
        const int64_t live = min(rows, (int64_t)*dyn_rows);
        const int64_t pad_to = min(rows, (live + 127) / 128 * 128);
        rows = live;
        rows_per_chunk = max((int64_t)2, (live + gridDim.y - 1) / gridDim.y);
        if (dst16 && blockIdx.y == 0 && col < pad_cols)
            for (int64_t r = live + sub; r < pad_to; r += 2) dst16[r * ld_dst + col] = __float2bfloat16(0.f);
    }
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk;
    const int64_t r1 = min(rows, r0 + rows_per_chunk);
    float acc = 0.f;
    if (col < c) {
        for (int64_t r = r0 + sub; r < r1; r += 2) {
            float v;
            if constexpr (sizeof(TIn) == 4) v = (float)src[r * ld_src + col];
            else v = __bfloat162float(src[r * ld_src + col]);
            acc += v;
            if (dst16) dst16[r * ld_dst + col] = __float2bfloat16(v);
        }
    } else if (dst16 && col < pad_cols) {
        for (int64_t r = r0 + sub; r < r1; r += 2) dst16[r * ld_dst + col] = __float2bfloat16(0.f);
    }
    s_part[threadIdx.x] = acc;
    __syncthreads();
    if (sub == 0 && col < c) {
        const float total = acc + s_part[threadIdx.x + 128];
        if (atomic_out) atomicAdd(atomic_out + col, total);
        else if (partial) partial[(int64_t)blockIdx.y * c + col] = total;
    }
}

static int colsum_chunks(int64_t rows, int c) {
    const int64_t col_blocks = cdiv(c, 128);
    int64_t chunks = std::max<int64_t>(1, std::min<int64_t>(cdiv(2 * (int64_t)num_sms(), col_blocks), cdiv(rows, 16)));
    return (int)std::min<int64_t>(chunks, 256);
}

size_t colsum_ws_bytes(int64_t rows, int c) { return align_up((size_t)colsum_chunks(rows, c) * c * sizeof(float), 256) + 256; }

template <typename TIn>
static int colsum_impl(const TIn* src, int64_t ld_src, __nv_bfloat16* dst16, int64_t ld_dst, float* colsum, int accumulate,
                       int64_t rows, int c, void* ws, size_t ws_bytes, cudaStream_t st, const int* dyn_rows) {
    PIO_CHECK_ARG(src && c > 0 && rows >= 0, "colsum: bad arguments");
    if (rows == 0) return PIO_OK;
    const int chunks = colsum_chunks(rows, c);
    float* partial = nullptr;
    const bool atomic = (accumulate == 2);  // add straight into the (pre-zeroed / running) destination
    if (colsum && !atomic) {
        Workspace W(ws, ws_bytes);
        partial = W.takef((size_t)chunks * c);
        if (!partial) {
            set_error("colsum: workspace too small (%zu bytes)", ws_bytes);
            return PIO_ERR_WORKSPACE;
        }
    }
    const int pad_cols = dst16 ? (int)ld_dst : c;
    dim3 grid((unsigned)cdiv(std::max(pad_cols, c), 128), (unsigned)chunks);
    PIO_CUDA(launch_pdl(colsum_partial_kernel<TIn>, grid, dim3(256), 0, st, src, ld_src, dst16, ld_dst, partial, rows, c,
                        cdiv(rows, chunks), pad_cols, (colsum && atomic) ? colsum : (float*)nullptr, dyn_rows));
    count_launch();
    if (colsum && !atomic) PIO_TRY(reduce_partials(partial, chunks, c, colsum, c, nullptr, 0, accumulate, st));
    return PIO_OK;
}

int cast_colsum_f32(const float* src, int64_t ld_src, __nv_bfloat16* dst16, int64_t ld_dst, float* colsum, int accumulate,
                    int64_t rows, int c, void* ws, size_t ws_bytes, cudaStream_t st, const int* dyn_rows) {
    return colsum_impl<float>(src, ld_src, dst16, ld_dst, colsum, accumulate, rows, c, ws, ws_bytes, st, dyn_rows);
}

int colsum_bf16(const __nv_bfloat16* src, int64_t ld_src, float* colsum, int accumulate, int64_t rows, int c, void* ws,
                size_t ws_bytes, cudaStream_t st, const int* dyn_rows) {
    return colsum_impl<__nv_bfloat16>(src, ld_src, nullptr, 0, colsum, accumulate, rows, c, ws, ws_bytes, st, dyn_rows);
}

}  // namespace pio
