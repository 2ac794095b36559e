// Device helpers shared by the fused latent-block kernels (pio_latent_tc.cu: one CTA per 128-row tile;
// pio_latent_cl.cu: a cluster of four CTAs per tile): TMA stores, tensor-memory chunk loads / stores, bf16 rows into
// 128-byte-swizzled operand atoms, running LayerNorm statistics.
#pragma once
#include <cuda.h>

#include "pio_ops_common.cuh"
#include "pio_tc_common.cuh"

namespace pio {
namespace lb {

using bf16 = __nv_bfloat16;

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"((uint64_t)map),
                 "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// 32 fp32 of row `row` (0..31) into a 32-row x 128-byte swizzled staging slice
__device__ __forceinline__ void stage_row_f32(uint8_t* slice, int row, const float (&v)[32]) {
    uint8_t* prow = slice + row * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(prow + ((j ^ (row & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}

__device__ __forceinline__ void ld_chunk(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    tmem_ld32(taddr, r);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void st_chunk(uint32_t taddr, const float (&v)[32]) {
    uint32_t r[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
    tmem_st32(taddr, r);
}

// 32 fp32 -> bf16 into row `row`, 64-byte half `hh` of a 128-row x 128-byte swizzled (SW128, K-major) atom
__device__ __forceinline__ void store_half_row_smem(uint8_t* atom, int row, int hh, const float (&v)[32]) {
    uint8_t* prow = atom + row * 128;
#pragma unroll
    for (int piece = 0; piece < 4; ++piece) {
        uint32_t w[4];
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
            __nv_bfloat162 h2 = __floats2bfloat162_rn(v[piece * 8 + 2 * e2], v[piece * 8 + 2 * e2 + 1]);
            w[e2] = *reinterpret_cast<uint32_t*>(&h2);
        }
        const int ci = hh * 4 + piece;
        *reinterpret_cast<uint4*>(prow + ((ci ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}
__device__ __forceinline__ void store32_bf16_global(bf16* dst, const float (&v)[32]) {
#pragma unroll
    for (int d = 0; d < 32; d += 8) {
        uint32_t w[4];
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
            __nv_bfloat162 h2 = __floats2bfloat162_rn(v[d + 2 * e2], v[d + 2 * e2 + 1]);
            w[e2] = *reinterpret_cast<uint32_t*>(&h2);
        }
        *reinterpret_cast<uint4*>(dst + d) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}
__device__ __forceinline__ void store32_f32_global(float* dst, const float (&v)[32]) {
#pragma unroll
    for (int d = 0; d < 32; d += 4) *reinterpret_cast<float4*>(dst + d) = make_float4(v[d], v[d + 1], v[d + 2], v[d + 3]);
}
// v[i] += p[i], p in shared memory and warp-uniform (a broadcast read)
__device__ __forceinline__ void add_vec32(float (&v)[32], const float* p) {
#pragma unroll
    for (int d = 0; d < 32; d += 4) {
        const float4 t = *reinterpret_cast<const float4*>(p + d);
        v[d] += t.x; v[d + 1] += t.y; v[d + 2] += t.z; v[d + 3] += t.w;
    }
}

// running (count, mean, sum of squared deviations); chunks merge with Chan's update, so the statistics
// are as robust as a two-pass LayerNorm although every value is seen once
struct Stat {
    float n, mean, m2;
};
__device__ __forceinline__ void stat_merge(Stat& a, float nb, float mean_b, float m2_b) {
    const float nt = a.n + nb;
    const float d = mean_b - a.mean;
    const float f = nb / nt;
    a.mean = fmaf(d, f, a.mean);
    a.m2 = a.m2 + m2_b + d * d * a.n * f;
    a.n = nt;
}
__device__ __forceinline__ void stat_chunk(Stat& a, const float (&v)[32]) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) s += v[i];
    const float mean = s * (1.0f / 32.0f);
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const float d = v[i] - mean;
        m2 = fmaf(d, d, m2);
    }
    if (a.n == 0.f) {
        a.n = 32.f; a.mean = mean; a.m2 = m2;
    } else {
        stat_merge(a, 32.f, mean, m2);
    }
}

// 32 bf16 of row `row`, 64-byte half `hh` of a 128-row x 128-byte swizzled atom -> fp32
__device__ __forceinline__ void load_half_row_smem(const uint8_t* atom, int row, int hh, float (&v)[32]) {
    const uint8_t* prow = atom + row * 128;
#pragma unroll
    for (int piece = 0; piece < 4; ++piece) {
        const int ci = hh * 4 + piece;
        const uint4 w = *reinterpret_cast<const uint4*>(prow + ((ci ^ (row & 7)) << 4));
        const uint32_t u[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
            const __nv_bfloat162 h2 = *reinterpret_cast<const __nv_bfloat162*>(&u[e2]);
            const float2 f = __bfloat1622float2(h2);
            v[piece * 8 + 2 * e2] = f.x;
            v[piece * 8 + 2 * e2 + 1] = f.y;
        }
    }
}
// 32 fp32 of row `row` of a 128-row x 128-byte swizzled box (TMA box of 32 fp32 columns)
__device__ __forceinline__ void load_row_f32_smem(const uint8_t* box, int row, float (&v)[32]) {
    const uint8_t* prow = box + row * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 t = *reinterpret_cast<const float4*>(prow + ((j ^ (row & 7)) << 4));
        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
    }
}

namespace cl {
// ------------------------------------------------------------------ thread-block cluster PTX
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// the shared::cluster address of `local` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t local, uint32_t rank) {
    uint32_t a;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(local), "r"(rank));
    return a;
}
// 8 bytes into a peer's shared memory, completing on the peer's mbarrier like a TMA write: the receiver learns of the
// data through the barrier's transaction count, so neither side needs a fence (a release / acquire at cluster scope
// compiles to MEMBAR.ALL.GPU / CCTL.IVALL, ~1 K cycles each: the first version spent a third of the layer in them)
__device__ __forceinline__ void st_async_f32x2(uint32_t addr, float a, float b, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(addr), "f"(a), "f"(b),
                 "r"(bar)
                 : "memory");
}
// arrive on a peer's mbarrier; relaxed: it orders no data, only tells the peer that products of this CTA have completed
__device__ __forceinline__ void mbar_arrive_peer(uint32_t addr) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(addr) : "memory");
}
// 16-byte-granular copy from this CTA's shared memory into a peer's; the bytes complete on the peer's mbarrier
__device__ __forceinline__ void bulk_copy_to_peer(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
                 "r"(src_cta), "r"(bytes), "r"(bar_cluster)
                 : "memory");
}
}  // namespace cl

}  // namespace lb

// (cols, rows, layers) tensor map with byte strides (pitch, layer_stride); box = box_cols x box_rows x 1, 128-byte swizzle
inline int make_out_map(CUtensorMap* map, bool f32, const void* base, int64_t cols, int64_t rows, int64_t nl, int64_t pitch,
                        int64_t layer_stride, int box_cols, int box_rows, bool swizzle = true) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled entry point not available from the driver");
        return PIO_ERR_CUDA;
    }
    PIO_CHECK_ARG((((uintptr_t)base) & 15) == 0 && pitch % 16 == 0 && layer_stride % 16 == 0 && layer_stride > 0,
                  "latent_block_fwd_fused: saved-state tensor not 16-byte aligned (base %p pitch %lld layer stride %lld)", base,
                  (long long)pitch, (long long)layer_stride);
    cuuint64_t gdim[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)nl};
    cuuint64_t gstr[2] = {(cuuint64_t)pitch, (cuuint64_t)layer_stride};
    cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1u};
    cuuint32_t estr[3] = {1u, 1u, 1u};
    CUresult r = fn(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim,
                    gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (latent block output) failed with CUresult %d", (int)r);
        return PIO_ERR_CUDA;
    }
    return PIO_OK;
}

}  // namespace pio
