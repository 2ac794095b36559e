// Batched weight-gradient products: dW_i[m_i x 256] += dY_i^T X_i (+ db_i[m_i] += column sums of dY_i) for a table of
// problems that all contract over the same `rows`, in ONE launch.
//
// Reference: the weight / bias gradients of the Linear layers inside self_attention_layer (perceiver/model.py:20-26,
// :102-116), which the reference leaves to autograd.  The fused backward chain of a latent block
// (pio_latent_bwd_cl.cu) leaves dY and X of 4 products per layer in memory; launched one by one those 24 products of
// a 6-layer block took 24 x ~11 us on a side stream (each is 0.5-1.6 GFLOP: set-up, a 64-block K loop on few CTAs
// and an atomic epilogue), 2.5 x the chain kernel itself.  Here every (problem, 128-row tile of dW, K split) is one
// CTA of one grid sized to a single wave:
//   * both operands are read in place, MN-major (dY[rows][m] and X[rows][256] are row-major with the contraction
//     index outermost), by TMA into a 4-stage ring of 16 KB + 32 KB;
//   * tcgen05.mma M = 128, N = 256, fp32 accumulator in 256 TMEM columns; a second, N = 16 product of the same A tile
//     with a tile of ones accumulates the tile's bias gradient (column sums of dY) in 16 more columns for free;
//   * epilogue: accumulator rows staged through the idle ring so that a warp adds 512 contiguous bytes per
//     instruction (red.global.add.v4.f32) into the gradient, which the caller has zero-filled or is accumulating into.
#include <cuda.h>

#include <atomic>

#include "pio_ops_common.cuh"
#include "pio_tc_common.cuh"

namespace pio {

int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

namespace dwb {

constexpr int BM = 128, BN = 256, BK = 64;
constexpr int STAGES = 4;
constexpr int A_STAGE = BM * BK * 2;  // 16 KB: two boxes of 64 k-rows x 64 m
constexpr int B_STAGE = BN * BK * 2;  // 32 KB: four boxes of 64 k-rows x 64 n
constexpr int ONES_BYTES = 8192;
constexpr int THREADS = 320;  // warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2..9 epilogue
constexpr int MAX_PROB = 32, MAX_TILES = 160;
constexpr int PITCH = 128 + 4;  // floats per staged row (a warp finishes 128 columns)
constexpr size_t SMEM_BYTES = STAGES * (A_STAGE + B_STAGE) + ONES_BYTES + 16 * sizeof(uint64_t) + 1024;
static_assert(8 * 32 * PITCH * sizeof(float) <= STAGES * (A_STAGE + B_STAGE), "epilogue staging must fit in the ring");

struct Problem {
    CUtensorMap tmA, tmB;
    float* dw;
    float* db;
    int m;
    int pad;
};
struct Batch {
    Problem prob[MAX_PROB];
    unsigned char tile_prob[MAX_TILES];
    unsigned char tile_m[MAX_TILES];
    int num_kb, kb_per_split;
    const int* dyn_rows;  // device-side count of live rows (the contraction then ends at its next multiple of 64), or null
    int dyn_off;
    int store;  // A_MN = false with a single K split: the tile is stored, not added
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// A_MN = true: the weight-gradient form above (A = dY^T, the contraction over rows, the device-side count bounds K).
// A_MN = false: C (m x 256, fp32, accumulated into) += A (m x k, bf16, K-major) B (k x 256, bf16, row-major) split over K —
// the loss head's dx = dlogits W (m = labelled rows of a chunk, k = vocabulary): 128 x 256 tiles so that A is read once,
// K split to fill the GPU; the device-side count then bounds M (tiles past it exit).
template <bool A_MN>
__global__ void __launch_bounds__(THREADS, 1) dw_batched_kernel(const __grid_constant__ Batch g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB = smem + STAGES * A_STAGE;
    uint8_t* sOnes = sB + STAGES * B_STAGE;
    uint64_t* full_bar = (uint64_t*)(sOnes + ONES_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full = empty_bar + STAGES;
    uint32_t* tmem_slot = (uint32_t*)(tmem_full + 1);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Problem& P = g.prob[A_MN ? g.tile_prob[blockIdx.x] : 0];
    const int m0 = (A_MN ? (int)g.tile_m[blockIdx.x] : (int)blockIdx.x) * BM;  // (the single product of the A_MN = false form needs no tile table)
    int num_kb = g.num_kb, kb_per_split = g.kb_per_split, m_live = P.m;
    if (g.dyn_rows != nullptr) {
        pdl_wait();
        const int live = max(0, __ldg(g.dyn_rows) - g.dyn_off);
        if (A_MN) {  // (block-uniform) the live rows are re-divided among the K splits
            num_kb = min(num_kb, (live + BK - 1) / BK);
            kb_per_split = (num_kb + (int)gridDim.y - 1) / (int)gridDim.y;
        } else {
            m_live = min(m_live, live);
            if (m0 >= m_live) return;
        }
    }
    const int kb_begin = blockIdx.y * kb_per_split;
    const int kb_end = min(num_kb, kb_begin + kb_per_split);
    const int nkb = kb_end - kb_begin;
    if (nkb <= 0) return;

    if (threadIdx.x == 0) {
        prefetch_tmap(&P.tmA);
        prefetch_tmap(&P.tmB);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
    }
    for (int i = threadIdx.x; i < ONES_BYTES / 4; i += THREADS) reinterpret_cast<uint32_t*>(sOnes)[i] = 0x3F803F80u;  // bf16 1.0 pairs
    fence_proxy_async();
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                mbar_wait(&empty_bar[s], ((uint32_t)(i / STAGES) & 1u) ^ 1u);
                mbar_expect_tx(&full_bar[s], A_STAGE + B_STAGE);
                const int k0 = (kb_begin + i) * BK;
                if (A_MN) {
#pragma unroll
                    for (int j = 0; j < BM / 64; ++j) tma_load_2d(sA + s * A_STAGE + j * 8192, &P.tmA, m0 + j * 64, k0, &full_bar[s]);
                } else {
                    tma_load_2d(sA + s * A_STAGE, &P.tmA, k0, m0, &full_bar[s]);
                }
#pragma unroll
                for (int j = 0; j < BN / 64; ++j) tma_load_2d(sB + s * B_STAGE + j * 8192, &P.tmB, j * 64, k0, &full_bar[s]);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(BN, A_MN, true);
            const uint32_t idesc_b = make_idesc(16, A_MN, true);
            const uint32_t ones = smem_u32(sOnes);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                mbar_wait(&full_bar[s], (uint32_t)(i / STAGES) & 1u);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(sA + s * A_STAGE), b_addr = smem_u32(sB + s * B_STAGE);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                    const uint64_t da = A_MN ? make_smem_desc(a_addr + k * 2048, 8192, 1024) : make_smem_desc(a_addr + k * 32, 16, 1024);
                    umma_bf16(tmem_base, da, make_smem_desc(b_addr + k * 2048, 8192, 1024), idesc, (i > 0 || k > 0) ? 1u : 0u);
                    if (A_MN && P.db != nullptr)
                        umma_bf16(tmem_base + BN, da, make_smem_desc(ones + k * 2048, 8192, 1024), idesc_b, (i > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(tmem_full);
        }
        __syncwarp();
    } else {
        const int q = warp & 3;            // TMEM lane quadrant: rows m0 + 32 q .. + 32 of dW
        const int half = (warp - 2) >> 2;  // which 128 of the 256 columns
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        float* stile = reinterpret_cast<float*>(smem) + (half * 4 + q) * (32 * PITCH);
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
#pragma unroll
        for (int c0 = 0; c0 < 128; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + lane_off + (uint32_t)(half * 128 + c0), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(stile + lane * PITCH + c0 + j) =
                    make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        }
        if (A_MN && half == 0 && P.db != nullptr) {  // the tile's bias gradient: every one of the 16 extra columns holds it
            uint32_t r[32];
            tmem_ld32(tmem_base + lane_off + (uint32_t)BN, r);
            tmem_ld_wait();
            const int m = m0 + q * 32 + lane;
            if (m < P.m) atomicAdd(P.db + m, __uint_as_float(r[0]));
        }
        __syncwarp();
        const int rows_here = min(32, m_live - (m0 + q * 32));
        for (int rr = 0; rr < rows_here; ++rr) {
            const float4 v = *reinterpret_cast<const float4*>(stile + rr * PITCH + lane * 4);
            float* o = P.dw + (int64_t)(m0 + q * 32 + rr) * BN + half * 128 + lane * 4;
            if (!A_MN && g.store) *reinterpret_cast<float4*>(o) = v;
            else red_add_v4(o, v.x, v.y, v.z, v.w);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace dwb

// items[i]: dw (m x 256, row-major, accumulated into), db (m, accumulated into; may be null), dy16 [rows][m] (pitch lddy),
// x16 [rows][256] (pitch ldx).  dyn_rows (device memory, nullable): only the first *dyn_rows - dyn_off rows are live; both
// operands must be zero from there to the next multiple of 64 rows.
int dw_batched(const DwBatchItem* items, int count, int64_t rows, cudaStream_t st, const int* dyn_rows, int dyn_off) {
    using namespace dwb;
    PIO_CHECK_ARG(items && count >= 1 && count <= MAX_PROB && rows > 0, "dw_batched: bad arguments (count %d)", count);
    static_assert(sizeof(Batch) <= 32000, "dw_batched: the problem table must fit the kernel parameter space");
    Batch g{};
    int tiles = 0;
    for (int i = 0; i < count; ++i) {
        const DwBatchItem& it = items[i];
        PIO_CHECK_ARG(it.dy16 && it.x16 && it.dw && it.m > 0 && it.n == BN && (((uintptr_t)it.dw) & 15) == 0,
                      "dw_batched: item %d not supported (m %d n %d)", i, it.m, it.n);
        PIO_TRY(make_tmap_bf16(&g.prob[i].tmA, it.dy16, rows, it.m, it.lddy, 64));
        PIO_TRY(make_tmap_bf16(&g.prob[i].tmB, it.x16, rows, BN, it.ldx, 64));
        g.prob[i].dw = it.dw;
        g.prob[i].db = it.db;
        g.prob[i].m = it.m;
        for (int t = 0; t < (int)cdiv(it.m, BM); ++t) {
            PIO_CHECK_ARG(tiles < MAX_TILES && t < 256, "dw_batched: more than %d output tiles", MAX_TILES);
            g.tile_prob[tiles] = (unsigned char)i;
            g.tile_m[tiles] = (unsigned char)t;
            ++tiles;
        }
    }
    g.dyn_rows = dyn_rows;
    g.dyn_off = dyn_off;
    g.num_kb = (int)cdiv(rows, (int64_t)BK);
    const int splits = std::max(1, std::min(g.num_kb, num_sms() / tiles));
    g.kb_per_split = (int)cdiv(g.num_kb, splits);
    const int used = (int)cdiv(g.num_kb, g.kb_per_split);
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(dw_batched_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        PIO_CUDA(cudaFuncSetAttribute(dw_batched_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    PIO_CUDA(launch_pdl(dw_batched_kernel<true>, dim3((unsigned)tiles, (unsigned)used), dim3(THREADS), SMEM_BYTES, st, g));
    count_launch();
    return PIO_OK;
}

// c32 (m x 256, fp32, row-major) (+)= a16 (m x k, bf16, pitch lda) b16 (k x 256, bf16, pitch ldb) in 128 x 256 tiles (a is read
// once), split over K when there are fewer tiles than SMs; dyn_rows (device memory, nullable): only the first
// *dyn_rows - dyn_off rows of a / c are live.  *stored = 1 when the tiles were stored (one K split: c32 need not be zeroed),
// 0 when they were ADDED with atomics (the caller zero-fills c32 first; xw_splitk_stores tells beforehand).
bool xw_splitk_supported(int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldb) {
    return n == dwb::BN && m > 0 && m < (1ll << 30) && k > 0 && lda % 8 == 0 && ldb % 8 == 0;
}
bool xw_splitk_stores(int64_t m, int64_t k) {
    const int64_t tiles = cdiv(m, (int64_t)dwb::BM);
    return std::max<int64_t>(1, std::min<int64_t>(cdiv(k, (int64_t)dwb::BK), num_sms() / tiles)) == 1;
}
int xw_splitk(const void* a16, int64_t lda, const void* b16, int64_t ldb, float* c32, int64_t m, int64_t k, cudaStream_t st,
              const int* dyn_rows, int dyn_off) {
    using namespace dwb;
    PIO_CHECK_ARG(a16 && b16 && c32 && xw_splitk_supported(m, BN, k, lda, ldb) && (((uintptr_t)c32) & 15) == 0, "xw_splitk: bad arguments");
    Batch g{};
    PIO_TRY(make_tmap_bf16(&g.prob[0].tmA, a16, m, k, lda, BM));
    PIO_TRY(make_tmap_bf16(&g.prob[0].tmB, b16, k, BN, ldb, 64));
    g.prob[0].dw = c32;
    g.prob[0].db = nullptr;
    g.prob[0].m = (int)m;
    const int tiles = (int)cdiv(m, (int64_t)BM);
    g.dyn_rows = dyn_rows;
    g.dyn_off = dyn_off;
    g.num_kb = (int)cdiv(k, (int64_t)BK);
    const int splits = std::max(1, std::min(g.num_kb, num_sms() / tiles));
    g.store = splits == 1 ? 1 : 0;
    g.kb_per_split = (int)cdiv(g.num_kb, splits);
    const int used = (int)cdiv(g.num_kb, g.kb_per_split);
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(dw_batched_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        PIO_CUDA(cudaFuncSetAttribute(dw_batched_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    PIO_CUDA(launch_pdl(dw_batched_kernel<false>, dim3((unsigned)tiles, (unsigned)used), dim3(THREADS), SMEM_BYTES, st, g));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
