// Persistent vocabulary-logits product for the fused linear + cross-entropy head:
//   C[m][n] (bf16) = A[m][k] W[n][k]^T + bias[n]   for the first *count - off rows of a chunk, k <= 256.
//
// Reference: the nn.Linear of PerceiverDecoder / TextOutputAdapter (perceiver/adapter.py:101-109) under
// LitMaskedLanguageModel's cross-entropy (lightning.py:223-226).  pio_linear_ce_fwd_bwd launches every kernel for the
// worst case (all rows labelled) and lets CTAs past the device-side row count exit; with one CTA per 128 x 128 tile
// (gemm_tc_kernel) a chunk without a single live row still cost 15 us (2686 CTAs started and retired), and a live
// chunk ran its four-k-block mainloops at 18 % tensor-pipe activity because every CTA paid set-up, a cold pipeline
// and a serial epilogue.  Here the grid is one CTA per SM and each CTA walks a contiguous range of the LIVE tiles
// (an empty chunk retires 148 CTAs at once):
//   * the A tile (128 rows x k) stays in shared memory while the CTA moves along n; W tiles (128 x k, 64 KB) stream
//     through a 2-stage TMA ring;
//   * two 128-column accumulators in tensor memory: the eight epilogue warps turn tile i into bf16 (bias added,
//     staged per warp so that global stores are whole 128-byte lines) while the MMA thread already issues tile i + 1;
//   * rows between the live count and the end of the last live 128-row tile are written as zeros (the weight-gradient
//     product contracts over the row axis in 64-row blocks).
#include <cuda.h>

#include <atomic>

#include "pio_ops_common.cuh"
#include "pio_tc_common.cuh"

namespace pio {

int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

namespace lg {

constexpr int BM = 128, BN = 128;
constexpr int MAX_KB = 4;             // k <= 256
constexpr int KB_BYTES = 128 * 128;   // one 64-wide k-block of a 128-row tile
constexpr int A_BYTES = MAX_KB * KB_BYTES;
constexpr int B_BYTES = MAX_KB * KB_BYTES;
constexpr int STAGES = 2;
constexpr int THREADS = 320;  // warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2..9 epilogue
constexpr int STAGING = 8 * 32 * 128;
constexpr size_t SMEM_BYTES = A_BYTES + STAGES * B_BYTES + STAGING + 16 * sizeof(uint64_t) + 1024;
static_assert(SMEM_BYTES <= 232448, "logits kernel: more than 227 KB of shared memory");

struct Params {
    __nv_bfloat16* c;
    int64_t ldc;
    const float* bias;
    int M, N, num_kb;
    const int* dyn_rows;
    int dyn_off;
};

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(THREADS, 1)
    logits_persistent_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB = sA + A_BYTES;
    uint8_t* sStage = sB + STAGES * B_BYTES;
    uint64_t* bars = (uint64_t*)(sStage + STAGING);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + 1;
    uint64_t* b_full = bars + 2;     // [STAGES]
    uint64_t* b_empty = bars + 4;    // [STAGES]
    uint64_t* acc_full = bars + 6;   // [2]
    uint64_t* acc_empty = bars + 8;  // [2]
    uint32_t* tmem_slot = (uint32_t*)(bars + 10);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_wait();
    int m_eff = p.M;
    if (p.dyn_rows != nullptr) m_eff = min(m_eff, max(0, __ldg(p.dyn_rows) - p.dyn_off));
    const int n_tiles = (p.N + BN - 1) / BN;
    const int live = ((m_eff + BM - 1) / BM) * n_tiles;
    const int per = (live + (int)gridDim.x - 1) / (int)gridDim.x;
    const int t_begin = min(live, (int)blockIdx.x * per), t_end = min(live, t_begin + per);
    if (t_begin >= t_end) return;  // (block-uniform) nothing to do: leave before any set-up

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&b_full[s], 1);
            mbar_init(&b_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&acc_full[s], 1);
            mbar_init(&acc_empty[s], 8);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 2 * BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int nkb = p.num_kb;

    if (warp == 0) {
        if (lane == 0) {
            int cur_m = -1, a_loads = 0;
            for (int t = t_begin, i = 0; t < t_end; ++t, ++i) {
                const int mt = t / n_tiles, nt = t - mt * n_tiles;
                const int s = i % STAGES;
                mbar_wait(&b_empty[s], ((uint32_t)(i / STAGES) & 1u) ^ 1u);
                mbar_expect_tx(&b_full[s], (uint32_t)nkb * KB_BYTES);
                for (int kb = 0; kb < nkb; ++kb) tma_load_2d(sB + s * B_BYTES + kb * KB_BYTES, &tmB, kb * 64, nt * BN, &b_full[s]);
                if (mt != cur_m) {  // a new row tile: the products of the previous one must have read sA
                    if (a_loads > 0) mbar_wait(a_empty, (uint32_t)(a_loads - 1) & 1u);
                    mbar_expect_tx(a_full, (uint32_t)nkb * KB_BYTES);
                    for (int kb = 0; kb < nkb; ++kb) tma_load_2d(sA + kb * KB_BYTES, &tmA, kb * 64, mt * BM, a_full);
                    cur_m = mt;
                    ++a_loads;
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(BN, false, false);
            const uint32_t a_addr = smem_u32(sA);
            int cur_m = -1, a_loads = 0;
            for (int t = t_begin, i = 0; t < t_end; ++t, ++i) {
                const int mt = t / n_tiles;
                const bool new_m = mt != cur_m;
                if (new_m) {
                    mbar_wait(a_full, (uint32_t)a_loads & 1u);
                    cur_m = mt;
                    ++a_loads;
                }
                const int s = i % STAGES, acc = i & 1;
                mbar_wait(&b_full[s], (uint32_t)(i / STAGES) & 1u);
                mbar_wait(&acc_empty[acc], ((uint32_t)(i >> 1) & 1u) ^ 1u);  // the epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t b_addr = smem_u32(sB + s * B_BYTES);
                for (int kb = 0; kb < nkb; ++kb)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem_base + (uint32_t)(acc * BN), make_smem_desc(a_addr + kb * KB_BYTES + k * 32, 16, 1024),
                                  make_smem_desc(b_addr + kb * KB_BYTES + k * 32, 16, 1024), idesc, (kb > 0 || k > 0) ? 1u : 0u);
                umma_commit(&b_empty[s]);
                umma_commit(&acc_full[acc]);
                // the last tile of this row tile: once its products are done sA may be overwritten
                const bool last_of_m = (t + 1 == t_end) || ((t + 1) / n_tiles != mt);
                if (last_of_m) umma_commit(a_empty);
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3;            // TMEM lane quadrant: rows 32 q .. 32 q + 31 of the tile
        const int half = (warp - 2) >> 2;  // columns 64 half .. + 63 of the tile
        uint8_t* stg = sStage + (warp - 2) * (32 * 128);
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        for (int t = t_begin, i = 0; t < t_end; ++t, ++i) {
            const int mt = t / n_tiles, nt = t - mt * n_tiles;
            const int acc = i & 1;
            const int n0 = nt * BN + half * 64;
            float bias[2];  // lane l adds the bias of columns n0 + l and n0 + 32 + l (shuffled to the column's owner below)
            bias[0] = (n0 + lane < p.N && p.bias) ? __ldg(p.bias + n0 + lane) : 0.f;
            bias[1] = (n0 + 32 + lane < p.N && p.bias) ? __ldg(p.bias + n0 + 32 + lane) : 0.f;
            mbar_wait(&acc_full[acc], (uint32_t)(i >> 1) & 1u);
            tc_fence_after();
            uint32_t r0[32], r1[32];
            tmem_ld32(tmem_base + lane_off + (uint32_t)(acc * BN + half * 64), r0);
            tmem_ld32(tmem_base + lane_off + (uint32_t)(acc * BN + half * 64 + 32), r1);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);  // the accumulator is in registers: the next tile but one may overwrite it
            // thread = row: 64 bf16 (128 bytes) into the warp's staging tile, 16-byte pieces rotated by the row to spread the banks
            uint8_t* prow = stg + lane * 128;
#pragma unroll
            for (int pc = 0; pc < 8; ++pc) {
                uint32_t w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int col = pc * 8 + 2 * e;
                    const float b0 = __shfl_sync(0xffffffffu, bias[col >> 5], col & 31);
                    const float b1 = __shfl_sync(0xffffffffu, bias[(col + 1) >> 5], (col + 1) & 31);
                    const float v0 = __uint_as_float(col < 32 ? r0[col] : r1[col - 32]) + b0;
                    const float v1 = __uint_as_float(col + 1 < 32 ? r0[col + 1] : r1[col + 1 - 32]) + b1;
                    __nv_bfloat162 h2 = __floats2bfloat162_rn(v0, v1);
                    w[e] = *reinterpret_cast<uint32_t*>(&h2);
                }
                *reinterpret_cast<uint4*>(prow + (((pc + lane) & 7) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
            }
            __syncwarp();
            // lanes 8 j .. 8 j + 7 write the eight 16-byte pieces of one row: four whole rows (lines) per instruction
            const int64_t row_base = (int64_t)mt * BM + q * 32;
#pragma unroll
            for (int rr0 = 0; rr0 < 32; rr0 += 4) {
                const int rr = rr0 + (lane >> 3), pc = lane & 7;
                const int64_t gm = row_base + rr;
                const int col0 = n0 + pc * 8;
                if (gm < p.M && col0 < p.N) {
                    uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 128 + (((pc + rr) & 7) << 4));
                    if (gm >= m_eff) v = make_uint4(0u, 0u, 0u, 0u);  // operand padding for the product over the row axis
                    *reinterpret_cast<uint4*>(p.c + gm * p.ldc + col0) = v;
                }
            }
            __syncwarp();
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * BN);
    }
}

}  // namespace lg

bool logits_persistent_supported(int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldb, int64_t ldc) {
    return getenv("PIO_NO_PERSISTENT_LOGITS") == nullptr && m > 0 && m < (1ll << 30) && n >= 128 && n < (1ll << 30) && k >= 64 &&
           k <= 64 * lg::MAX_KB && k % 64 == 0 && lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0 && ldc >= (n + 7) / 8 * 8;
}

// c (m x n, bf16, pitch ldc >= n rounded up to 8) = a (m x k, bf16) w (n x k, bf16)^T + bias, rows past *dyn_rows - dyn_off
// skipped (zeros up to the end of the last live 128-row tile)
int logits_persistent(const void* a, int64_t lda, const void* w, int64_t ldb, const float* bias, void* c, int64_t ldc, int64_t m,
                      int64_t n, int64_t k, const int* dyn_rows, int dyn_off, cudaStream_t st) {
    using namespace lg;
    PIO_CHECK_ARG(a && w && c && logits_persistent_supported(m, n, k, lda, ldb, ldc) && (((uintptr_t)c) & 15) == 0,
                  "logits_persistent: unsupported shape (m %lld n %lld k %lld)", (long long)m, (long long)n, (long long)k);
    CUtensorMap tmA, tmB;
    PIO_TRY(make_tmap_bf16(&tmA, a, m, k, lda, BM));
    PIO_TRY(make_tmap_bf16(&tmB, w, n, k, ldb, BN));
    Params p{};
    p.c = (__nv_bfloat16*)c;
    p.ldc = ldc;
    p.bias = bias;
    p.M = (int)m;
    p.N = (int)n;
    p.num_kb = (int)(k / 64);
    p.dyn_rows = dyn_rows;
    p.dyn_off = dyn_off;
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(logits_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    const int64_t tiles = cdiv(m, (int64_t)BM) * cdiv(n, (int64_t)BN);
    const unsigned grid = (unsigned)std::min<int64_t>(tiles, num_sms());
    PIO_CUDA(launch_pdl(logits_persistent_kernel, dim3(grid), dim3(THREADS), SMEM_BYTES, st, tmA, tmB, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
