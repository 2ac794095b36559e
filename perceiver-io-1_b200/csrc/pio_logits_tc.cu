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
constexpr int EPI_WARPS = 16;  // four per scheduler: the epilogue (and the LayerNorm prologue) is a chain of latencies, ncu: 8.4 cycles per issue with two
constexpr int THREADS = 64 + EPI_WARPS * 32;  // warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2..17 epilogue
constexpr int STAGING = EPI_WARPS * 32 * 64;
constexpr size_t SMEM_BYTES = A_BYTES + STAGES * B_BYTES + STAGING + 16 * sizeof(uint64_t) + 1024;
static_assert(SMEM_BYTES <= 232448, "logits kernel: more than 227 KB of shared memory");

struct Params {
    __nv_bfloat16* c;
    int64_t ldc;
    const float* bias;
    int M, N, num_kb;
    const int* dyn_rows;
    int dyn_off;
    // LN = true: the A tile is LayerNorm(x) formed in the kernel
    const float* x;       // (M, k) fp32, row pitch k
    const float* gamma;
    const float* beta;
    __nv_bfloat16* h0;    // LayerNorm output as bf16 (M, ldh): saved for the backward pass
    int64_t ldh;
    float* stats;         // (M, 2): mean, rstd
    float eps;
};

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// LN = true (k = 256 only): C = LayerNorm(x) W^T + bias with the LayerNorm inside — the kv_norm + K/V in-projection of a
// cross-attention layer (reference model.py:59-77, :80-99) as ONE kernel: the eight epilogue warps first normalise the
// 128 rows of a row tile (warp = 16 rows, lane = 8 consecutive channels: coalesced fp32 reads, statistics by warp
// shuffles), write them as bf16 straight into the swizzled A tile (and to h0 / stats for the backward pass), then serve
// as the epilogue of that row tile's column tiles.  x is read once from HBM and the normalised operand never makes the
// round trip through global memory that a separate LayerNorm kernel + GEMM pays.
template <bool LN>
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
    int per = (live + (int)gridDim.x - 1) / (int)gridDim.x;
    if (LN) per = (per + n_tiles - 1) / n_tiles * n_tiles;  // whole row tiles: a row tile is normalised by one CTA only
    const int t_begin = min(live, (int)blockIdx.x * per), t_end = min(live, t_begin + per);
    if (t_begin >= t_end) return;  // (block-uniform) nothing to do: leave before any set-up

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        mbar_init(a_full, LN ? EPI_WARPS : 1);
        mbar_init(a_empty, 1);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&b_full[s], 1);
            mbar_init(&b_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&acc_full[s], 1);
            mbar_init(&acc_empty[s], EPI_WARPS);
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
                if (!LN && mt != cur_m) {  // a new row tile: the products of the previous one must have read sA
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
        const int q = warp & 3;           // TMEM lane quadrant: rows 32 q .. 32 q + 31 of the tile
        const int cq = (warp - 2) >> 2;   // columns 32 cq .. + 31 of the tile
        uint8_t* stg = sStage + (warp - 2) * (32 * 64);
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        int cur_m = -1, a_loads = 0;
        float gam[8], bet[8];
        if (LN) {
#pragma unroll
            for (int j = 0; j < 8; j += 4) {
                const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.gamma + lane * 8 + j));
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.beta + lane * 8 + j));
                gam[j] = g4.x; gam[j + 1] = g4.y; gam[j + 2] = g4.z; gam[j + 3] = g4.w;
                bet[j] = b4.x; bet[j + 1] = b4.y; bet[j + 2] = b4.z; bet[j + 3] = b4.w;
            }
        }
        for (int t = t_begin, i = 0; t < t_end; ++t, ++i) {
            const int mt = t / n_tiles, nt = t - mt * n_tiles;
            const int acc = i & 1;
            if (LN && mt != cur_m) {
                // ---- LayerNorm of the row tile into sA: this warp's 8 rows, four at a time in flight
                if (a_loads > 0) mbar_wait(a_empty, (uint32_t)(a_loads - 1) & 1u);  // the previous row tile's products have read sA
                const int w16 = warp - 2;
                const int kb = lane >> 3, piece = lane & 7;
#pragma unroll 1
                for (int r4 = 0; r4 < 8; r4 += 4) {
                    float v[4][8];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int64_t gm = (int64_t)mt * BM + w16 * 8 + r4 + u;
                        if (gm < p.M) {
                            const float4 a = __ldg(reinterpret_cast<const float4*>(p.x + gm * 256 + lane * 8));
                            const float4 b = __ldg(reinterpret_cast<const float4*>(p.x + gm * 256 + lane * 8 + 4));
                            v[u][0] = a.x; v[u][1] = a.y; v[u][2] = a.z; v[u][3] = a.w;
                            v[u][4] = b.x; v[u][5] = b.y; v[u][6] = b.z; v[u][7] = b.w;
                        } else {
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[u][j] = 0.f;
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int row = w16 * 8 + r4 + u;
                        const int64_t gm = (int64_t)mt * BM + row;
                        float sum = 0.f;
#pragma unroll
                        for (int j = 0; j < 8; ++j) sum += v[u][j];
                        const float mean = warp_sum(sum) * (1.0f / 256.0f);
                        float sq = 0.f;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float d = v[u][j] - mean;
                            sq = fmaf(d, d, sq);
                        }
                        const float rstd = rsqrtf(warp_sum(sq) * (1.0f / 256.0f) + p.eps);
                        uint32_t w[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float y0 = fmaf((v[u][2 * e] - mean) * rstd, gam[2 * e], bet[2 * e]);
                            const float y1 = fmaf((v[u][2 * e + 1] - mean) * rstd, gam[2 * e + 1], bet[2 * e + 1]);
                            __nv_bfloat162 h2 = __floats2bfloat162_rn(gm < p.M ? y0 : 0.f, gm < p.M ? y1 : 0.f);
                            w[e] = *reinterpret_cast<uint32_t*>(&h2);
                        }
                        const uint4 pk = make_uint4(w[0], w[1], w[2], w[3]);
                        *reinterpret_cast<uint4*>(sA + kb * KB_BYTES + row * 128 + ((piece ^ (row & 7)) << 4)) = pk;
                        if (gm < p.M) {
                            *reinterpret_cast<uint4*>(p.h0 + gm * p.ldh + lane * 8) = pk;
                            if (lane == 0) *reinterpret_cast<float2*>(p.stats + 2 * gm) = make_float2(mean, rstd);
                        }
                    }
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_full);
                cur_m = mt;
                ++a_loads;
            }
            const int n0 = nt * BN + cq * 32;
            const float bias = (n0 + lane < p.N && p.bias) ? __ldg(p.bias + n0 + lane) : 0.f;  // of column n0 + lane (shuffled to the rows below)
            mbar_wait(&acc_full[acc], (uint32_t)(i >> 1) & 1u);
            tc_fence_after();
            uint32_t r0[32];
            tmem_ld32(tmem_base + lane_off + (uint32_t)(acc * BN + cq * 32), r0);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);  // the accumulator is in registers: the next tile but one may overwrite it
            // thread = row: 32 bf16 (64 bytes) into the warp's staging tile, 16-byte pieces rotated by the row to spread the banks
            uint8_t* prow = stg + lane * 64;
#pragma unroll
            for (int pc = 0; pc < 4; ++pc) {
                uint32_t w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int col = pc * 8 + 2 * e;
                    const float b0 = __shfl_sync(0xffffffffu, bias, col);
                    const float b1 = __shfl_sync(0xffffffffu, bias, col + 1);
                    __nv_bfloat162 h2 = __floats2bfloat162_rn(__uint_as_float(r0[col]) + b0, __uint_as_float(r0[col + 1]) + b1);
                    w[e] = *reinterpret_cast<uint32_t*>(&h2);
                }
                *reinterpret_cast<uint4*>(prow + (((pc + (lane >> 1)) & 3) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
            }
            __syncwarp();
            // lanes 4 j .. 4 j + 3 write the four 16-byte pieces of one row: eight rows per instruction
            const int64_t row_base = (int64_t)mt * BM + q * 32;
#pragma unroll
            for (int rr0 = 0; rr0 < 32; rr0 += 8) {
                const int rr = rr0 + (lane >> 2), pc = lane & 3;
                const int64_t gm = row_base + rr;
                const int col0 = n0 + pc * 8;
                if (gm < p.M && col0 < p.N) {
                    uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 64 + (((pc + (rr >> 1)) & 3) << 4));
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
        PIO_CUDA(cudaFuncSetAttribute(logits_persistent_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        PIO_CUDA(cudaFuncSetAttribute(logits_persistent_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    const int64_t tiles = cdiv(m, (int64_t)BM) * cdiv(n, (int64_t)BN);
    const unsigned grid = (unsigned)std::min<int64_t>(tiles, num_sms());
    PIO_CUDA(launch_pdl(logits_persistent_kernel<false>, dim3(grid), dim3(THREADS), SMEM_BYTES, st, tmA, tmB, p));
    count_launch();
    return PIO_OK;
}

bool ln_linear_supported(int64_t m, int64_t n, int64_t k, int64_t ldw, int64_t ldc, int64_t ldh) {
    return getenv("PIO_NO_LN_LINEAR") == nullptr && k == 256 && m > 0 && m < (1ll << 30) && n >= 128 && n < (1ll << 30) && ldw % 8 == 0 &&
           ldc % 8 == 0 && ldc >= (n + 7) / 8 * 8 && ldh % 8 == 0 && ldh >= k;
}

// c (m x n, bf16) = LayerNorm(x; gamma, beta) w^T + bias with x (m x 256, fp32, dense); also writes the LayerNorm output
// h0 (m x 256, bf16, pitch ldh) and stats (m x 2: mean, rstd) that the backward pass reads
int ln_linear(const float* x, const float* gamma, const float* beta, float eps, const void* w, int64_t ldw, const float* bias, void* c,
              int64_t ldc, void* h0, int64_t ldh, float* stats, int64_t m, int64_t n, cudaStream_t st) {
    using namespace lg;
    PIO_CHECK_ARG(x && gamma && beta && w && c && h0 && stats && ln_linear_supported(m, n, 256, ldw, ldc, ldh) &&
                      ((((uintptr_t)c) | ((uintptr_t)h0) | ((uintptr_t)x)) & 15) == 0,
                  "ln_linear: unsupported arguments (m %lld n %lld)", (long long)m, (long long)n);
    CUtensorMap tmA, tmB;
    PIO_TRY(make_tmap_bf16(&tmB, w, n, 256, ldw, BN));
    tmA = tmB;  // (unused)
    Params p{};
    p.c = (__nv_bfloat16*)c;
    p.ldc = ldc;
    p.bias = bias;
    p.M = (int)m;
    p.N = (int)n;
    p.num_kb = 4;
    p.x = x; p.gamma = gamma; p.beta = beta; p.h0 = (__nv_bfloat16*)h0; p.ldh = ldh; p.stats = stats; p.eps = eps;
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(logits_persistent_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    // whole row tiles per CTA, the same number on every CTA when possible (256 row tiles: 128 CTAs x 2)
    const int64_t m_tiles = cdiv(m, (int64_t)BM), n_tiles = cdiv(n, (int64_t)BN);
    const int64_t per_cta = cdiv(m_tiles, (int64_t)num_sms());
    const unsigned grid = (unsigned)cdiv(m_tiles, per_cta);
    (void)n_tiles;
    PIO_CUDA(launch_pdl(logits_persistent_kernel<true>, dim3(grid), dim3(THREADS), SMEM_BYTES, st, tmA, tmB, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
