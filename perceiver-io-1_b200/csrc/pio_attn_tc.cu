// Flash-style attention on the 5th-gen tensor cores (PIO_MODE_BF16, head dim 32 / 64 / 128): forward, dQ and dK/dV.
//
// Forward: one CTA owns 128 query rows of one (batch, head) and streams the key/value axis in tiles of 64 keys:
//   * Q, K and V tiles arrive by 3-D TMA (cp.async.bulk.tensor, 128-byte swizzle; batch and row are separate tensor
//     dimensions, so rows past the end of a sample are zero-filled), K/V through a 3-stage ring;
//   * S = Q K^T is a tcgen05.mma (M=128, N=64) into one of two TMEM buffers; four softmax warps read it with tcgen05.ld
//     (one query row per thread), apply scale, padding mask and the online softmax with ex2.approx against a LAZILY
//     updated reference maximum (O and the running sum are rescaled only when the maximum grew by more than 2^8), and
//     hand P to the P.V product as bf16 pairs THROUGH TENSOR MEMORY (tcgen05.st; the MMA takes its A operand from
//     TMEM: no shared-memory round trip);
//   * O += P V accumulates in TMEM across all key tiles (V is consumed MN-major, exactly as TMA delivered it); the
//     schedule is software-pipelined: S(j+1) is issued before P(j) is awaited;
//   * two CTAs per SM (one at dh = 128: 512 TMEM columns).
// Backward (P recomputed from the saved log-sum-exp): a dQ kernel per 128 query rows and a dK/dV kernel per 128 key
// rows on the transposed tile; dS / P^T / dS^T reach their products through tensor memory as well (see below).
// Warp roles in all three: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..5 = softmax / math.
#include <cuda.h>

#include <atomic>
#include <stdlib.h>

#include "pio_ops_common.cuh"
#include "pio_tc_common.cuh"

namespace pio {

constexpr int FA_BQ = 128;
constexpr int FA_THREADS = 192;

struct FaParams {
    const uint8_t* pad;  // (batch, m) or null
    float* o32;          // optional
    __nv_bfloat16* o16;  // optional
    float* lse;          // (batch, heads, n)
    int64_t o_bs, o_rs;
    int batch, heads, n, m;
    float scale_log2e;
    DropCfg drop;  // dropout on the attention probabilities (threshold 0 = off)
};

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// descriptor of k-step t (16 elements) of a K-major operand tile of `rows` rows whose K extent may span several
// 64-column swizzle atoms (each atom = rows x 128 bytes, stored one after the other)
__device__ __forceinline__ uint64_t kmajor_desc(uint32_t base, int rows, int t) {
    return make_smem_desc(base + (uint32_t)(t >> 2) * (uint32_t)(rows * 128) + (uint32_t)(t & 3) * 32, 16, 1024);
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// a bf16 tile of `rows` rows as ATOMS swizzle atoms of rows x 64 columns, starting at column col0.  Head size 32
// loads the 64-column box that holds the head and its neighbour; the products then address the head's half of
// each 128-byte row (HeadGeom::sub) and run two k-steps instead of four.
template <int DH>
struct HeadGeom {
    static constexpr int ATOMS = (DH + 63) / 64;
    static constexpr int TILE_COLS = ATOMS * 64;
    static constexpr int KS = DH / 16;
    __device__ static int col0(int h) { return DH == 32 ? (h >> 1) * 64 : h * DH; }
    __device__ static uint32_t sub(int h) { return DH == 32 ? (uint32_t)(h & 1) * 64u : 0u; }
};
template <int DH>
__device__ __forceinline__ void tma_load_tile(uint8_t* dst, const CUtensorMap* map, int col0, int row0, int b, int rows, uint64_t* bar) {
#pragma unroll
    for (int a = 0; a < HeadGeom<DH>::ATOMS; ++a) tma_load_3d(dst + a * rows * 128, map, col0 + a * 64, row0, b, bar);
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// 32 fp32 values -> bf16, into row `row` of a 128-row x 128-byte swizzled (K-major, SW128) region;
// half = which 64-byte half of the 128-byte row (columns 0..31 or 32..63 of the 64-column region)
__device__ __forceinline__ void store_row_half_bf16(uint8_t* region, int row, int half, const float (&v)[32]) {
    uint8_t* prow = region + row * 128;
#pragma unroll
    for (int piece = 0; piece < 4; ++piece) {
        uint32_t w[4];
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
            __nv_bfloat162 h2 = __floats2bfloat162_rn(v[piece * 8 + 2 * e2], v[piece * 8 + 2 * e2 + 1]);
            w[e2] = *reinterpret_cast<uint32_t*>(&h2);
        }
        const int ci = half * 4 + piece;
        *reinterpret_cast<uint4*>(prow + ((ci ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

__device__ __forceinline__ void store_row64_bf16_global(__nv_bfloat16* dst, const float (&v)[32], int c0) {
#pragma unroll
    for (int d = 0; d < 32; d += 8) {
        uint32_t w[4];
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
            __nv_bfloat162 h2 = __floats2bfloat162_rn(v[d + 2 * e2], v[d + 2 * e2 + 1]);
            w[e2] = *reinterpret_cast<uint32_t*>(&h2);
        }
        *reinterpret_cast<uint4*>(dst + c0 + d) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// v[0..32) += the 32 bf16 values at dst + c0
__device__ __forceinline__ void add_row64_bf16_global(const __nv_bfloat16* dst, float (&v)[32], int c0) {
#pragma unroll
    for (int d = 0; d < 32; d += 8) {
        const uint4 w = *reinterpret_cast<const uint4*>(dst + c0 + d);
        const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
            const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ws[e2]));
            v[d + 2 * e2] += f.x;
            v[d + 2 * e2 + 1] += f.y;
        }
    }
}

template <int BKV, bool DROP, int DH>
__global__ void __launch_bounds__(FA_THREADS, DH == 128 ? 1 : 2)
    attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const FaParams p) {
    static_assert(BKV == 64, "the pipelined schedule is laid out for 64-key tiles");
    static_assert(DH == 32 || DH == 64 || DH == 128, "head size 32, 64 or 128");
    using G = HeadGeom<DH>;
    constexpr int KV_STAGES = 3;
    constexpr int Q_BYTES = FA_BQ * G::TILE_COLS * 2;   // 16 KB per atom
    constexpr int KV_BYTES = BKV * G::TILE_COLS * 2;    // 8 KB per atom
    // S0 | S1 (64 columns each) | O (DH columns) | P0 | P1 (bf16 pairs, 32 columns each)
    constexpr int TMEM_COLS = DH == 128 ? 512 : 256;
    constexpr uint32_t TM_O = 128, TM_P = 128 + DH;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + Q_BYTES;                  // [KV_STAGES][KV_BYTES]
    uint8_t* sV = sK + KV_STAGES * KV_BYTES;     // [KV_STAGES][KV_BYTES]
    float* sBias = (float*)(sV + KV_STAGES * KV_BYTES);  // [2][BKV]
    uint64_t* bars = (uint64_t*)(sBias + 2 * BKV);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;               // [3]
    uint64_t* kv_empty = bars + 4;              // [3]
    uint64_t* s_full = bars + 7;                // [2]
    uint64_t* o_full = bars + 9;                // [2]
    uint64_t* p_full = bars + 11;
    uint32_t* tmem_slot = (uint32_t*)(bars + 12);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * FA_BQ, h = blockIdx.y, b = blockIdx.z;
    const int T = (p.m + BKV - 1) / BKV;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmK);
        prefetch_tmap(&tmV);
        mbar_init(q_full, 1);
        for (int s = 0; s < KV_STAGES; ++s) {
            mbar_init(&kv_full[s], 1);
            mbar_init(&kv_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&s_full[s], 1);
            mbar_init(&o_full[s], 1);
        }
        mbar_init(p_full, 4);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // prologue done; everything below depends on earlier kernels' results

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(q_full, Q_BYTES);
            tma_load_tile<DH>(sQ, &tmQ, G::col0(h), q0, b, FA_BQ, q_full);
            for (int j = 0; j < T; ++j) {
                const int s = j % KV_STAGES;
                const uint32_t ph = (uint32_t)(j / KV_STAGES) & 1u;
                mbar_wait(&kv_empty[s], ph ^ 1u);
                mbar_expect_tx(&kv_full[s], 2 * KV_BYTES);
                tma_load_tile<DH>(sK + s * KV_BYTES, &tmK, G::col0(h), j * BKV, b, BKV, &kv_full[s]);
                tma_load_tile<DH>(sV + s * KV_BYTES, &tmV, G::col0(h), j * BKV, b, BKV, &kv_full[s]);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            // Software pipeline: S(j+1) = Q K(j+1)^T is issued before waiting for P(j), so the tensor core works
            // on the next score tile while the softmax warps are busy with the current one, and O(j) = P(j) V(j)
            // lands in the TMEM buffer the softmax warps only read one tile later.
            const uint32_t idesc_s = make_idesc(BKV, false, false);
            const uint32_t idesc_o = make_idesc(DH, false, true);
            const uint32_t q_addr = smem_u32(sQ) + G::sub(h);
            auto issue_s = [&](int j) {
                const int s = j % KV_STAGES;
                mbar_wait(&kv_full[s], (uint32_t)(j / KV_STAGES) & 1u);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(sK + s * KV_BYTES) + G::sub(h);
#pragma unroll
                for (int t = 0; t < G::KS; ++t)
                    umma_bf16(tmem_base + (uint32_t)(j & 1) * 64, kmajor_desc(q_addr, FA_BQ, t), kmajor_desc(k_addr, BKV, t), idesc_s,
                              t > 0 ? 1u : 0u);
                umma_commit(&s_full[j & 1]);
            };
            mbar_wait(q_full, 0);
            issue_s(0);
            for (int j = 0; j < T; ++j) {
                if (j + 1 < T) issue_s(j + 1);
                mbar_wait(p_full, (uint32_t)j & 1u);
                tc_fence_after();
                const int s = j % KV_STAGES;
                const uint32_t v_addr = smem_u32(sV + s * KV_BYTES) + G::sub(h);
                const uint32_t p_tmem = tmem_base + TM_P + (uint32_t)(j & 1) * 32;
#pragma unroll
                for (int t = 0; t < BKV / 16; ++t)  // P read from TMEM (no shared-memory A traffic); O accumulates in TMEM
                    umma_bf16_ts(tmem_base + TM_O, p_tmem + (uint32_t)t * 8, make_smem_desc(v_addr + t * 2048, BKV * 128, 1024),
                                 idesc_o, (j > 0 || t > 0) ? 1u : 0u);
                umma_commit(&o_full[j & 1]);
                umma_commit(&kv_empty[s]);
            }
        }
        __syncwarp();
    } else {
        // ---------------- softmax warps: thread = query row, TMEM lane quadrant = warp id mod 4
        const int qd = warp & 3;
        const int row = qd * 32 + lane;          // 0..127 within the tile
        const int tid = (warp - 2) * 32 + lane;  // 0..127 for cooperative loads
        const uint32_t lane_off = (uint32_t)(qd * 32) << 16;
        const uint8_t* padb = p.pad ? p.pad + (int64_t)b * p.m : nullptr;
        // Online softmax with a lazily updated reference maximum (m_ref, in log2 units): P = 2^(s - m_ref) may
        // grow to 2^8 before the row's O accumulator (in TMEM) and running sum are rescaled, so after the first
        // tiles almost no tile pays for a rescale and none pays for folding O through registers.
        constexpr float LAZY = 8.0f;
        float m_ref = -INFINITY, l_run = 0.f;
        const uint32_t to = tmem_base + TM_O + lane_off;
        for (int j = 0; j < T; ++j) {
            const int j0 = j * BKV;
            const bool masked = (padb != nullptr) || (j0 + BKV > p.m);  // CTA-uniform
            float* bias = sBias + (j & 1) * BKV;
            if (masked) {
                if (tid < BKV) {
                    const int key = j0 + tid;
                    const bool dead = (key >= p.m) || (padb != nullptr && padb[key] != 0);
                    bias[tid] = dead ? -INFINITY : 0.f;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            mbar_wait(&s_full[j & 1], (uint32_t)(j >> 1) & 1u);
            tc_fence_after();
            const uint32_t ts = tmem_base + (uint32_t)(j & 1) * 64 + lane_off;
            // the whole 64-column score row lives in registers: one TMEM read, max, then exp
            uint32_t r0[32], r1[32];
            tmem_ld32(ts, r0);
            tmem_ld32(ts + 32, r1);
            tmem_ld_wait();
            float sv[64];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                sv[i] = __uint_as_float(r0[i]);
                sv[32 + i] = __uint_as_float(r1[i]);
            }
            if (masked) {
#pragma unroll
                for (int i = 0; i < 64; ++i) sv[i] += bias[i];
            }
            float mx = sv[0];
#pragma unroll
            for (int i = 1; i < 64; ++i) mx = fmaxf(mx, sv[i]);
            mx *= p.scale_log2e;  // scale > 0: max commutes with it
            // P(j) goes to the buffer PV(j-2) read: that product must be done
            if (j >= 2) mbar_wait(&o_full[j & 1], (uint32_t)((j - 2) >> 1) & 1u);
            const bool grow = mx > m_ref + LAZY || (m_ref == -INFINITY && mx != -INFINITY);
            if (__any_sync(0xffffffffu, grow)) {
                const float m_new = grow ? mx : m_ref;
                const float alpha = (m_ref == -INFINITY) ? 0.f : fast_exp2(m_ref - m_new);  // 1 for rows that keep m_ref
                if (j > 0) {  // O holds the sum over tiles < j: rescale this row of it
                    mbar_wait(&o_full[(j - 1) & 1], (uint32_t)((j - 1) >> 1) & 1u);
                    tc_fence_after();
#pragma unroll
                    for (int c0 = 0; c0 < DH; c0 += 32) {
                        uint32_t r[32];
                        tmem_ld32(to + (uint32_t)c0, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        tmem_st32(to + (uint32_t)c0, r);
                    }
                    tmem_st_wait();
                }
                l_run *= alpha;
                m_ref = m_new;
            }
            const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;  // fully masked so far: every p below is 0
            const float neg_m = -m_use;
            float l_tile = 0.f;
            uint32_t pw[32];  // this row of P as bf16 pairs
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                float pv[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    pv[i] = fast_exp2(fmaf(sv[half * 32 + i], p.scale_log2e, neg_m));
                    l_tile += pv[i];
                }
                if constexpr (DROP) {  // the normaliser sees every key; only P.V sees the mask
                    const uint64_t base = (((uint64_t)b * p.heads + h) * p.n + (uint64_t)(q0 + row)) * (uint64_t)p.m + (uint64_t)(j0 + half * 32);
#pragma unroll
                    for (int i = 0; i < 32; ++i) pv[i] = drop_keep(p.drop, base + i) ? pv[i] * p.drop.inv_keep : 0.f;
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    __nv_bfloat162 h2 = __floats2bfloat162_rn(pv[2 * i], pv[2 * i + 1]);
                    pw[half * 16 + i] = *reinterpret_cast<uint32_t*>(&h2);
                }
            }
            tmem_st32(tmem_base + TM_P + (uint32_t)(j & 1) * 32 + lane_off, pw);
            tmem_st_wait();
            l_run += l_tile;
            tc_fence_before();  // P is in TMEM, the reads of S(j) (and the rescale of O) are done: the MMA warp may go on
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
        }
        mbar_wait(&o_full[(T - 1) & 1], (uint32_t)((T - 1) >> 1) & 1u);
        tc_fence_after();
        const float m_run = m_ref;
        const int qi = q0 + row;
        const float inv = 1.0f / l_run;  // l = 0 (every key padded): inf * 0 = NaN, like the reference
        const int64_t off = (int64_t)b * p.o_bs + (int64_t)qi * p.o_rs + (int64_t)h * DH;
#pragma unroll
        for (int c0 = 0; c0 < DH; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(to + (uint32_t)c0, r);  // warp-collective: every lane takes part, valid row or not
            tmem_ld_wait();
            if (qi < p.n) {
                float acc[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(r[i]) * inv;
                if (p.o32) {
                    float* o = p.o32 + off + c0;
#pragma unroll
                    for (int d = 0; d < 32; d += 4) *reinterpret_cast<float4*>(o + d) = make_float4(acc[d], acc[d + 1], acc[d + 2], acc[d + 3]);
                }
                if (p.o16) store_row64_bf16_global(p.o16 + off, acc, c0);
            }
        }
        tc_fence_before();
        if (qi < p.n) {
            p.lse[((int64_t)b * p.heads + h) * p.n + qi] = (m_run + log2f(l_run)) * 0.69314718055994530942f;
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------
// Backward.  P is recomputed from the saved log-sum-exp (nothing of size N x M was stored).  Two
// kernels, mirroring the forward: one owns query rows and produces dQ (and the row term
// delta = dO.O), one owns key rows and produces dK and dV — so no gradient needs atomics.
// All five products run on tcgen05 with their accumulators in TMEM; dQ, dK and dV accumulate across
// the streamed tiles directly in TMEM (no online rescaling exists in the backward).
// ---------------------------------------------------------------------------------------------
struct FaBwdParams {
    const uint8_t* pad;      // (batch, m) or null
    const float* o32;        // forward output (for delta)
    const __nv_bfloat16* o16;  // the same as bf16: used instead of o32 when given (the fused latent block saves only this)
    const __nv_bfloat16* do16;
    const float* lse;        // (batch, heads, n), natural log
    float* delta;            // (batch, heads, n): written by the dQ kernel, read by the dK/dV kernel
    __nv_bfloat16* dq16;
    __nv_bfloat16* dk16;
    __nv_bfloat16* dv16;
    int64_t o_bs, o_rs, q_bs, q_rs, k_bs, k_rs, v_bs, v_rs;
    int batch, heads, n, m;
    float scale, scale_log2e;
    DropCfg drop;
    int dkv_add;             // dK / dV are added to what dk16 / dv16 hold (a second consumer of the same keys and values)
};

// dQ: CTA = 128 query rows of one (batch, head); streams key tiles of 64.
template <bool DROP, int DH>
__global__ void __launch_bounds__(FA_THREADS, DH == 128 ? 1 : 2)
    attn_bwd_dq_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                          const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const FaBwdParams p) {
    static_assert(DH == 32 || DH == 64 || DH == 128, "head size 32, 64 or 128");
    using G = HeadGeom<DH>;
    constexpr int BKV = 64;
    constexpr int Q_BYTES = FA_BQ * G::TILE_COLS * 2, KV_BYTES = BKV * G::TILE_COLS * 2, KV_STAGES = 3;
    constexpr int TMEM_COLS = DH == 128 ? 512 : 256;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sdO = sQ + Q_BYTES;
    uint8_t* sK = sdO + Q_BYTES;              // [KV_STAGES]
    uint8_t* sV = sK + KV_STAGES * KV_BYTES;  // [KV_STAGES]
    float* sBias = (float*)(sV + KV_STAGES * KV_BYTES);  // [2][BKV]
    uint64_t* bars = (uint64_t*)(sBias + 2 * BKV);
    uint64_t* qdo_full = bars;
    uint64_t* kv_full = bars + 1;   // [3]
    uint64_t* kv_empty = bars + 4;  // [3]
    uint64_t* sdp_full = bars + 7;
    uint64_t* sdp_read = bars + 8;  // the math warps hold S(j), dP(j) in registers: the next pair may be issued
    uint64_t* ds_full = bars + 9;
    uint64_t* dq_done = bars + 10;  // [2], one per dS buffer
    uint32_t* tmem_slot = (uint32_t*)(bars + 12);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * FA_BQ, h = blockIdx.y, b = blockIdx.z;
    const int T = (p.m + BKV - 1) / BKV;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ); prefetch_tmap(&tmdO); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
        mbar_init(qdo_full, 1);
        for (int s = 0; s < KV_STAGES; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
        mbar_init(sdp_full, 1);
        mbar_init(sdp_read, 4);
        mbar_init(ds_full, 4);
        mbar_init(&dq_done[0], 1);
        mbar_init(&dq_done[1], 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // prologue done; everything below depends on earlier kernels' results
    // S | dP (64 columns each) | dQ (DH columns) | dS0 | dS1 (bf16 pairs, 32 columns each)
    const uint32_t tmem_s = tmem_base, tmem_dp = tmem_base + 64, tmem_dq = tmem_base + 128, tmem_ds = tmem_base + 128 + DH;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(qdo_full, 2 * Q_BYTES);
            tma_load_tile<DH>(sQ, &tmQ, G::col0(h), q0, b, FA_BQ, qdo_full);
            tma_load_tile<DH>(sdO, &tmdO, G::col0(h), q0, b, FA_BQ, qdo_full);
            for (int j = 0; j < T; ++j) {
                const int s = j % KV_STAGES;
                const uint32_t ph = (uint32_t)(j / KV_STAGES) & 1u;
                mbar_wait(&kv_empty[s], ph ^ 1u);
                mbar_expect_tx(&kv_full[s], 2 * KV_BYTES);
                tma_load_tile<DH>(sK + s * KV_BYTES, &tmK, G::col0(h), j * BKV, b, BKV, &kv_full[s]);
                tma_load_tile<DH>(sV + s * KV_BYTES, &tmV, G::col0(h), j * BKV, b, BKV, &kv_full[s]);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc_kk = make_idesc(BKV, false, false);  // S, dP: both operands K-major
            const uint32_t idesc_dq = make_idesc(DH, false, true);    // dQ: dS K-major, K tile MN-major
            mbar_wait(qdo_full, 0);
            const uint32_t q_addr = smem_u32(sQ) + G::sub(h), do_addr = smem_u32(sdO) + G::sub(h);
            // Software pipeline: S(j+1), dP(j+1) are issued as soon as the math warps have S(j), dP(j) in registers,
            // so the tensor core works on the next score tile while they turn the current one into dS(j).
            auto issue_sdp = [&](int j) {
                const int s = j % KV_STAGES;
                mbar_wait(&kv_full[s], (uint32_t)(j / KV_STAGES) & 1u);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(sK + s * KV_BYTES) + G::sub(h), v_addr = smem_u32(sV + s * KV_BYTES) + G::sub(h);
#pragma unroll
                for (int t = 0; t < G::KS; ++t)
                    umma_bf16(tmem_s, kmajor_desc(q_addr, FA_BQ, t), kmajor_desc(k_addr, BKV, t), idesc_kk, t > 0 ? 1u : 0u);
#pragma unroll
                for (int t = 0; t < G::KS; ++t)
                    umma_bf16(tmem_dp, kmajor_desc(do_addr, FA_BQ, t), kmajor_desc(v_addr, BKV, t), idesc_kk, t > 0 ? 1u : 0u);
                umma_commit(sdp_full);
            };
            issue_sdp(0);
            for (int j = 0; j < T; ++j) {
                const int s = j % KV_STAGES;
                if (j + 1 < T) {
                    mbar_wait(sdp_read, (uint32_t)j & 1u);
                    tc_fence_after();
                    issue_sdp(j + 1);
                }
                mbar_wait(ds_full, (uint32_t)j & 1u);
                tc_fence_after();
                const uint32_t k_addr = smem_u32(sK + s * KV_BYTES) + G::sub(h);
                const uint32_t ds_tmem = tmem_ds + (uint32_t)(j & 1) * 32;
#pragma unroll
                for (int t = 0; t < BKV / 16; ++t)  // A = dS from tensor memory
                    umma_bf16_ts(tmem_dq, ds_tmem + (uint32_t)t * 8, make_smem_desc(k_addr + t * 2048, BKV * 128, 1024), idesc_dq,
                                 (j > 0 || t > 0) ? 1u : 0u);
                umma_commit(&dq_done[j & 1]);
                umma_commit(&kv_empty[s]);
            }
        }
        __syncwarp();
    } else {
        const int qd = warp & 3;
        const int row = qd * 32 + lane;
        const int tid = (warp - 2) * 32 + lane;
        const uint32_t lane_off = (uint32_t)(qd * 32) << 16;
        const uint8_t* padb = p.pad ? p.pad + (int64_t)b * p.m : nullptr;
        const int qi = q0 + row;
        const bool q_ok = qi < p.n;
        // delta = dO . O for this row; lse in the base-2 domain
        float delta = 0.f, lse2 = INFINITY;  // invalid rows: p = exp2(-inf) = 0
        if (q_ok) {
            const int64_t off = (int64_t)b * p.o_bs + (int64_t)qi * p.o_rs + (int64_t)h * DH;
            const __nv_bfloat16* g = p.do16 + off;
#pragma unroll
            for (int d = 0; d < DH; d += 8) {
                float4 o0, o1;
                if (p.o16 != nullptr) {
                    const uint4 uo = *reinterpret_cast<const uint4*>(p.o16 + off + d);
                    const __nv_bfloat162* ob = reinterpret_cast<const __nv_bfloat162*>(&uo);
                    o0 = make_float4(__low2float(ob[0]), __high2float(ob[0]), __low2float(ob[1]), __high2float(ob[1]));
                    o1 = make_float4(__low2float(ob[2]), __high2float(ob[2]), __low2float(ob[3]), __high2float(ob[3]));
                } else {
                    o0 = *reinterpret_cast<const float4*>(p.o32 + off + d);
                    o1 = *reinterpret_cast<const float4*>(p.o32 + off + d + 4);
                }
                const uint4 u = *reinterpret_cast<const uint4*>(g + d);
                const __nv_bfloat162* gb = reinterpret_cast<const __nv_bfloat162*>(&u);
                delta += o0.x * __low2float(gb[0]) + o0.y * __high2float(gb[0]) + o0.z * __low2float(gb[1]) + o0.w * __high2float(gb[1]) +
                         o1.x * __low2float(gb[2]) + o1.y * __high2float(gb[2]) + o1.z * __low2float(gb[3]) + o1.w * __high2float(gb[3]);
            }
            const int64_t si = ((int64_t)b * p.heads + h) * p.n + qi;
            lse2 = p.lse[si] * 1.44269504088896340736f;
            p.delta[si] = delta;
        }
        for (int j = 0; j < T; ++j) {
            const int j0 = j * BKV;
            const bool masked = (padb != nullptr) || (j0 + BKV > p.m);
            float* bias = sBias + (j & 1) * BKV;
            if (masked) {
                if (tid < BKV) {
                    const int key = j0 + tid;
                    bias[tid] = ((key >= p.m) || (padb != nullptr && padb[key] != 0)) ? -INFINITY : 0.f;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            mbar_wait(sdp_full, (uint32_t)j & 1u);
            tc_fence_after();
            if (T == 1) {
                // One key tile holds the whole row (latent self-attention): delta = sum_k p_k dP_k from the probabilities that
                // are recomputed below, so that sum_k dS_k = 0 to fp32 rounding.  (dO . O from a bf16 copy of O leaves a
                // common-mode error along the mean key in dQ; with several key tiles the sum would need a pass of its own and
                // the fp32 / bf16 output is used, above.)
                float dsum = 0.f;
#pragma unroll
                for (int c0 = 0; c0 < BKV; c0 += 32) {
                    uint32_t rs[32], rp[32];
                    tmem_ld32(tmem_s + lane_off + (uint32_t)c0, rs);
                    tmem_ld32(tmem_dp + lane_off + (uint32_t)c0, rp);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        float e = __uint_as_float(rs[i]) * p.scale_log2e - lse2;
                        if (masked) e += bias[c0 + i];
                        float dpi = __uint_as_float(rp[i]);
                        if constexpr (DROP) {
                            const uint64_t idx = (((uint64_t)b * p.heads + h) * p.n + (uint64_t)qi) * (uint64_t)p.m + (uint64_t)(j0 + c0 + i);
                            dpi = drop_keep(p.drop, idx) ? dpi * p.drop.inv_keep : 0.f;
                        }
                        dsum = fmaf(fast_exp2(e), dpi, dsum);
                    }
                }
                delta = dsum;
                if (q_ok) p.delta[((int64_t)b * p.heads + h) * p.n + qi] = delta;
            }
            if (j >= 2) mbar_wait(&dq_done[j & 1], (uint32_t)((j - 2) >> 1) & 1u);  // dQ(j-2) has consumed this dS buffer
            uint32_t dw[32];  // this row of dS as bf16 pairs
#pragma unroll
            for (int c0 = 0; c0 < BKV; c0 += 32) {
                uint32_t rs[32], rp[32];
                tmem_ld32(tmem_s + lane_off + (uint32_t)c0, rs);
                tmem_ld32(tmem_dp + lane_off + (uint32_t)c0, rp);
                tmem_ld_wait();
                if (c0 + 32 == BKV) {  // S(j), dP(j) are out of tensor memory: let the MMA warp issue the next pair
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(sdp_read);
                }
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float ds2[2];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        float e = __uint_as_float(rs[i + u]) * p.scale_log2e - lse2;
                        if (masked) e += bias[c0 + i + u];
                        const float pv = fast_exp2(e);
                        float dpi = __uint_as_float(rp[i + u]);
                        if constexpr (DROP) {
                            const uint64_t idx = (((uint64_t)b * p.heads + h) * p.n + (uint64_t)qi) * (uint64_t)p.m + (uint64_t)(j0 + c0 + i + u);
                            dpi = drop_keep(p.drop, idx) ? dpi * p.drop.inv_keep : 0.f;
                        }
                        ds2[u] = pv * (dpi - delta) * p.scale;
                    }
                    __nv_bfloat162 hd = __floats2bfloat162_rn(ds2[0], ds2[1]);
                    dw[(c0 + i) >> 1] = *reinterpret_cast<uint32_t*>(&hd);
                }
            }
            tmem_st32(tmem_ds + (uint32_t)(j & 1) * 32 + lane_off, dw);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(ds_full);
        }
        mbar_wait(&dq_done[(T - 1) & 1], (uint32_t)((T - 1) >> 1) & 1u);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < DH; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_dq + lane_off + (uint32_t)c0, r);
            tmem_ld_wait();
            if (q_ok) {
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                store_row64_bf16_global(p.dq16 + (int64_t)b * p.q_bs + (int64_t)qi * p.q_rs + (int64_t)h * DH, v, c0);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// dK, dV: CTA = 128 key rows of one (batch, head); streams query tiles of 64.  Works on the
// transposed score tile S^T = K Q^T (thread = key row, column = query), so P^T and dS^T come out in
// exactly the K-major layout the dV = P^T dO and dK = dS^T Q products read.
template <bool DROP, int DH>
__global__ void __launch_bounds__(FA_THREADS, DH == 128 ? 1 : 2)
    attn_bwd_dkv_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                           const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const FaBwdParams p) {
    static_assert(DH == 32 || DH == 64 || DH == 128, "head size 32, 64 or 128");
    using G = HeadGeom<DH>;
    constexpr int BQ2 = 64, BK2 = 128;
    constexpr int K_BYTES = BK2 * G::TILE_COLS * 2, Q_BYTES = BQ2 * G::TILE_COLS * 2;
    constexpr int TMEM_COLS = DH == 128 ? 512 : 256;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sK = smem;
    uint8_t* sV = sK + K_BYTES;
    uint8_t* sQ = sV + K_BYTES;        // [2]
    uint8_t* sdO = sQ + 2 * Q_BYTES;   // [2]
    float* sLse = (float*)(sdO + 2 * Q_BYTES);  // [2][BQ2]
    float* sDelta = sLse + 2 * BQ2;         // [2][BQ2]
    uint64_t* bars = (uint64_t*)(sDelta + 2 * BQ2);
    uint64_t* kv_full = bars;
    uint64_t* q_full = bars + 1;   // [2]
    uint64_t* q_empty = bars + 3;  // [2]
    uint64_t* sdp_full = bars + 5;
    uint64_t* pds_full = bars + 6;
    uint64_t* acc_done = bars + 7;
    uint32_t* tmem_slot = (uint32_t*)(bars + 8);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k0 = blockIdx.x * BK2, h = blockIdx.y, b = blockIdx.z;
    const int T = (p.n + BQ2 - 1) / BQ2;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ); prefetch_tmap(&tmdO); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
        mbar_init(kv_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
        mbar_init(sdp_full, 1);
        mbar_init(pds_full, 4);
        mbar_init(acc_done, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // prologue done; everything below depends on earlier kernels' results
    // S^T | dP^T (64 columns each) | dV | dK (DH columns each)
    const uint32_t tmem_st = tmem_base, tmem_dpt = tmem_base + 64, tmem_dv = tmem_base + 128, tmem_dk = tmem_base + 128 + DH;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(kv_full, 2 * K_BYTES);
            tma_load_tile<DH>(sK, &tmK, G::col0(h), k0, b, BK2, kv_full);
            tma_load_tile<DH>(sV, &tmV, G::col0(h), k0, b, BK2, kv_full);
            for (int i = 0; i < T; ++i) {
                const int s = i & 1;
                const uint32_t ph = (uint32_t)(i >> 1) & 1u;
                mbar_wait(&q_empty[s], ph ^ 1u);
                mbar_expect_tx(&q_full[s], 2 * Q_BYTES);
                tma_load_tile<DH>(sQ + s * Q_BYTES, &tmQ, G::col0(h), i * BQ2, b, BQ2, &q_full[s]);
                tma_load_tile<DH>(sdO + s * Q_BYTES, &tmdO, G::col0(h), i * BQ2, b, BQ2, &q_full[s]);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc_kk = make_idesc(BQ2, false, false);  // S^T, dP^T: K-major x K-major, N = 64 queries
            const uint32_t idesc_acc = make_idesc(DH, false, true);   // dV, dK: P^T / dS^T K-major, dO / Q tile MN-major
            mbar_wait(kv_full, 0);
            const uint32_t k_addr = smem_u32(sK) + G::sub(h), v_addr = smem_u32(sV) + G::sub(h);
            for (int i = 0; i < T; ++i) {
                const int s = i & 1;
                const uint32_t ph = (uint32_t)(i >> 1) & 1u;
                mbar_wait(&q_full[s], ph);
                tc_fence_after();
                const uint32_t q_addr = smem_u32(sQ + s * Q_BYTES) + G::sub(h), do_addr = smem_u32(sdO + s * Q_BYTES) + G::sub(h);
#pragma unroll
                for (int t = 0; t < G::KS; ++t)
                    umma_bf16(tmem_st, kmajor_desc(k_addr, BK2, t), kmajor_desc(q_addr, BQ2, t), idesc_kk, t > 0 ? 1u : 0u);
#pragma unroll
                for (int t = 0; t < G::KS; ++t)
                    umma_bf16(tmem_dpt, kmajor_desc(v_addr, BK2, t), kmajor_desc(do_addr, BQ2, t), idesc_kk, t > 0 ? 1u : 0u);
                umma_commit(sdp_full);
                mbar_wait(pds_full, (uint32_t)i & 1u);
                tc_fence_after();
#pragma unroll
                for (int t = 0; t < BQ2 / 16; ++t)  // A = P^T from tensor memory (it overwrote S^T in place)
                    umma_bf16_ts(tmem_dv, tmem_st + (uint32_t)t * 8, make_smem_desc(do_addr + t * 2048, BQ2 * 128, 1024), idesc_acc,
                                 (i > 0 || t > 0) ? 1u : 0u);
#pragma unroll
                for (int t = 0; t < BQ2 / 16; ++t)  // A = dS^T from tensor memory (over dP^T)
                    umma_bf16_ts(tmem_dk, tmem_dpt + (uint32_t)t * 8, make_smem_desc(q_addr + t * 2048, BQ2 * 128, 1024), idesc_acc,
                                 (i > 0 || t > 0) ? 1u : 0u);
                umma_commit(acc_done);
                umma_commit(&q_empty[s]);
            }
        }
        __syncwarp();
    } else {
        const int qd = warp & 3;
        const int row = qd * 32 + lane;  // key row within the tile
        const int tid = (warp - 2) * 32 + lane;
        const uint32_t lane_off = (uint32_t)(qd * 32) << 16;
        const int key = k0 + row;
        const bool key_ok = key < p.m;
        const bool dead_key = !key_ok || (p.pad != nullptr && p.pad[(int64_t)b * p.m + key] != 0);
        const int64_t stat_base = ((int64_t)b * p.heads + h) * p.n;
        // lse / delta of a query tile are staged through shared memory one tile ahead, so their global loads are
        // off the MMA -> math -> MMA chain
        auto stage_stats = [&](int i) {
            if (tid < BQ2) {
                const int qi = i * BQ2 + tid;
                sLse[(i & 1) * BQ2 + tid] = (qi < p.n) ? p.lse[stat_base + qi] * 1.44269504088896340736f : INFINITY;  // invalid query: p = 0
                sDelta[(i & 1) * BQ2 + tid] = (qi < p.n) ? p.delta[stat_base + qi] : 0.f;
            }
        };
        stage_stats(0);
        for (int i = 0; i < T; ++i) {
            const float* lse_s = sLse + (i & 1) * BQ2;
            const float* del_s = sDelta + (i & 1) * BQ2;
            asm volatile("bar.sync 1, 128;" ::: "memory");  // stats of tile i visible; everyone is done with tile i-1's
            if (i + 1 < T) stage_stats(i + 1);
            mbar_wait(sdp_full, (uint32_t)i & 1u);
            tc_fence_after();
            // P^T and dS^T go back into the TMEM columns S^T / dP^T came from (as bf16 pairs, 32 columns each) and feed
            // the dV / dK products from there.  The next S^T / dP^T products are issued after those, on the same
            // in-order tensor pipe, so nothing is overwritten early.
            uint32_t pw[32], dw[32];
#pragma unroll
            for (int c0 = 0; c0 < BQ2; c0 += 32) {
                uint32_t rs[32], rp[32];
                tmem_ld32(tmem_st + lane_off + (uint32_t)c0, rs);
                tmem_ld32(tmem_dpt + lane_off + (uint32_t)c0, rp);
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 32; q += 2) {
                    float pv2[2], ds2[2];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const float pr = dead_key ? 0.f : fast_exp2(__uint_as_float(rs[q + u]) * p.scale_log2e - lse_s[c0 + q + u]);
                        float keep_scale = 1.f;
                        if constexpr (DROP) {
                            const uint64_t idx = (((uint64_t)b * p.heads + h) * p.n + (uint64_t)(i * BQ2 + c0 + q + u)) * (uint64_t)p.m + (uint64_t)key;
                            keep_scale = drop_keep(p.drop, idx) ? p.drop.inv_keep : 0.f;
                        }
                        pv2[u] = pr * keep_scale;
                        ds2[u] = pr * (__uint_as_float(rp[q + u]) * keep_scale - del_s[c0 + q + u]) * p.scale;
                    }
                    __nv_bfloat162 hp = __floats2bfloat162_rn(pv2[0], pv2[1]), hd = __floats2bfloat162_rn(ds2[0], ds2[1]);
                    pw[(c0 + q) >> 1] = *reinterpret_cast<uint32_t*>(&hp);
                    dw[(c0 + q) >> 1] = *reinterpret_cast<uint32_t*>(&hd);
                }
            }
            tmem_st32(tmem_st + lane_off, pw);
            tmem_st32(tmem_dpt + lane_off, dw);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(pds_full);
        }
        mbar_wait(acc_done, (uint32_t)(T - 1) & 1u);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < DH; c0 += 32) {
            uint32_t rv[32], rk[32];
            tmem_ld32(tmem_dv + lane_off + (uint32_t)c0, rv);
            tmem_ld32(tmem_dk + lane_off + (uint32_t)c0, rk);
            tmem_ld_wait();
            if (key_ok) {
                float v[32];
#pragma unroll
                for (int q = 0; q < 32; ++q) v[q] = __uint_as_float(rv[q]);
                __nv_bfloat16* dv_row = p.dv16 + (int64_t)b * p.v_bs + (int64_t)key * p.v_rs + (int64_t)h * DH;
                if (p.dkv_add) add_row64_bf16_global(dv_row, v, c0);
                store_row64_bf16_global(dv_row, v, c0);
#pragma unroll
                for (int q = 0; q < 32; ++q) v[q] = __uint_as_float(rk[q]);
                __nv_bfloat16* dk_row = p.dk16 + (int64_t)b * p.k_bs + (int64_t)key * p.k_rs + (int64_t)h * DH;
                if (p.dkv_add) add_row64_bf16_global(dk_row, v, c0);
                store_row64_bf16_global(dk_row, v, c0);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// 3-D bf16 tensor map over (batch, rows, cols) with element strides (bs, rs, 1); box = 64 cols x box_rows x 1
static int make_tmap3(CUtensorMap* map, const void* base, int64_t cols, int64_t rows, int64_t batch, int64_t rs, int64_t bs,
                      int box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled entry point not available from the driver");
        return PIO_ERR_CUDA;
    }
    PIO_CHECK_ARG((((uintptr_t)base) & 15) == 0 && rs % 8 == 0 && bs % 8 == 0,
                  "attention (bf16): operand base/strides must be 16-byte aligned (base %p, row stride %lld, batch stride %lld)",
                  base, (long long)rs, (long long)bs);
    cuuint64_t gdim[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)batch};
    cuuint64_t gstr[2] = {(cuuint64_t)rs * 2, (cuuint64_t)bs * 2};
    cuuint32_t box[3] = {64u, (cuuint32_t)box_rows, 1u};
    cuuint32_t estr[3] = {1u, 1u, 1u};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim, gstr, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (3-D) failed with CUresult %d", (int)r);
        return PIO_ERR_CUDA;
    }
    return PIO_OK;
}

bool attn_fwd_tc_supported(const pio_attn_desc* d) {
    static const bool disabled = getenv("PIO_DISABLE_TC_ATTN") != nullptr;  // debugging aid: force the fp32-arithmetic kernel
    return !disabled && (d->dh == 32 || d->dh == 64 || d->dh == 128) && d->n > 0 && d->m > 0 && d->batch > 0;
}

template <int DH>
static int launch_fwd(dim3 grid, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const FaParams& p) {
    constexpr int BKV = 64;
    constexpr int TC = HeadGeom<DH>::TILE_COLS;
    const size_t smem = (size_t)FA_BQ * TC * 2 + 6 * (size_t)BKV * TC * 2 + 2 * BKV * 4 + 128 + 1024;
    static std::atomic<bool> configured[64];  // per head size
    int dev_ = 0;
    PIO_CUDA(cudaGetDevice(&dev_));  // the attribute is per function AND per device
    dev_ = (dev_ >= 0 && dev_ < 64) ? dev_ : 0;
    if (!configured[dev_].load()) {
        PIO_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel<BKV, false, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PIO_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel<BKV, true, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev_].store(true);
    }
    if (p.drop.threshold != 0) PIO_CUDA(launch_pdl(attn_fwd_tc_kernel<BKV, true, DH>, grid, dim3(FA_THREADS), smem, st, tq, tk, tv, p));
    else PIO_CUDA(launch_pdl(attn_fwd_tc_kernel<BKV, false, DH>, grid, dim3(FA_THREADS), smem, st, tq, tk, tv, p));
    return PIO_OK;
}

template <int DH>
static int launch_bwd_dq(dim3 grid, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tdo, const CUtensorMap& tk,
                         const CUtensorMap& tv, const FaBwdParams& p) {
    constexpr int TC = HeadGeom<DH>::TILE_COLS;
    const size_t smem = 2 * (size_t)FA_BQ * TC * 2 + 6 * (size_t)64 * TC * 2 + 2 * 64 * 4 + 128 + 1024;
    static std::atomic<bool> configured[64];
    int dev_ = 0;
    PIO_CUDA(cudaGetDevice(&dev_));  // the attribute is per function AND per device
    dev_ = (dev_ >= 0 && dev_ < 64) ? dev_ : 0;
    if (!configured[dev_].load()) {
        PIO_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<false, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PIO_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<true, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev_].store(true);
    }
    if (p.drop.threshold != 0) PIO_CUDA(launch_pdl(attn_bwd_dq_tc_kernel<true, DH>, grid, dim3(FA_THREADS), smem, st, tq, tdo, tk, tv, p));
    else PIO_CUDA(launch_pdl(attn_bwd_dq_tc_kernel<false, DH>, grid, dim3(FA_THREADS), smem, st, tq, tdo, tk, tv, p));
    return PIO_OK;
}

template <int DH>
static int launch_bwd_dkv(dim3 grid, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tdo, const CUtensorMap& tk,
                          const CUtensorMap& tv, const FaBwdParams& p) {
    constexpr int TC = HeadGeom<DH>::TILE_COLS;
    const size_t smem = 2 * (size_t)128 * TC * 2 + 4 * (size_t)64 * TC * 2 + 4 * 64 * 4 + 128 + 1024;
    static std::atomic<bool> configured[64];
    int dev_ = 0;
    PIO_CUDA(cudaGetDevice(&dev_));  // the attribute is per function AND per device
    dev_ = (dev_ >= 0 && dev_ < 64) ? dev_ : 0;
    if (!configured[dev_].load()) {
        PIO_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<false, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PIO_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<true, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev_].store(true);
    }
    if (p.drop.threshold != 0) PIO_CUDA(launch_pdl(attn_bwd_dkv_tc_kernel<true, DH>, grid, dim3(FA_THREADS), smem, st, tq, tdo, tk, tv, p));
    else PIO_CUDA(launch_pdl(attn_bwd_dkv_tc_kernel<false, DH>, grid, dim3(FA_THREADS), smem, st, tq, tdo, tk, tv, p));
    return PIO_OK;
}

// q, k, v of the descriptor are bf16 (element strides as given); writes o (fp32, optional via d->o),
// o16 (optional) and lse.
int attn_fwd_tc(const pio_attn_desc* d, __nv_bfloat16* o16, cudaStream_t st, const DropCfg* drop) {
    PIO_CHECK_ARG(d && d->q && d->k && d->v && d->lse && (d->o || o16), "attn_fwd_tc: null pointer");
    PIO_CHECK_ARG(attn_fwd_tc_supported(d), "attn_fwd_tc: unsupported shape (head dim %d)", d->dh);
    PIO_CHECK_ARG(d->heads <= 65535 && d->batch <= 65535, "attn_fwd_tc: batch/heads exceed grid limits");
    constexpr int BKV = 64;
    const int64_t cols = (int64_t)d->heads * d->dh;
    CUtensorMap tq, tk, tv;
    PIO_TRY(make_tmap3(&tq, d->q, cols, d->n, d->batch, d->q_rs, d->q_bs, FA_BQ));
    PIO_TRY(make_tmap3(&tk, d->k, cols, d->m, d->batch, d->k_rs, d->k_bs, BKV));
    PIO_TRY(make_tmap3(&tv, d->v, cols, d->m, d->batch, d->v_rs, d->v_bs, BKV));
    FaParams p{};
    p.pad = d->pad;
    p.o32 = d->o;
    p.o16 = o16;
    p.lse = d->lse;
    p.o_bs = d->o_bs; p.o_rs = d->o_rs;
    p.batch = d->batch; p.heads = d->heads; p.n = d->n; p.m = d->m;
    p.scale_log2e = d->scale * 1.44269504088896340736f;
    p.drop = drop ? *drop : DropCfg{0, 0, 1.f};
    dim3 grid((unsigned)cdiv(d->n, FA_BQ), (unsigned)d->heads, (unsigned)d->batch);
    if (d->dh == 32) PIO_TRY((launch_fwd<32>(grid, st, tq, tk, tv, p)));
    else if (d->dh == 64) PIO_TRY((launch_fwd<64>(grid, st, tq, tk, tv, p)));
    else PIO_TRY((launch_fwd<128>(grid, st, tq, tk, tv, p)));
    count_launch();
    return PIO_OK;
}

// q, k, v, dO bf16; o fp32; lse fp32 -> dq16, dk16, dv16 (bf16, strides of q / k / v); delta scratch (batch, heads, n)
int attn_bwd_tc(const pio_attn_desc* d, const __nv_bfloat16* do16, float* delta, __nv_bfloat16* dq16, __nv_bfloat16* dk16,
                __nv_bfloat16* dv16, cudaStream_t st, const DropCfg* drop, const __nv_bfloat16* o16, bool dkv_add) {
    PIO_CHECK_ARG(d && d->q && d->k && d->v && (d->o || o16) && d->lse && do16 && delta && dq16 && dk16 && dv16,
                  "attn_bwd_tc: null pointer");
    PIO_CHECK_ARG(attn_fwd_tc_supported(d), "attn_bwd_tc: unsupported shape (head dim %d)", d->dh);
    PIO_CHECK_ARG(d->heads <= 65535 && d->batch <= 65535, "attn_bwd_tc: batch/heads exceed grid limits");
    if (xattn_bwd_tc_supported(d, drop)) return xattn_bwd_tc(d, do16, dq16, dk16, dv16, st, o16, dkv_add);  // (delta stays unused)
    const int64_t cols = (int64_t)d->heads * d->dh;
    FaBwdParams p{};
    p.pad = d->pad; p.o32 = d->o; p.o16 = o16; p.do16 = do16; p.lse = d->lse; p.delta = delta;
    p.dq16 = dq16; p.dk16 = dk16; p.dv16 = dv16;
    p.o_bs = d->o_bs; p.o_rs = d->o_rs; p.q_bs = d->q_bs; p.q_rs = d->q_rs;
    p.k_bs = d->k_bs; p.k_rs = d->k_rs; p.v_bs = d->v_bs; p.v_rs = d->v_rs;
    p.batch = d->batch; p.heads = d->heads; p.n = d->n; p.m = d->m;
    p.scale = d->scale;
    p.scale_log2e = d->scale * 1.44269504088896340736f;
    p.drop = drop ? *drop : DropCfg{0, 0, 1.f};
    p.dkv_add = dkv_add ? 1 : 0;
    {   // dQ + delta: 128-row query tiles, 64-row key tiles
        CUtensorMap tq, tdo, tk, tv;
        PIO_TRY(make_tmap3(&tq, d->q, cols, d->n, d->batch, d->q_rs, d->q_bs, FA_BQ));
        PIO_TRY(make_tmap3(&tdo, do16, cols, d->n, d->batch, d->o_rs, d->o_bs, FA_BQ));
        PIO_TRY(make_tmap3(&tk, d->k, cols, d->m, d->batch, d->k_rs, d->k_bs, 64));
        PIO_TRY(make_tmap3(&tv, d->v, cols, d->m, d->batch, d->v_rs, d->v_bs, 64));
        dim3 grid((unsigned)cdiv(d->n, FA_BQ), (unsigned)d->heads, (unsigned)d->batch);
        if (d->dh == 32) PIO_TRY((launch_bwd_dq<32>(grid, st, tq, tdo, tk, tv, p)));
        else if (d->dh == 64) PIO_TRY((launch_bwd_dq<64>(grid, st, tq, tdo, tk, tv, p)));
        else PIO_TRY((launch_bwd_dq<128>(grid, st, tq, tdo, tk, tv, p)));
        count_launch();
    }
    {   // dK, dV: 128-row key tiles, 64-row query tiles
        CUtensorMap tq, tdo, tk, tv;
        PIO_TRY(make_tmap3(&tq, d->q, cols, d->n, d->batch, d->q_rs, d->q_bs, 64));
        PIO_TRY(make_tmap3(&tdo, do16, cols, d->n, d->batch, d->o_rs, d->o_bs, 64));
        PIO_TRY(make_tmap3(&tk, d->k, cols, d->m, d->batch, d->k_rs, d->k_bs, 128));
        PIO_TRY(make_tmap3(&tv, d->v, cols, d->m, d->batch, d->v_rs, d->v_bs, 128));
        dim3 grid((unsigned)cdiv(d->m, 128), (unsigned)d->heads, (unsigned)d->batch);
        if (d->dh == 32) PIO_TRY((launch_bwd_dkv<32>(grid, st, tq, tdo, tk, tv, p)));
        else if (d->dh == 64) PIO_TRY((launch_bwd_dkv<64>(grid, st, tq, tdo, tk, tv, p)));
        else PIO_TRY((launch_bwd_dkv<128>(grid, st, tq, tdo, tk, tv, p)));
        count_launch();
    }
    return PIO_OK;
}

}  // namespace pio

extern "C" int pio_attn_bwd_bf16(const pio_attn_desc* fwd, const void* d_o16, void* dq16, void* dk16, void* dv16, float* delta,
                                 pio_stream_t stream) {
    PIO_CHECK_ARG(fwd != nullptr, "pio_attn_bwd_bf16: null descriptor");
    if (!pio::attn_fwd_tc_supported(fwd)) {
        pio::set_error("pio_attn_bwd_bf16: head dim %d has no tensor-core backward (use pio_attn_bwd)", fwd->dh);
        return PIO_ERR_UNSUPPORTED;
    }
    return pio::attn_bwd_tc(fwd, (const __nv_bfloat16*)d_o16, delta, (__nv_bfloat16*)dq16, (__nv_bfloat16*)dk16,
                            (__nv_bfloat16*)dv16, (cudaStream_t)stream, nullptr, nullptr, false);
}
