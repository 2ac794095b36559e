// Tensor-core attention backward for the decoder on the labelled rows (pio_decode_rows_bwd), head size 64, up to 128
// keys per sample.
//
// Reference: PerceiverDecoder.cross_attention (perceiver/model.py:203-206, :229-237) — every decoder query attends to
// the n latents of its own sample; the rows here are the COMPACTED labelled (sample, position) pairs, sample by sample
// (seg[b] .. seg[b + 1]).  The SIMT kernel this replaces (dr_attn_bwd_kernel, warp = row, dK / dV in registers of
// key-owning warps) took 63 us at the benchmark shape for 0.16 GFLOP.  Here a CTA owns one (sample, head):
//   * K_h, V_h (n keys x 64, zero rows up to 128 by TMA out-of-bounds fill) stay in shared memory; the sample's rows
//     arrive 128 at a time (Q_h and dO_h tiles by TMA; rows past the sample's last belong to the next sample and
//     are masked);
//   * per row tile: S = Q K^T and dP = dO V^T on tcgen05 (N = 128), then thread = row: p = exp(s / 8 - lse),
//     delta = sum_k p dP (the form consistent with the recomputed probabilities, as in attn_bwd_dq_tc_kernel for one
//     key tile), dS = p (dP - delta) / 8, both written as bf16 tiles that serve dQ = dS K as the K-major A operand
//     and dK += dS^T Q, dV += P^T dO as MN-major A operands (accumulating in tensor memory over the row tiles);
//   * dQ rows leave from registers (only the sample's own rows are written), dK / dV after the last tile.
#include <cuda.h>

#include <atomic>

#include "pio_latent_common.cuh"

namespace pio {

int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

namespace drt {

using lb::bf16;

constexpr int DH = 64;
constexpr int ROWS = 128;
constexpr int ATOM = ROWS * 128;
constexpr int EW_WARPS = 8;
constexpr int THREADS = EW_WARPS * 32 + 64;
constexpr size_t SMEM_BYTES = 8 * ATOM + 2 * ROWS * sizeof(float) + 16 * sizeof(uint64_t) + 1024;

struct Params {
    const float* lse;    // (rows, heads), natural log
    const int* seg;      // (batch + 1): first compact row of every sample
    const int* count;    // live rows
    bf16* dq16;          // (max_rows, e)
    bf16* dkv;           // (batch, n, 2 e)
    int n, e, heads;
    int64_t max_rows;
    float scale, scale_log2e, log2e;
};

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

__global__ void __launch_bounds__(THREADS, 1)
    dr_attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                          const __grid_constant__ CUtensorMap tmKV, const Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sDO = sQ + ATOM;
    uint8_t* sK = sDO + ATOM;
    uint8_t* sV = sK + ATOM;
    uint8_t* sP = sV + ATOM;         // [128 rows][128 keys] bf16 as two 64-key atoms
    uint8_t* sDS = sP + 2 * ATOM;
    float* sDelta = (float*)(sDS + 2 * ATOM);  // [2 column halves][128 rows]
    uint64_t* bars = (uint64_t*)(sDelta + 2 * ROWS);
    uint64_t* kv_full = bars;
    uint64_t* qdo_full = bars + 1;
    uint64_t* qdo_free = bars + 2;
    uint64_t* mma_done = bars + 3;
    uint64_t* ew_done = bars + 4;
    uint32_t* tmem_slot = (uint32_t*)(bars + 5);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, b = blockIdx.y;
    pdl_wait();
    const int cnt = __ldg(p.count);
    if (h == 0 && b == 0) {  // rows between the count and the next multiple of 128: zero (operand padding of dWq = dQ^T H)
        const int64_t upto = min(p.max_rows, ((int64_t)cnt + 127) / 128 * 128);
        for (int64_t t = (int64_t)cnt * p.e + threadIdx.x; t < upto * p.e; t += blockDim.x) p.dq16[t] = __float2bfloat16(0.f);
    }
    const int lo = __ldg(p.seg + b), hi = __ldg(p.seg + b + 1);
    const int tiles = (hi - lo + ROWS - 1) / ROWS;
    const int n = p.n;
    if (tiles <= 0) {  // no labelled position in this sample: its K_h, V_h get no gradient
        for (int t = threadIdx.x; t < n * (DH / 8); t += blockDim.x) {
            const int key = t / (DH / 8), pc = t % (DH / 8);
            bf16* row = p.dkv + ((int64_t)b * n + key) * (2 * p.e) + h * DH + pc * 8;
            *reinterpret_cast<uint4*>(row) = make_uint4(0u, 0u, 0u, 0u);
            *reinterpret_cast<uint4*>(row + p.e) = make_uint4(0u, 0u, 0u, 0u);
        }
        return;
    }

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmDO);
        prefetch_tmap(&tmKV);
        mbar_init(kv_full, 1);
        mbar_init(qdo_full, 1);
        mbar_init(qdo_free, 1);
        mbar_init(mma_done, 1);
        mbar_init(ew_done, EW_WARPS);
        fence_barrier_init();
    }
    if (warp == EW_WARPS + 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t T_S = tmem_base;         // S, 128 keys
    const uint32_t T_P = tmem_base + 128;   // dP
    const uint32_t T_DQ = tmem_base + 256;  // dQ | dK | dV: 64 columns each
    const uint32_t T_DK = tmem_base + 320;
    const uint32_t T_DV = tmem_base + 384;

    if (warp == EW_WARPS) {
        if (lane == 0) {
            mbar_expect_tx(kv_full, 2 * ATOM);
            tma_load_3d(sK, &tmKV, h * DH, 0, b, kv_full);
            tma_load_3d(sV, &tmKV, p.e + h * DH, 0, b, kv_full);
            for (int t = 0; t < tiles; ++t) {
                if (t > 0) mbar_wait(qdo_free, (uint32_t)(t - 1) & 1u);
                mbar_expect_tx(qdo_full, 2 * ATOM);
                tma_load_2d(sQ, &tmQ, h * DH, lo + t * ROWS, qdo_full);
                tma_load_2d(sDO, &tmDO, h * DH, lo + t * ROWS, qdo_full);
            }
        }
        __syncwarp();
    } else if (warp == EW_WARPS + 1) {
        if (lane == 0) {
            const uint32_t idesc_s = make_idesc(128, false, false);
            const uint32_t idesc_dq = make_idesc(DH, false, true);
            const uint32_t idesc_dkv = make_idesc(DH, true, true);
            const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sDO), k_addr = smem_u32(sK), v_addr = smem_u32(sV);
            const uint32_t p_addr = smem_u32(sP), ds_addr = smem_u32(sDS);
            uint32_t ew_phase = 0;
            mbar_wait(kv_full, 0);
            for (int t = 0; t < tiles; ++t) {
                mbar_wait(qdo_full, (uint32_t)t & 1u);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k)  // S = Q K^T
                    umma_bf16(T_S, make_smem_desc(q_addr + k * 32, 16, 1024), make_smem_desc(k_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < 4; ++k)  // dP = dO V^T
                    umma_bf16(T_P, make_smem_desc(do_addr + k * 32, 16, 1024), make_smem_desc(v_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
                umma_commit(mma_done);
                mbar_wait(ew_done, ew_phase);  // P, dS tiles written
                ew_phase ^= 1u;
                tc_fence_after();
#pragma unroll
                for (int s = 0; s < 8; ++s)  // dQ = dS K: 16 keys per step
                    umma_bf16(T_DQ, make_smem_desc(ds_addr + (s >> 2) * ATOM + (s & 3) * 32, 16, 1024), make_smem_desc(k_addr + s * 2048, 8192, 1024),
                              idesc_dq, s > 0 ? 1u : 0u);
#pragma unroll
                for (int s = 0; s < 8; ++s)  // dK += dS^T Q: 16 rows per step, A read MN-major (two 64-key atoms 16 KB apart)
                    umma_bf16(T_DK, make_smem_desc(ds_addr + s * 2048, ATOM, 1024), make_smem_desc(q_addr + s * 2048, 8192, 1024), idesc_dkv,
                              (t > 0 || s > 0) ? 1u : 0u);
#pragma unroll
                for (int s = 0; s < 8; ++s)  // dV += P^T dO
                    umma_bf16(T_DV, make_smem_desc(p_addr + s * 2048, ATOM, 1024), make_smem_desc(do_addr + s * 2048, 8192, 1024), idesc_dkv,
                              (t > 0 || s > 0) ? 1u : 0u);
                umma_commit(mma_done);
                umma_commit(qdo_free);
                mbar_wait(ew_done, ew_phase);  // dQ has been drained (the next tile's dQ product overwrites it)
                ew_phase ^= 1u;
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3, cg = warp >> 2;
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        uint32_t mma_phase = 0;
        auto wait_mma = [&]() {
            mbar_wait(mma_done, mma_phase);
            mma_phase ^= 1u;
            tc_fence_after();
        };
        auto signal = [&]() {
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) lb::mbar_arrive(ew_done);
        };
        for (int t = 0; t < tiles; ++t) {
            const int gi = lo + t * ROWS + row;
            const bool row_ok = gi < hi;
            const float lse2 = row_ok ? __ldg(p.lse + (int64_t)gi * p.heads + h) * p.log2e : 0.f;
            wait_mma();
            {   // 32-key chunks cg and cg + 2; keys >= n and rows of other samples contribute nothing
                float part = 0.f;
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int key0 = (cg + 2 * cc) * 32;
                    if (key0 < n) {  // (block-uniform)
                        float s[32], dp[32];
                        lb::ld_chunk(T_S + lane_off + (uint32_t)key0, s);
                        lb::ld_chunk(T_P + lane_off + (uint32_t)key0, dp);
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (key0 + i < n) part = fmaf(lb::ex2(fmaf(s[i], p.scale_log2e, -lse2)), dp[i], part);
                    }
                }
                sDelta[cg * ROWS + row] = part;
                asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
                const float delta = sDelta[row] + sDelta[ROWS + row];
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int chunk = cg + 2 * cc;
                    const int key0 = chunk * 32;
                    float s[32], dp[32];
                    if (key0 < n) {
                        lb::ld_chunk(T_S + lane_off + (uint32_t)key0, s);
                        lb::ld_chunk(T_P + lane_off + (uint32_t)key0, dp);
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const bool ok = row_ok && key0 + i < n;
                            s[i] = ok ? lb::ex2(fmaf(s[i], p.scale_log2e, -lse2)) : 0.f;
                            dp[i] = ok ? s[i] * (dp[i] - delta) * p.scale : 0.f;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            s[i] = 0.f;
                            dp[i] = 0.f;
                        }
                    }
                    lb::store_half_row_smem(sP + (chunk >> 1) * ATOM, row, chunk & 1, s);
                    lb::store_half_row_smem(sDS + (chunk >> 1) * ATOM, row, chunk & 1, dp);
                }
                asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");  // (sDelta is rewritten by the next tile)
                if (!row_ok) {
                    // rows past the sample's last: P and dS are zero there, but the Q / dO rows the tile loaded belong to
                    // the next sample or lie beyond the live rows (anything, NaN included), and 0 x NaN in dK += dS^T Q,
                    // dV += P^T dO is NaN: zero them (a 128-byte row whatever the swizzle) before those products run
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        *reinterpret_cast<uint4*>(sQ + row * 128 + cg * 64 + i * 16) = make_uint4(0u, 0u, 0u, 0u);
                        *reinterpret_cast<uint4*>(sDO + row * 128 + cg * 64 + i * 16) = make_uint4(0u, 0u, 0u, 0u);
                    }
                }
            }
            signal();
            wait_mma();
            {   // dQ rows of this sample -> global
                float v[32];
                lb::ld_chunk(T_DQ + lane_off + (uint32_t)(cg * 32), v);
                if (row_ok) lb::store32_bf16_global(p.dq16 + (int64_t)gi * p.e + h * DH + cg * 32, v);
            }
            if (t + 1 == tiles) {  // dK_h, dV_h: rows = keys
                float v[32];
                bf16* out = p.dkv + ((int64_t)b * n + row) * (2 * p.e) + h * DH + cg * 32;
                lb::ld_chunk(T_DK + lane_off + (uint32_t)(cg * 32), v);
                if (row < n) lb::store32_bf16_global(out, v);
                lb::ld_chunk(T_DV + lane_off + (uint32_t)(cg * 32), v);
                if (row < n) lb::store32_bf16_global(out + p.e, v);
            }
            signal();
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == EW_WARPS + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// Forward: o16[i, h] = softmax(q_i K_h^T / 8) V_h and lse[i, h] for the labelled rows, same CTA layout: S = Q K^T on tcgen05,
// thread = row softmax (the two threads of a row split the keys and exchange maximum and sum through shared memory),
// P as a bf16 K-major tile, O = P V with V read MN-major in place.
struct FwdParams {
    float* lse;
    const int* seg;
    const int* count;
    bf16* o16;  // (max_rows, e)
    int n, e, heads;
    int64_t max_rows;
    float scale_log2e;
};

__global__ void __launch_bounds__(THREADS, 1)
    dr_attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const FwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + ATOM;
    uint8_t* sV = sK + ATOM;
    uint8_t* sP = sV + ATOM;  // [128 rows][128 keys] bf16 as two 64-key atoms
    float* sRed = (float*)(sP + 2 * ATOM);  // [2 column halves][128 rows]
    uint64_t* bars = (uint64_t*)(sRed + 2 * ROWS);
    uint64_t* kv_full = bars;
    uint64_t* q_full = bars + 1;
    uint64_t* q_free = bars + 2;
    uint64_t* mma_done = bars + 3;
    uint64_t* ew_done = bars + 4;
    uint32_t* tmem_slot = (uint32_t*)(bars + 5);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, b = blockIdx.y;
    pdl_wait();
    const int cnt = __ldg(p.count);
    if (h == 0 && b == 0) {  // rows between the count and the next multiple of 128: zero (operand padding of dWo = dY^T O)
        const int64_t upto = min(p.max_rows, ((int64_t)cnt + 127) / 128 * 128);
        for (int64_t t = (int64_t)cnt * p.e + threadIdx.x; t < upto * p.e; t += blockDim.x) p.o16[t] = __float2bfloat16(0.f);
    }
    const int lo = __ldg(p.seg + b), hi = __ldg(p.seg + b + 1);
    const int tiles = (hi - lo + ROWS - 1) / ROWS;
    const int n = p.n;
    if (tiles <= 0) return;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmKV);
        mbar_init(kv_full, 1);
        mbar_init(q_full, 1);
        mbar_init(q_free, 1);
        mbar_init(mma_done, 1);
        mbar_init(ew_done, EW_WARPS);
        fence_barrier_init();
    }
    if (warp == EW_WARPS + 1) tmem_alloc(tmem_slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t T_S = tmem_base;        // S, 128 keys
    const uint32_t T_O = tmem_base + 128;  // O, 64 columns

    if (warp == EW_WARPS) {
        if (lane == 0) {
            mbar_expect_tx(kv_full, 2 * ATOM);
            tma_load_3d(sK, &tmKV, h * DH, 0, b, kv_full);
            tma_load_3d(sV, &tmKV, p.e + h * DH, 0, b, kv_full);
            for (int t = 0; t < tiles; ++t) {
                if (t > 0) mbar_wait(q_free, (uint32_t)(t - 1) & 1u);
                mbar_expect_tx(q_full, ATOM);
                tma_load_2d(sQ, &tmQ, h * DH, lo + t * ROWS, q_full);
            }
        }
        __syncwarp();
    } else if (warp == EW_WARPS + 1) {
        if (lane == 0) {
            const uint32_t idesc_s = make_idesc(128, false, false);
            const uint32_t idesc_pv = make_idesc(DH, false, true);
            const uint32_t q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV), p_addr = smem_u32(sP);
            uint32_t ew_phase = 0;
            mbar_wait(kv_full, 0);
            for (int t = 0; t < tiles; ++t) {
                mbar_wait(q_full, (uint32_t)t & 1u);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(T_S, make_smem_desc(q_addr + k * 32, 16, 1024), make_smem_desc(k_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
                umma_commit(mma_done);
                umma_commit(q_free);
                mbar_wait(ew_done, ew_phase);  // P written
                ew_phase ^= 1u;
                tc_fence_after();
#pragma unroll
                for (int s = 0; s < 8; ++s)  // O = P V: 16 keys per step
                    umma_bf16(T_O, make_smem_desc(p_addr + (s >> 2) * ATOM + (s & 3) * 32, 16, 1024), make_smem_desc(v_addr + s * 2048, 8192, 1024),
                              idesc_pv, s > 0 ? 1u : 0u);
                umma_commit(mma_done);
                mbar_wait(ew_done, ew_phase);  // O drained (the next tile's products overwrite S and O)
                ew_phase ^= 1u;
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3, cg = warp >> 2;
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        uint32_t mma_phase = 0;
        auto wait_mma = [&]() {
            mbar_wait(mma_done, mma_phase);
            mma_phase ^= 1u;
            tc_fence_after();
        };
        auto signal = [&]() {
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) lb::mbar_arrive(ew_done);
        };
        for (int t = 0; t < tiles; ++t) {
            const int gi = lo + t * ROWS + row;
            const bool row_ok = gi < hi;
            wait_mma();
            float inv_l;
            {
                float mx = -INFINITY;
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int key0 = (cg + 2 * cc) * 32;
                    if (key0 < n) {
                        float s[32];
                        lb::ld_chunk(T_S + lane_off + (uint32_t)key0, s);
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (key0 + i < n) mx = fmaxf(mx, s[i]);
                    }
                }
                sRed[cg * ROWS + row] = mx;
                asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
                mx = fmaxf(sRed[row], sRed[ROWS + row]);
                asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
                const float m2 = mx * p.scale_log2e;
                float lsum = 0.f;
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int chunk = cg + 2 * cc;
                    const int key0 = chunk * 32;
                    float s[32];
                    if (key0 < n) {
                        lb::ld_chunk(T_S + lane_off + (uint32_t)key0, s);
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            s[i] = (key0 + i < n) ? lb::ex2(fmaf(s[i], p.scale_log2e, -m2)) : 0.f;
                            lsum += s[i];
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) s[i] = 0.f;
                    }
                    lb::store_half_row_smem(sP + (chunk >> 1) * ATOM, row, chunk & 1, s);
                }
                sRed[cg * ROWS + row] = lsum;
                asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
                lsum = sRed[row] + sRed[ROWS + row];
                asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
                inv_l = 1.0f / lsum;
                if (cg == 0 && row_ok) p.lse[(int64_t)gi * p.heads + h] = (m2 + log2f(lsum)) * 0.69314718055994530942f;
            }
            signal();
            wait_mma();
            {
                float v[32];
                lb::ld_chunk(T_O + lane_off + (uint32_t)(cg * 32), v);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] *= inv_l;
                if (row_ok) lb::store32_bf16_global(p.o16 + (int64_t)gi * p.e + h * DH + cg * 32, v);
            }
            signal();
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == EW_WARPS + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 256);
    }
}

}  // namespace drt

bool dr_attn_bwd_tc_supported(int n, int e, int heads) {
    return getenv("PIO_NO_DR_ATTN_TC") == nullptr && heads > 0 && e == heads * drt::DH && n >= 1 && n <= 128 && e % 8 == 0;
}

// q16, do16 (max_rows, e) bf16; kv (batch, n, 2 e) bf16; lse (max_rows, heads); seg (batch + 1); count: live rows (device);
// dq16 (max_rows, e) and dkv (batch, n, 2 e) are written
int dr_attn_bwd_tc(const void* q16, const void* kv, const void* do16, const float* lse, const int* seg, const int* count, void* dq16,
                   void* dkv, int batch, int n, int e, int heads, int64_t max_rows, cudaStream_t st) {
    using namespace drt;
    PIO_CHECK_ARG(q16 && kv && do16 && lse && seg && count && dq16 && dkv && dr_attn_bwd_tc_supported(n, e, heads) && batch > 0 &&
                      batch <= 65535 && heads <= 65535,
                  "dr_attn_bwd_tc: unsupported arguments");
    CUtensorMap tq, tdo, tkv;
    PIO_TRY(make_tmap_bf16(&tq, q16, max_rows, e, e, ROWS));
    PIO_TRY(make_tmap_bf16(&tdo, do16, max_rows, e, e, ROWS));
    PIO_TRY(make_out_map(&tkv, false, kv, 2 * (int64_t)e, n, batch, 2 * (int64_t)e * 2, (int64_t)n * 2 * e * 2, 64, ROWS));
    Params p{};
    p.lse = lse; p.seg = seg; p.count = count; p.dq16 = (bf16*)dq16; p.dkv = (bf16*)dkv;
    p.n = n; p.e = e; p.heads = heads; p.max_rows = max_rows;
    p.scale = 1.0f / sqrtf((float)DH);
    p.log2e = 1.44269504088896340736f;
    p.scale_log2e = p.scale * p.log2e;
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(dr_attn_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    PIO_CUDA(launch_pdl(dr_attn_bwd_tc_kernel, dim3((unsigned)heads, (unsigned)batch), dim3(THREADS), SMEM_BYTES, st, tq, tdo, tkv, p));
    count_launch();
    return PIO_OK;
}

int dr_attn_fwd_tc(const void* q16, const void* kv, const int* seg, const int* count, void* o16, float* lse, int batch, int n, int e,
                   int heads, int64_t max_rows, cudaStream_t st) {
    using namespace drt;
    PIO_CHECK_ARG(q16 && kv && seg && count && o16 && lse && dr_attn_bwd_tc_supported(n, e, heads) && batch > 0 && batch <= 65535 &&
                      heads <= 65535,
                  "dr_attn_fwd_tc: unsupported arguments");
    CUtensorMap tq, tkv;
    PIO_TRY(make_tmap_bf16(&tq, q16, max_rows, e, e, ROWS));
    PIO_TRY(make_out_map(&tkv, false, kv, 2 * (int64_t)e, n, batch, 2 * (int64_t)e * 2, (int64_t)n * 2 * e * 2, 64, ROWS));
    FwdParams p{};
    p.lse = lse; p.seg = seg; p.count = count; p.o16 = (bf16*)o16;
    p.n = n; p.e = e; p.heads = heads; p.max_rows = max_rows;
    p.scale_log2e = (1.0f / sqrtf((float)DH)) * 1.44269504088896340736f;
    const size_t smem = 5 * ATOM + 2 * ROWS * sizeof(float) + 16 * sizeof(uint64_t) + 1024;
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(dr_attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev].store(true);
    }
    PIO_CUDA(launch_pdl(dr_attn_fwd_tc_kernel, dim3((unsigned)heads, (unsigned)batch), dim3(THREADS), smem, st, tq, tkv, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
