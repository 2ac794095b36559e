// Persistent fused latent self-attention block, forward (PIO_MODE_BF16; C = 256 channels, 4 heads of 64).
//
// Reference: perceiver/model.py:36-44 (self_attention_layer / self_attention_block), :102-116
// (SelfAttention), :20-26 (mlp), :47-56 (Residual).  One launch runs ALL layers of a block; one CTA owns
// 128 latent rows (128 / N whole samples) from the first LayerNorm of the first layer to the last
// residual add of the last layer:
//
//   * the fp32 residual stream of the 128 rows lives in 256 TMEM columns for the whole kernel; the
//     attention out-projection and the second MLP product ACCUMULATE INTO it on the tensor core
//     (D += A B with D = x), so both residual adds are free and the stream never leaves the SM between
//     layers except as the copies the backward pass needs;
//   * LayerNorm, bias adds, softmax and GELU are thread-local: thread = row (tcgen05.ld hands a thread
//     32 columns of its own row), two warps per TMEM lane quadrant split the columns;
//   * every operand of every product is staged on chip: LayerNorm / GELU / attention outputs are written
//     as bf16 straight into the 128-byte-swizzled K-major layout the MMA reads, Q_h / K_h / V_h of one
//     head likewise; the probabilities go to the P.V product through tensor memory;
//   * the layer's 768 KB of bf16 weights stream from L2 through a 3-stage, 32 KB-per-stage TMA ring in
//     exactly the order the products consume them (per head: 4 k-blocks of [Wq_h;Wk_h;Wv_h], then the
//     head's k-block of Wo; then W1 and W2 k-block by k-block); they cannot be resident (227 KB of
//     shared memory), but the ring runs ahead of the math, so the loads are off the critical path;
//   * self-attention of the 128 / N samples of a tile is one 128 x 128 score product per head whose
//     off-diagonal (cross-sample) blocks are never read; P is written with zeros there.
//
// Warp roles: warps 0..15 element-wise math (TMEM lane quadrant = warp & 3, column group = warp >> 2: a thread
// owns 64 of its row's 256 channels, i.e. its row of one 64-column operand atom), warp 16 = TMA producer,
// warp 17 = TMEM allocator + MMA issuer.  The element-wise warps and the MMA thread alternate strictly
// (mbarriers ew_done / mma_done).  Measured on the way (tools/latent_trace.py, cycles per layer and CTA): with
// 8 element-wise warps the element-wise phases took 42.7 K and the MMA side 30.1 K of 73 K cycles, the
// element-wise warps issuing at 0.4 IPC per scheduler in chains of tensor-memory loads, shared-memory
// stores and fences: hence four warps per scheduler, and rows kept in registers across the LayerNorm
// reduction instead of a second pass over tensor memory.
//
// What it writes for the backward pass is the layout of SaSaved / MlpSaved (pio_ops_bf16.cu), so the
// unfused backward operators run unchanged on its output.  Thread = row is the wrong shape for global
// stores (32 lanes = 32 different rows: every 16-byte store instruction touched 32 cache lines and the first
// version spent a third of its time in the LSU), so nothing but the per-row scalars is stored from
// registers: the bf16 tensors leave by TMA straight from the swizzled operand tiles the MMAs read (the MMA
// thread issues the store when the element-wise warps hand a tile over, and waits for the store's read of
// shared memory before it lets them overwrite it), the fp32 residual copies through the 4 KB of the
// LayerNorm-output atom that the warp itself is about to fill (32 rows x 128 bytes, swizzled: private to
// the warp, so only its own lane 0 has to wait for the store's read).  GELU' has no tile to leave from and
// is the one tensor still stored from registers.
#include <cuda.h>
#include <stdlib.h>

#include "pio_latent_common.cuh"

namespace pio {

int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

namespace lb {

constexpr int C = 256;
constexpr int DH = 64;
constexpr int H = 4;
constexpr int ROWS = 128;
constexpr int EW_WARPS = 16;
constexpr int EW_THREADS = EW_WARPS * 32;
constexpr int THREADS = EW_THREADS + 64;
constexpr int STAGES = 3;
constexpr int STAGE_BYTES = 32768;
constexpr int QKV_STAGE_BYTES = 3 * 64 * 128;  // 24 KB: 64 rows each of Wq_h, Wk_h, Wv_h x one 64-column k-block
constexpr int A_BYTES = 4 * ROWS * 128;        // 64 KB: 128 rows x 256 bf16 as four 64-column swizzle atoms
constexpr int HEAD_BYTES = ROWS * 128;         // 16 KB: 128 rows x 64 bf16
constexpr int TMEM_COLS = 512;
// per-layer parameter block in shared memory (floats)
constexpr int PAR_LN1_W = 0, PAR_LN1_B = 256, PAR_BQKV = 512, PAR_BO = 1280, PAR_LN2_W = 1536, PAR_LN2_B = 1792, PAR_B1 = 2048,
              PAR_B2 = 2304, PAR_FLOATS = 2560;

struct LayerPtrs {
    const float *ln1_w, *ln1_b, *bqkv, *bo, *ln2_w, *ln2_b, *b1, *b2;
    float* sa_stats;
    bf16* sa_h0;
    bf16* qkv;
    float* o32;
    bf16* o16;
    float* lse;
    float* xmid;
    float* mlp_stats;
    bf16* mlp_h0;
    bf16* mlp_h1;
    bf16* mlp_g;
    float* xout;
};

// output tensor maps: (columns, rows, layer) with the layer stride of the saved buffer
enum OutMap : int { OM_SA_H0 = 0, OM_QKV, OM_O16, OM_MLP_H0, OM_MLP_G, OM_MLP_H1, OM_XMID, OM_XNEXT, OM_Y, OM_COUNT };

struct Params {
    CUtensorMap maps[PIO_LATENT_FUSED_MAX_LAYERS][4];  // Wqkv (768 x 256), Wo, W1, W2 (256 x 256), bf16 row-major
    CUtensorMap out[OM_COUNT];
    LayerPtrs layer[PIO_LATENT_FUSED_MAX_LAYERS];
    const float* x;
    int layers;
    int n;  // latents per sample: 32, 64 or 128
    int64_t total_rows;
    float scale_log2e;
};

// Phase timeline for tuning (only in builds with -DPIO_LB_TRACE, tools/latent_trace.py): CTA 0's MMA thread (who 0)
// and lane 0 of element-wise warp 0 (who 1) log (tag, clock64) pairs.
#ifdef PIO_LB_TRACE
__device__ long long g_trace[2][2][8192];
__device__ int g_trace_n[2];
#define LB_TRACE(who, tag)                                              \
    do {                                                                \
        if (blockIdx.x == 0) {                                          \
            const int _i = g_trace_n[who];                              \
            if (_i < 8192) {                                            \
                g_trace[who][0][_i] = (tag);                            \
                g_trace[who][1][_i] = clock64();                        \
                g_trace_n[who] = _i + 1;                                \
            }                                                           \
        }                                                               \
    } while (0)
#else
#define LB_TRACE(who, tag) do { } while (0)
#endif

__device__ __forceinline__ void bar_ew() { asm volatile("bar.sync 1, 512;" ::: "memory"); }

__global__ void __launch_bounds__(THREADS, 1) latent_block_fwd_kernel(const __grid_constant__ Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                      // LayerNorm / GELU output: A operand of the QKV, W1, W2 products
    uint8_t* sQ = sA + A_BYTES;              // Q_h, later O_h (A operand of the out-projection)
    uint8_t* sK = sQ + HEAD_BYTES;
    uint8_t* sV = sK + HEAD_BYTES;
    uint8_t* ring = sV + HEAD_BYTES;         // [STAGES][32 KB] weight tiles
    float2* sRed = (float2*)(ring + STAGES * STAGE_BYTES);  // [4 column groups][128 rows]; reused only after an ew_done / mma_done round trip
    // the current layer's LayerNorm parameters and biases (read by every thread for every chunk: from L2 each read
    // cost a full memory round trip in a dependent chain, ncu: 40 % of the stall samples)
    float* sPar = (float*)(sRed + 4 * ROWS);
    uint64_t* bars = (uint64_t*)(sPar + PAR_FLOATS);
    uint64_t* full = bars;                   // [STAGES]
    uint64_t* empty = bars + STAGES;         // [STAGES]
    uint64_t* mma_done = bars + 2 * STAGES;
    uint64_t* ew_done = bars + 2 * STAGES + 1;
    uint64_t* sa_free = bars + 2 * STAGES + 2;  // the TMA stores out of sA issued by the MMA thread have read it
    uint32_t* tmem_slot = (uint32_t*)(bars + 2 * STAGES + 3);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(mma_done, 1);
        mbar_init(ew_done, EW_WARPS);
        mbar_init(sa_free, 1);
        fence_barrier_init();
    }
    if (warp == EW_WARPS + 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t T_RES = tmem_base;        // residual stream, 256 columns
    const uint32_t T_S1 = tmem_base + 256;   // scratch: [Q_h | K_h | V_h] (192) / S (128) + P (64 packed) + O_h (64) / MLP hidden (256)
    pdl_wait();  // everything below reads what earlier kernels wrote (activations, bf16 weight shadows)

    if (warp == EW_WARPS) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int prod = 0;
            auto acquire = [&](uint32_t bytes) -> uint8_t* {
                const int s = prod % STAGES;
                mbar_wait(&empty[s], ((uint32_t)(prod / STAGES) & 1u) ^ 1u);
                mbar_expect_tx(&full[s], bytes);
                return ring + s * STAGE_BYTES;
            };
            for (int l = 0; l < p.layers; ++l) {
                const CUtensorMap* mq = &p.maps[l][0];
                const CUtensorMap* mo = &p.maps[l][1];
                const CUtensorMap* m1 = &p.maps[l][2];
                const CUtensorMap* m2 = &p.maps[l][3];
                if (l == 0) {
                    prefetch_tmap(mq); prefetch_tmap(mo); prefetch_tmap(m1); prefetch_tmap(m2);
                }
                for (int h = 0; h < H; ++h) {
                    for (int kb = 0; kb < 4; ++kb) {
                        uint8_t* dst = acquire(QKV_STAGE_BYTES);
                        uint64_t* bar = &full[prod % STAGES];
#pragma unroll
                        for (int j = 0; j < 3; ++j) tma_load_2d(dst + j * 8192, mq, kb * 64, j * C + h * DH, bar);
                        ++prod;
                    }
                    uint8_t* dst = acquire(STAGE_BYTES);
                    tma_load_2d(dst, mo, h * DH, 0, &full[prod % STAGES]);
                    ++prod;
                }
                for (int kb = 0; kb < 4; ++kb) {
                    uint8_t* dst = acquire(STAGE_BYTES);
                    tma_load_2d(dst, m1, kb * 64, 0, &full[prod % STAGES]);
                    ++prod;
                }
                for (int kb = 0; kb < 4; ++kb) {
                    uint8_t* dst = acquire(STAGE_BYTES);
                    tma_load_2d(dst, m2, kb * 64, 0, &full[prod % STAGES]);
                    ++prod;
                }
            }
        }
        __syncwarp();
    } else if (warp == EW_WARPS + 1) {
        // ------------------------------------------------------------ MMA issuer (+ TMA stores of the operand tiles)
        if (lane == 0) {
            const uint32_t idesc_qkv = make_idesc(192, false, false);
            const uint32_t idesc_s = make_idesc(128, false, false);
            const uint32_t idesc_pv = make_idesc(DH, false, true);
            const uint32_t idesc_c = make_idesc(C, false, false);
            const uint32_t a_addr = smem_u32(sA), q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV);
            int cons = 0;
            uint32_t ew_phase = 0;
            int tr = 0;
            auto wait_ew = [&]() {
                LB_TRACE(0, 1000 + tr);
                mbar_wait(ew_done, ew_phase);
                ew_phase ^= 1u;
                tc_fence_after();
                LB_TRACE(0, 2000 + tr);
                ++tr;
            };
            auto ring_wait = [&]() -> uint32_t {
                const int s = cons % STAGES;
                mbar_wait(&full[s], (uint32_t)(cons / STAGES) & 1u);
                tc_fence_after();
                return smem_u32(ring + s * STAGE_BYTES);
            };
            auto ring_release = [&]() {
                umma_commit(&empty[cons % STAGES]);
                ++cons;
            };
            auto out_proj = [&]() {  // x += O_h Wo[:, head columns]^T
                const uint32_t b = ring_wait();
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(T_RES, make_smem_desc(q_addr + k * 32, 16, 1024), make_smem_desc(b + k * 32, 16, 1024), idesc_c, 1u);
                ring_release();
            };
            auto full_k = [&](uint32_t tmem_d, bool into_residual) {  // D (+)= sA W^T over the four k-blocks of the ring
                for (int kb = 0; kb < 4; ++kb) {
                    LB_TRACE(0, 3000 + kb);
                    const uint32_t b = ring_wait();
                    LB_TRACE(0, 3100 + kb);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem_d, make_smem_desc(a_addr + kb * (ROWS * 128) + k * 32, 16, 1024),
                                  make_smem_desc(b + k * 32, 16, 1024), idesc_c, (into_residual || kb > 0 || k > 0) ? 1u : 0u);
                    ring_release();
                }
            };
            const int row0 = blockIdx.x * ROWS;
            auto store_tile = [&](int which, const uint8_t* tile, int col0, int l) {  // one 128-row x 64-column bf16 atom
                tma_store_3d(&p.out[which], tile, col0, row0, l);
            };
            for (int l = 0; l < p.layers; ++l) {
                for (int h = 0; h < H; ++h) {
                    wait_ew();  // h = 0: LayerNorm output in sA; h > 0: O_{h-1} in sQ
                    if (h == 0) {
#pragma unroll
                        for (int a = 0; a < 4; ++a) store_tile(OM_SA_H0, sA + a * (ROWS * 128), a * 64, l);
                        tma_store_commit();
                    } else {
                        store_tile(OM_O16, sQ, (h - 1) * DH, l);
                        tma_store_commit();
                        out_proj();
                    }
                    for (int kb = 0; kb < 4; ++kb) {  // [Q_h | K_h | V_h] = LN(x) [Wq_h; Wk_h; Wv_h]^T
                        const uint32_t b = ring_wait();
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16(T_S1, make_smem_desc(a_addr + kb * (ROWS * 128) + k * 32, 16, 1024),
                                      make_smem_desc(b + k * 32, 16, 1024), idesc_qkv, (kb > 0 || k > 0) ? 1u : 0u);
                        ring_release();
                    }
                    if (h > 0) tma_store_wait_read();  // O_{h-1} has left sQ: the element-wise warps may write Q_h over it
                    umma_commit(mma_done);
                    wait_ew();  // Q_h, K_h, V_h as bf16 in shared memory
                    store_tile(OM_QKV, sQ, h * DH, l);
                    store_tile(OM_QKV, sK, C + h * DH, l);
                    store_tile(OM_QKV, sV, 2 * C + h * DH, l);
                    tma_store_commit();
#pragma unroll
                    for (int k = 0; k < 4; ++k)  // S = Q_h K_h^T for all 128 x 128 (row, key) pairs of the tile
                        umma_bf16(T_S1, make_smem_desc(q_addr + k * 32, 16, 1024), make_smem_desc(k_addr + k * 32, 16, 1024), idesc_s,
                                  k > 0 ? 1u : 0u);
                    umma_commit(mma_done);
                    wait_ew();  // P (bf16 pairs, zero outside the row's own sample) in tensor memory
#pragma unroll
                    for (int t = 0; t < 8; ++t)  // O_h = P V_h
                        umma_bf16_ts(T_S1 + 192, T_S1 + 128 + (uint32_t)t * 8, make_smem_desc(v_addr + t * 2048, ROWS * 128, 1024),
                                     idesc_pv, t > 0 ? 1u : 0u);
                    tma_store_wait_read();  // Q_h has left sQ (O_h goes there next); K_h, V_h and (h = 0) the LayerNorm copy likewise
                    umma_commit(mma_done);
                }
                wait_ew();
                store_tile(OM_O16, sQ, (H - 1) * DH, l);
                tma_store_commit();
                out_proj();
                umma_commit(mma_done);
                wait_ew();  // LayerNorm 2 output in sA
#pragma unroll
                for (int a = 0; a < 4; ++a) store_tile(OM_MLP_H0, sA + a * (ROWS * 128), a * 64, l);
                tma_store_commit();
                full_k(T_S1, false);
                umma_commit(mma_done);
                LB_TRACE(0, 3200);
                tma_store_wait_read();  // off the critical path: the element-wise warps first turn the hidden tile into GELU values
                LB_TRACE(0, 3201);
                mbar_arrive(sa_free);
                wait_ew();  // GELU output in sA
#pragma unroll
                for (int a = 0; a < 4; ++a) store_tile(OM_MLP_G, sA + a * (ROWS * 128), a * 64, l);
                tma_store_commit();
                full_k(T_RES, true);
                umma_commit(mma_done);
                tma_store_wait_read();
                mbar_arrive(sa_free);
            }
            tma_store_wait_all();
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------ element-wise warps
        const int q = warp & 3, cg = warp >> 2;   // TMEM lane quadrant, column group (= operand atom) of this warp
        const int row = q * 32 + lane;
        const int col_base = cg * 64;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const int64_t grow = (int64_t)blockIdx.x * ROWS + row;
        const bool row_ok = grow < p.total_rows;
        const int n = p.n;
        const int64_t sample = grow / n;
        const int qi = (int)(grow - sample * n);
        const int kbase = (row / n) * n;  // first key column of this row's sample (warp-uniform: n >= 32)
        uint32_t mma_phase = 0, sa_phase = 0;
        // this warp's rows of its own LayerNorm-output atom: 32 rows x 128 bytes = 4 KB nobody else touches; doubles
        // as the staging slice of the fp32 residual copies (x_mid, x_next) that leave by TMA
        uint8_t* atom = sA + cg * (ROWS * 128);
        uint8_t* slice = atom + q * 4096;
        const int slice_row0 = blockIdx.x * ROWS + q * 32;
        int tr = 0;
        auto wait_mma = [&]() {
            if (threadIdx.x == 0) LB_TRACE(1, 1000 + tr);
            mbar_wait(mma_done, mma_phase);
            mma_phase ^= 1u;
            tc_fence_after();
            if (threadIdx.x == 0) LB_TRACE(1, 2000 + tr);
            ++tr;
        };
        auto wait_sa_free = [&]() {
            mbar_wait(sa_free, sa_phase);
            sa_phase ^= 1u;
        };
        auto signal = [&]() {  // everything this warp wrote (shared memory, tensor memory) is visible to the MMA thread
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(ew_done);
        };
        // 32 fp32 per row -> the warp's slice -> TMA store; an earlier store out of the slice must have read it
        auto stage_store_f32 = [&](int which, int col0, int l, const float* v) {
            if (lane == 0) tma_store_wait_read();
            __syncwarp();
            uint8_t* prow = slice + lane * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(prow + ((j ^ (lane & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                tma_store_3d(&p.out[which], slice, col0, slice_row0, l);
                tma_store_commit();
            }
        };
        // v = this thread's 64 channels of its row, st their statistics: combine the four column groups' statistics,
        // write (mean, rstd), normalise in registers and write the bf16 row into the warp's atom of sA
        auto ln_finish = [&](float (&v)[64], const Stat& st, const float* gamma, const float* beta, float* stats_out) {
            float2* red = sRed;
            red[cg * ROWS + row] = make_float2(st.mean, st.m2);
            if (threadIdx.x == 0) LB_TRACE(1, 6000);
            bar_ew();
            if (threadIdx.x == 0) LB_TRACE(1, 6001);
            Stat t{0.f, 0.f, 0.f};
#pragma unroll
            for (int g = 0; g < 4; ++g) {  // same order in every thread: the four warps of a row agree bit for bit
                const float2 o = red[g * ROWS + row];
                if (g == 0) {
                    t.n = 64.f; t.mean = o.x; t.m2 = o.y;
                } else {
                    stat_merge(t, 64.f, o.x, o.y);
                }
            }
            const float mean = t.mean;
            const float rstd = rsqrtf(t.m2 * (1.0f / C) + LN_EPS);
            if (cg == 0 && row_ok) *reinterpret_cast<float2*>(stats_out + 2 * grow) = make_float2(mean, rstd);
#pragma unroll
            for (int d = 0; d < 64; d += 4) {
                const float4 g = *reinterpret_cast<const float4*>(gamma + col_base + d);
                const float4 b = *reinterpret_cast<const float4*>(beta + col_base + d);
                v[d] = fmaf((v[d] - mean) * rstd, g.x, b.x);
                v[d + 1] = fmaf((v[d + 1] - mean) * rstd, g.y, b.y);
                v[d + 2] = fmaf((v[d + 2] - mean) * rstd, g.z, b.z);
                v[d + 3] = fmaf((v[d + 3] - mean) * rstd, g.w, b.w);
            }
            if (threadIdx.x == 0) LB_TRACE(1, 6002);
            if (lane == 0) tma_store_wait_read();  // the residual copy staged in this slice has left it
            __syncwarp();
            if (threadIdx.x == 0) LB_TRACE(1, 6003);
            store_half_row_smem(atom, row, 0, *reinterpret_cast<float(*)[32]>(&v[0]));
            store_half_row_smem(atom, row, 1, *reinterpret_cast<float(*)[32]>(&v[32]));
        };

        // parameters of a layer: five coalesced loads per thread into registers (prefetch), later into sPar
        float pre[5];
        auto prefetch_params = [&](int l) {
            const LayerPtrs& Lp = p.layer[l];
            const float* seg[10] = {Lp.ln1_w, Lp.ln1_b, Lp.bqkv, Lp.bqkv + 256, Lp.bqkv + 512, Lp.bo, Lp.ln2_w, Lp.ln2_b, Lp.b1, Lp.b2};
#pragma unroll
            for (int k = 0; k < 5; ++k)  // element k * 512 + tid of the 2560: segment 2k or 2k + 1 (constant indices: no local array)
                pre[k] = __ldg(((threadIdx.x >> 8) ? seg[2 * k + 1] : seg[2 * k]) + (threadIdx.x & 255));
        };
        auto commit_params = [&]() {
#pragma unroll
            for (int k = 0; k < 5; ++k) sPar[k * EW_THREADS + threadIdx.x] = pre[k];
        };
        prefetch_params(0);

        // ---- layer 0 input: global -> tensor memory, statistics on the way
        float v[64];
        Stat st{0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int col0 = col_base + c * 32;
            float(&vc)[32] = *reinterpret_cast<float(*)[32]>(&v[c * 32]);
            if (row_ok) {
                const float4* src = reinterpret_cast<const float4*>(p.x + grow * C + col0);
#pragma unroll
                for (int d = 0; d < 8; ++d) {
                    const float4 t = src[d];
                    vc[4 * d] = t.x; vc[4 * d + 1] = t.y; vc[4 * d + 2] = t.z; vc[4 * d + 3] = t.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) vc[i] = 0.f;
            }
            st_chunk(T_RES + lane_off + (uint32_t)col0, vc);
            stat_chunk(st, vc);
        }
        tmem_st_wait();
        commit_params();  // (the barrier inside ln_finish orders these writes before anyone's reads)

        for (int l = 0; l < p.layers; ++l) {
            const LayerPtrs& L = p.layer[l];
            ln_finish(v, st, sPar + PAR_LN1_W, sPar + PAR_LN1_B, L.sa_stats);
            signal();

            for (int h = 0; h < H; ++h) {
                // ---- [Q_h | K_h | V_h] + bias -> bf16 operand tiles (column groups 0, 1, 2 take Q, K, V)
                wait_mma();
                if (cg < 3) {
                    uint8_t* tile = cg == 0 ? sQ : (cg == 1 ? sK : sV);
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        float t[32];
                        ld_chunk(T_S1 + lane_off + (uint32_t)(cg * 64 + c * 32), t);
                        add_vec32(t, sPar + PAR_BQKV + cg * C + h * DH + c * 32);
                        store_half_row_smem(tile, row, c, t);
                    }
                }
                signal();

                // ---- softmax over the row's own sample (column group 0 warps)
                wait_mma();
                float inv_l = 0.f;
                if (cg == 0) {
                    float mx = -INFINITY;
                    for (int c = 0; c < (n >> 5); ++c) {
                        float t[32];
                        ld_chunk(T_S1 + lane_off + (uint32_t)(kbase + c * 32), t);
#pragma unroll
                        for (int i = 0; i < 32; ++i) mx = fmaxf(mx, t[i]);
                    }
                    const float m2 = mx * p.scale_log2e;
                    float lsum = 0.f;
#pragma unroll 1
                    for (int g = 0; g < 2; ++g) {
                        uint32_t pw[32];
#pragma unroll
                        for (int c2 = 0; c2 < 2; ++c2) {
                            const int key0 = g * 64 + c2 * 32;
                            if (key0 >= kbase && key0 < kbase + n) {  // warp-uniform
                                float t[32];
                                ld_chunk(T_S1 + lane_off + (uint32_t)key0, t);
#pragma unroll
                                for (int i = 0; i < 32; ++i) {
                                    t[i] = ex2(fmaf(t[i], p.scale_log2e, -m2));
                                    lsum += t[i];
                                }
#pragma unroll
                                for (int i = 0; i < 16; ++i) {
                                    __nv_bfloat162 h2 = __floats2bfloat162_rn(t[2 * i], t[2 * i + 1]);
                                    pw[c2 * 16 + i] = *reinterpret_cast<uint32_t*>(&h2);
                                }
                            } else {
#pragma unroll
                                for (int i = 0; i < 16; ++i) pw[c2 * 16 + i] = 0u;
                            }
                        }
                        tmem_st32(T_S1 + 128 + (uint32_t)(g * 32) + lane_off, pw);
                    }
                    tmem_st_wait();
                    inv_l = 1.0f / lsum;
                    if (row_ok) L.lse[(sample * H + h) * n + qi] = (m2 + log2f(lsum)) * 0.69314718055994530942f;
                }
                signal();

                // ---- O_h / l -> bf16 operand of the out-projection (over Q_h)
                wait_mma();
                if (cg == 0) {
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        float t[32];
                        ld_chunk(T_S1 + 192 + lane_off + (uint32_t)(c * 32), t);
#pragma unroll
                        for (int i = 0; i < 32; ++i) t[i] *= inv_l;
                        store_half_row_smem(sQ, row, c, t);
                    }
                }
                signal();
            }

            // ---- x_mid = x + attention + bo (back into the residual columns, fp32 copy out), LayerNorm 2
            wait_mma();
            st = Stat{0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int col0 = col_base + c * 32;
                float(&vc)[32] = *reinterpret_cast<float(*)[32]>(&v[c * 32]);
                ld_chunk(T_RES + lane_off + (uint32_t)col0, vc);
                add_vec32(vc, sPar + PAR_BO + col0);
                st_chunk(T_RES + lane_off + (uint32_t)col0, vc);
                if (threadIdx.x == 0) LB_TRACE(1, 5000 + c);
                stage_store_f32(OM_XMID, col0, l, vc);
                if (threadIdx.x == 0) LB_TRACE(1, 5010 + c);
                stat_chunk(st, vc);
            }
            tmem_st_wait();
            if (threadIdx.x == 0) LB_TRACE(1, 5020);
            ln_finish(v, st, sPar + PAR_LN2_W, sPar + PAR_LN2_B, L.mlp_stats);
            if (threadIdx.x == 0) LB_TRACE(1, 5021);
            signal();
            if (threadIdx.x == 0) LB_TRACE(1, 5022);

            // ---- GELU(hidden + b1) -> sA (bf16; leaves by TMA), GELU' (bf16) for the backward pass from registers
            wait_mma();
            if (threadIdx.x == 0) LB_TRACE(1, 4000);
            wait_sa_free();  // the LayerNorm-2 copy has left sA
            if (threadIdx.x == 0) LB_TRACE(1, 4001);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int col0 = col_base + c * 32;
                float t[32], dg[32];
                if (threadIdx.x == 0) LB_TRACE(1, 4010 + c);
                ld_chunk(T_S1 + lane_off + (uint32_t)col0, t);
                add_vec32(t, sPar + PAR_B1 + col0);
#pragma unroll
                for (int i = 0; i < 32; ++i) gelu_erf_both(t[i], t[i], dg[i]);
                store_half_row_smem(atom, row, c, t);
                if (row_ok) store32_bf16_global(L.mlp_h1 + grow * C + col0, dg);
            }
            if (threadIdx.x == 0) LB_TRACE(1, 4020);
            signal();
            if (threadIdx.x == 0) LB_TRACE(1, 4021);

            // ---- x_next = x_mid + mlp + b2: fp32 copy out; stays in the residual columns (and registers) for the next layer
            wait_mma();
            wait_sa_free();  // the GELU copy has left sA: the slices are free for staging
            const bool last = (l + 1 == p.layers);
            if (!last) prefetch_params(l + 1);
            st = Stat{0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int col0 = col_base + c * 32;
                float(&vc)[32] = *reinterpret_cast<float(*)[32]>(&v[c * 32]);
                ld_chunk(T_RES + lane_off + (uint32_t)col0, vc);
                add_vec32(vc, sPar + PAR_B2 + col0);
                if (!last) {
                    st_chunk(T_RES + lane_off + (uint32_t)col0, vc);
                    stat_chunk(st, vc);
                }
                stage_store_f32(last ? OM_Y : OM_XNEXT, col0, last ? 0 : l, vc);
            }
            if (!last) {
                tmem_st_wait();
                bar_ew();  // every warp is done with this layer's parameters
                commit_params();
            }
        }
        if (lane == 0) tma_store_wait_all();
        __syncwarp();
        tc_fence_before();
    }
    __syncthreads();
    if (warp == EW_WARPS + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace lb

#ifdef PIO_LB_TRACE
extern "C" int pio_debug_lb_trace(long long* out, int max_pairs, int who, int reset) {
    int n = 0;
    cudaMemcpyFromSymbol(&n, lb::g_trace_n, sizeof(int), who * sizeof(int));
    n = std::min(n, max_pairs);
    if (out && n > 0) {
        cudaMemcpyFromSymbol(out, lb::g_trace, (size_t)n * 8, (size_t)(who * 2 + 0) * 8192 * 8);
        cudaMemcpyFromSymbol(out + max_pairs, lb::g_trace, (size_t)n * 8, (size_t)(who * 2 + 1) * 8192 * 8);
    }
    if (reset) {
        int z[2] = {0, 0};
        cudaMemcpyToSymbol(lb::g_trace_n, z, sizeof(z));
    }
    return n;
}
#endif

// When the fused kernel is used.  One CTA runs a whole layer for 128 rows, so its time per layer barely depends on
// the batch (measured on B200, 6 layers, N = 64: 42.7 us per layer at 32 tiles, 44.8 at 64, 48.5 at 148), while the
// unfused operators spread every product over all SMs and scale with the rows (34.6 us per layer at 32 tiles, 51.5
// at 64, 80.0 at 148).  The crossover is near 50 tiles; below it the seven launches per layer are faster, because a
// tile's element-wise work (~80 K warp instructions per layer) and its 1.5 MB of weight + saved-state traffic
// per layer are then concentrated on a quarter of the SMs.  PIO_FUSED_LATENT_MIN_ROWS overrides the threshold
// (0 = always), PIO_NO_FUSED_LATENT switches the kernel off; both are read per call (tests compare the two paths).
bool latent_cluster_preferred(int64_t rows);
int latent_block_fwd_cluster(const float* x, const LatentFusedLayer* layers, int nlayers, int batch, int n, cudaStream_t st);

bool latent_fused_supported(int64_t rows, int n, int ch, int heads, float drop_p) {
    if (getenv("PIO_NO_FUSED_LATENT") != nullptr) return false;
    if (!(ch == lb::C && heads == lb::H && (n == 32 || n == 64 || n == 128) && !(drop_p > 0.f))) return false;
    if (latent_cluster_preferred(rows)) return true;  // the cluster kernel (pio_latent_cl.cu) wins at any tile count it takes
    int64_t min_rows = 56 * lb::ROWS;
    if (const char* e = getenv("PIO_FUSED_LATENT_MIN_ROWS")) min_rows = atoll(e);
    return rows >= min_rows;
}

int latent_block_fwd_fused(const float* x, const LatentFusedLayer* layers, int nlayers, int batch, int n, cudaStream_t st) {
    PIO_CHECK_ARG(x && layers && nlayers >= 1 && nlayers <= PIO_LATENT_FUSED_MAX_LAYERS, "latent_block_fwd_fused: bad arguments");
    if (latent_cluster_preferred((int64_t)batch * n)) return latent_block_fwd_cluster(x, layers, nlayers, batch, n, st);
    const int64_t rows = (int64_t)batch * n;
    if (rows == 0) return PIO_OK;
    lb::Params p{};
    for (int l = 0; l < nlayers; ++l) {
        const LatentFusedLayer& s = layers[l];
        PIO_TRY(make_tmap_bf16(&p.maps[l][0], s.wqkv, 3 * lb::C, lb::C, lb::C, 64));
        PIO_TRY(make_tmap_bf16(&p.maps[l][1], s.wo, lb::C, lb::C, lb::C, 256));
        PIO_TRY(make_tmap_bf16(&p.maps[l][2], s.w1, lb::C, lb::C, lb::C, 256));
        PIO_TRY(make_tmap_bf16(&p.maps[l][3], s.w2, lb::C, lb::C, lb::C, 256));
        lb::LayerPtrs& d = p.layer[l];
        d.ln1_w = s.ln1_w; d.ln1_b = s.ln1_b; d.bqkv = s.bqkv; d.bo = s.bo;
        d.ln2_w = s.ln2_w; d.ln2_b = s.ln2_b; d.b1 = s.b1; d.b2 = s.b2;
        d.sa_stats = s.sa_stats; d.sa_h0 = (lb::bf16*)s.sa_h0; d.qkv = (lb::bf16*)s.qkv; d.o32 = s.o32; d.o16 = (lb::bf16*)s.o16;
        d.lse = s.lse; d.xmid = s.xmid; d.mlp_stats = s.mlp_stats; d.mlp_h0 = (lb::bf16*)s.mlp_h0; d.mlp_h1 = (lb::bf16*)s.mlp_h1;
        d.mlp_g = (lb::bf16*)s.mlp_g; d.xout = s.xout;
    }
    {
        using namespace lb;
        const LatentFusedLayer& f0 = layers[0];
        auto stride_of = [&](const void* a0, const void* a1) -> int64_t {
            return nlayers > 1 ? (int64_t)((const char*)a1 - (const char*)a0) : (int64_t)256;
        };
        const LatentFusedLayer& f1 = layers[nlayers > 1 ? 1 : 0];
        for (int l = 2; l < nlayers; ++l)  // the layers' slabs must be equally spaced (they are: BlockLayout::per_layer)
            PIO_CHECK_ARG((const char*)layers[l].xmid - (const char*)layers[l - 1].xmid == (const char*)f1.xmid - (const char*)f0.xmid,
                          "latent_block_fwd_fused: saved slabs are not equally spaced");
        PIO_TRY(make_out_map(&p.out[OM_SA_H0], false, f0.sa_h0, C, rows, nlayers, C * 2, stride_of(f0.sa_h0, f1.sa_h0), 64, ROWS));
        PIO_TRY(make_out_map(&p.out[OM_QKV], false, f0.qkv, 3 * C, rows, nlayers, 3 * C * 2, stride_of(f0.qkv, f1.qkv), 64, ROWS));
        PIO_TRY(make_out_map(&p.out[OM_O16], false, f0.o16, C, rows, nlayers, C * 2, stride_of(f0.o16, f1.o16), 64, ROWS));
        PIO_TRY(make_out_map(&p.out[OM_MLP_H0], false, f0.mlp_h0, C, rows, nlayers, C * 2, stride_of(f0.mlp_h0, f1.mlp_h0), 64, ROWS));
        PIO_TRY(make_out_map(&p.out[OM_MLP_G], false, f0.mlp_g, C, rows, nlayers, C * 2, stride_of(f0.mlp_g, f1.mlp_g), 64, ROWS));
        PIO_TRY(make_out_map(&p.out[OM_MLP_H1], false, f0.mlp_h1, C, rows, nlayers, C * 2, stride_of(f0.mlp_h1, f1.mlp_h1), 64, 32));
        PIO_TRY(make_out_map(&p.out[OM_XMID], true, f0.xmid, C, rows, nlayers, C * 4, stride_of(f0.xmid, f1.xmid), 32, 32));
        if (nlayers > 1)  // (the last layer's output is the OM_Y map; the slabs are spaced like every other saved tensor)
            PIO_TRY(make_out_map(&p.out[OM_XNEXT], true, f0.xout, C, rows, nlayers - 1, C * 4, stride_of(f0.xmid, f1.xmid), 32, 32));
        PIO_TRY(make_out_map(&p.out[OM_Y], true, layers[nlayers - 1].xout, C, rows, 1, C * 4, 256, 32, 32));
    }
    p.x = x;
    p.layers = nlayers;
    p.n = n;
    p.total_rows = rows;
    p.scale_log2e = (1.0f / sqrtf((float)lb::DH)) * 1.44269504088896340736f;
    const size_t smem = (size_t)lb::A_BYTES + 3 * lb::HEAD_BYTES + lb::STAGES * lb::STAGE_BYTES + 4 * lb::ROWS * sizeof(float2) +
                        lb::PAR_FLOATS * sizeof(float) + 16 * sizeof(uint64_t) + 1024;
    static_assert((size_t)lb::A_BYTES + 3 * lb::HEAD_BYTES + lb::STAGES * lb::STAGE_BYTES + 4 * lb::ROWS * sizeof(float2) +
                          lb::PAR_FLOATS * sizeof(float) + 16 * sizeof(uint64_t) + 1024 <= 232448,
                  "latent block kernel: more than 227 KB of shared memory");
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static bool configured[64] = {};
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        PIO_CUDA(cudaFuncSetAttribute(lb::latent_block_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev] = true;
    }
    const unsigned grid = (unsigned)cdiv(rows, lb::ROWS);
    PIO_CUDA(launch_pdl(lb::latent_block_fwd_kernel, dim3(grid), dim3(lb::THREADS), smem, st, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
