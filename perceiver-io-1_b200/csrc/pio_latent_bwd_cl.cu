// Fused latent self-attention block, BACKWARD input-gradient chain, one cluster of four CTAs per 128-row tile
// (PIO_MODE_BF16; C = 256 channels, 4 heads of 64).
//
// Reference: the backward of perceiver/model.py:36-44 (self_attention_layer / self_attention_block), :102-116
// (SelfAttention), :20-26 (mlp), :47-56 (Residual), which the reference leaves to autograd.  One launch walks ALL
// layers of a block from the last to the first and produces, per layer, every gradient with respect to an
// ACTIVATION; the gradients with respect to parameters are contractions over all rows of the batch and stay with the
// kernels that already form them (split-K tcgen05 GEMMs, column sums), which read the bf16 tensors this kernel leaves
// behind.  Per layer, with g = dL/d(layer output) held as fp32 in tensor memory (the CTA's 64 columns):
//
//   MLP:        dG = g W2            dH = dG * gelu'(h)         dN2 = dH W1
//               g_mid = g + LayerNorm2-backward(dN2; x_mid)
//   attention:  dO = g_mid Wo        (head r:)  S = Q K^T, P = exp(S / 8 - lse), dP = dO V^T,
//               dS = P * (dP - sum_k P dP) / 8,   dQ = dS K,  dK = dS^T Q,  dV = P^T dO
//               dN1 = [dQ | dK | dV] Wqkv        g_in = g_mid + LayerNorm1-backward(dN1; x)
//
// Work split as in the forward cluster kernel (pio_latent_cl.cu): CTA r of the cluster owns head r and channels
// [64 r, 64 r + 64).  dG, dN2 and dO contract over all 256 channels, so their A operands (g, dH, g_mid as bf16) are
// all-gathered over distributed shared memory, 16 KB per CTA and peer; the attention backward of head r is local to
// CTA r (keys = the tile's own rows, so dK / dV rows are the CTA's rows); dN1 contracts over the 768 columns of
// [dQ | dK | dV], of which a CTA holds the 192 of its head: it forms the K-partial product for ALL 256 output columns
// and the four partials are reduce-scattered — each CTA sends the three foreign 64-column slices of its partial as
// bf16 (16 KB each) and adds the three it receives to its own fp32 slice.  LayerNorm backward needs two row sums over
// all 256 channels: exchanged like the forward's statistics (st.async, 8 bytes per thread and peer).
//
// Left in global memory per layer for the batched weight-gradient kernel (pio_dw_batched.cu): g (bf16; dW2 = g^T G, db2),
// dH (dW1, db1), g_mid (dWo, dbo), [dQ | dK | dV] (dWqkv, dbqkv).  The LayerNorm parameter gradients are formed here:
// the column sums of dN and dN * xhat over a warp's 32 rows by a transposing shuffle butterfly, one atomic per lane
// (a separate kernel over bf16 copies of dN took 36-60 us per block beside the attention backward of the next operator;
// the sums cost ~0.8 K cycles per layer on the chain: step 1.902 -> 1.877 ms).
//
// Warp roles: warps 0..7 element-wise (TMEM lane quadrant = warp & 3, column half = warp >> 2), warp 8 = TMA producer
// of the weights (6 x 8 KB ring), warp 9 = TMEM allocator + MMA issuer + TMA stores, warp 10 = TMA loads of the saved
// activations (gelu', x_mid, x, Q / K / V of the head).
#include <cuda.h>
#include <stdlib.h>

#include <atomic>

#include "pio_latent_common.cuh"

namespace pio {

int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

namespace lbb {

using lb::bf16;
using namespace lb::cl;

constexpr int R = 4;
constexpr int C = 256;
constexpr int DH = 64;
constexpr int H = 4;
constexpr int ROWS = 128;
constexpr int EW_WARPS = 8;
constexpr int EW_THREADS = EW_WARPS * 32;
constexpr int THREADS = EW_THREADS + 96;
constexpr int SLOTS = 6;
constexpr int SLOT_BYTES = 8192;  // one 64 x 64 k-block of a weight slice, or 16 k-rows x 256 columns of the head's Wqkv rows
constexpr int ATOM = ROWS * 128;  // 16 KB: 128 rows x 64 bf16 (or x 32 fp32), 128-byte swizzle
constexpr int TMEM_COLS = 512;

constexpr size_t SMEM_BYTES = 4 * ATOM + 3 * ATOM + ATOM + 2 * ATOM + SLOTS * SLOT_BYTES + 8 * ROWS * sizeof(float2) +
                              2 * ROWS * sizeof(float) + 40 * sizeof(uint64_t) + 1024;
static_assert(SMEM_BYTES <= 232448, "cluster latent block backward kernel: more than 227 KB of shared memory");

struct LayerPtrs {
    const float* ln2_w;
    const float* ln1_w;
    const float* mlp_stats;  // (mean, rstd) per row
    const float* sa_stats;
    const float* lse;        // [sample][head][latent]
    float *dln2_w, *dln2_b, *dln1_w, *dln1_b;  // LayerNorm parameter gradients (C each), added into with atomics
};

enum InMap : int { IM_H1 = 0, IM_XMID, IM_XNEXT, IM_X, IM_QKV, IM_COUNT };
enum OutMap : int { OM_GY = 0, OM_GH, OM_GMID, OM_GQKV, OM_COUNT };

struct Params {
    CUtensorMap w[PIO_LATENT_FUSED_MAX_LAYERS][4];  // W2, W1, Wo (256 x 256, boxes of 64 x 64), Wqkv (768 x 256, boxes of 64 columns x 16 rows)
    CUtensorMap in[IM_COUNT];                       // (columns, rows, layer)
    CUtensorMap out[OM_COUNT];
    LayerPtrs layer[PIO_LATENT_FUSED_MAX_LAYERS];
    const float* dy;
    float* dx;
    int layers;
    int n;  // latents per sample: 32, 64 or 128
    int64_t total_rows;
    float scale_log2e, scale, log2e;
    int first_mlp_only;  // layer 0 (processed last) has no attention half of its own: 1 = MLP only, 2 = out-projection + MLP
    bf16* pre_do16;      // mode 2: the gradient of the attention output (rows x 256), written
};

__device__ __forceinline__ void bar_ew() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// Phase timeline for tuning (only in builds with -DPIO_LB_TRACE, tools/latent_cluster_trace.py --bwd): CTA 0's MMA
// thread (who 0) and thread 0 of its element-wise warps (who 1) log (tag, clock64) pairs.
#ifdef PIO_LB_TRACE
__device__ long long gb_trace[2][2][8192];
__device__ int gb_trace_n[2];
#define BW_TRACE(who, tag)                                               \
    do {                                                                 \
        if (blockIdx.x == 0) {                                           \
            const int _i = gb_trace_n[who];                              \
            if (_i < 8192) {                                             \
                gb_trace[who][0][_i] = (tag);                            \
                gb_trace[who][1][_i] = clock64();                        \
                gb_trace_n[who] = _i + 1;                                \
            }                                                            \
        }                                                                \
    } while (0)
__device__ unsigned long long gb_cta_time[4][1024];  // per CTA: kernel entry, after set-up, work done, exit (globaltimer, ns)
#define BW_STAMP(i)                                                                        \
    do {                                                                                   \
        if (threadIdx.x == 0 && blockIdx.x < 1024) {                                       \
            unsigned long long _t;                                                         \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t));                         \
            gb_cta_time[i][blockIdx.x] = _t;                                               \
        }                                                                                  \
    } while (0)
#else
#define BW_TRACE(who, tag) do { } while (0)
#define BW_STAMP(i) do { } while (0)
#endif
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

__global__ void __cluster_dims__(R, 1, 1) __launch_bounds__(THREADS, 1) latent_block_bwd_cluster_kernel(const __grid_constant__ Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;              // gathered A operand (atom j from CTA j); between the dO product and the reduce-scatter: P | dS
    uint8_t* sP = sA;                // [128 queries][128 keys] bf16 as two 64-key atoms
    uint8_t* sDS = sA + 2 * ATOM;
    uint8_t* sQ = sA + 4 * ATOM;     // Q_r | K_r | V_r (TMA), then dQ | dK | dV, then the outgoing slices of the reduce-scatter
    uint8_t* sK = sQ + ATOM;
    uint8_t* sV = sK + ATOM;
    uint8_t* sDO = sV + ATOM;        // gelu' of the CTA's hidden columns (TMA), later dO_r
    uint8_t* sX = sDO + ATOM;        // x_mid, later x: the CTA's 64 fp32 columns as two 32-column boxes
    uint8_t* ring = sX + 2 * ATOM;   // [SLOTS][8 KB] weight tiles
    float2* sStat = (float2*)(ring + SLOTS * SLOT_BYTES);  // [8 partials = (CTA, column half)][128 rows]
    float* sDelta = (float*)(sStat + 8 * ROWS);            // [2 column halves][128 rows]
    uint64_t* bars = (uint64_t*)(sDelta + 2 * ROWS);
    uint64_t* full = bars;
    uint64_t* empty = bars + SLOTS;
    uint64_t* mma_done = bars + 2 * SLOTS;
    uint64_t* ew_done = mma_done + 1;
    uint64_t* gath_full = mma_done + 2;   // [R]
    uint64_t* gath_free = mma_done + 6;
    uint64_t* stat_full = mma_done + 7;
    uint64_t* sa_free = mma_done + 8;     // the TMA store out of the CTA's own atom has read it
    uint64_t* h1_full = mma_done + 9;
    uint64_t* x_full = mma_done + 10;
    uint64_t* qkv_full = mma_done + 11;
    uint64_t* do_free = mma_done + 12;    // the products that read sDO are done: the next layer's gelu' may land there
    uint64_t* x_free = mma_done + 13;     // the element-wise warps have read sX
    uint64_t* q_free = mma_done + 14;     // the reduce-scatter's copies out of sQ / sK / sV have been received
    uint64_t* qs_free = mma_done + 15;    // the TMA stores of dQ / dK / dV have read sQ / sK / sV
    uint64_t* w2_full = mma_done + 16;    // the second half of the head's Wqkv rows has landed in sA (free after the dK / dV products)
    uint64_t* sa_w_free = mma_done + 17;  // P | dS have been read: sA may take that half
    uint32_t* tmem_slot = (uint32_t*)(mma_done + 18);

    pdl_trigger();
    BW_STAMP(0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = (int)cluster_rank();
    const int tile = blockIdx.x / R;

    if (threadIdx.x == 0) {
        for (int s = 0; s < SLOTS; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(mma_done, 1);
        mbar_init(ew_done, EW_WARPS);
        for (int j = 0; j < R; ++j) mbar_init(&gath_full[j], j == r ? 4 : 1);
        mbar_init(gath_free, R);
        mbar_init(stat_full, 1);
        mbar_init(sa_free, 1);
        mbar_init(h1_full, 1);
        mbar_init(x_full, 1);
        mbar_init(qkv_full, 1);
        mbar_init(do_free, 1);
        mbar_init(x_free, EW_WARPS);
        mbar_init(q_free, 1);
        mbar_init(qs_free, 1);
        mbar_init(w2_full, 1);
        mbar_init(sa_w_free, 1);
        fence_barrier_init();
    }
    if (warp == EW_WARPS + 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t T_D = tmem_base;          // g: the gradient stream, the CTA's 64 columns
    const uint32_t T_G = tmem_base + 64;     // dG slice; later dQ
    const uint32_t T_N = tmem_base + 128;    // dN2 slice; later dK
    const uint32_t T_O = tmem_base + 192;    // dO_r; later dV
    const uint32_t T_S = tmem_base + 256;    // S (128 keys); later, with T_P, the 256-column partial of dN1
    const uint32_t T_P = tmem_base + 384;    // dP (128 keys)
    pdl_wait();
    BW_STAMP(1);

    if (warp == EW_WARPS) {
        // ------------------------------------------------------------ TMA producer: weights
        if (lane == 0) {
            int prod = 0;
            auto acquire = [&]() -> uint8_t* {
                const int s = prod % SLOTS;
                mbar_wait(&empty[s], ((uint32_t)(prod / SLOTS) & 1u) ^ 1u);
                mbar_expect_tx(&full[s], SLOT_BYTES);
                return ring + s * SLOT_BYTES;
            };
            for (int l = p.layers - 1; l >= 0; --l) {
                if (l == p.layers - 1)
                    for (int w = 0; w < 4; ++w) prefetch_tmap(&p.w[l][w]);
                const bool mlp_only = (l == 0 && p.first_mlp_only);
                for (int w = 0; w < ((mlp_only && p.first_mlp_only == 1) ? 2 : 3); ++w)  // W2, W1, Wo: k-blocks (output-feature rows) in the order the atoms of the operand arrive
                    for (int i = 0; i < R; ++i) {
                        const int a = (r + R - i) % R;
                        uint8_t* dst = acquire();
                        tma_load_2d(dst, &p.w[l][w], r * DH, a * 64, &full[prod % SLOTS]);
                        ++prod;
                    }
                // Wqkv rows of head r (q, k, v), 16 at a time, all 256 columns (96 KB): the first six steps through the ring; the
                // ring cannot hold the rest in flight (each refill is a ~2 K-cycle round trip: the product took 5.4 K cycles
                // waiting for them), so the last six go to sA, which is idle from the dK / dV products to the reduce-scatter
                if (mlp_only) break;
                for (int kk = 0; kk < 6; ++kk) {
                    uint8_t* dst = acquire();
                    const int row0 = (kk >> 2) * C + r * DH + (kk & 3) * 16;
#pragma unroll
                    for (int j = 0; j < 4; ++j) tma_load_2d(dst + j * 2048, &p.w[l][3], j * 64, row0, &full[prod % SLOTS]);
                    ++prod;
                }
                mbar_wait(sa_w_free, (uint32_t)(p.layers - 1 - l) & 1u);
                mbar_expect_tx(w2_full, 6 * SLOT_BYTES);
                for (int kk = 6; kk < 12; ++kk) {
                    const int row0 = (kk >> 2) * C + r * DH + (kk & 3) * 16;
#pragma unroll
                    for (int j = 0; j < 4; ++j) tma_load_2d(sA + (kk - 6) * SLOT_BYTES + j * 2048, &p.w[l][3], j * 64, row0, w2_full);
                }
            }
        }
        __syncwarp();
    } else if (warp == EW_WARPS + 2) {
        // ------------------------------------------------------------ TMA loads of the saved activations
        if (lane == 0) {
            const int row0 = tile * ROWS;
            uint32_t dof = 0, xf = 0, qf = 0;
            for (int i = 0; i < IM_COUNT; ++i) prefetch_tmap(&p.in[i]);
            for (int l = p.layers - 1; l >= 0; --l) {
                const bool first = (l == p.layers - 1);
                if (!first) {
                    mbar_wait(do_free, dof);
                    dof ^= 1u;
                }
                mbar_expect_tx(h1_full, ATOM);
                tma_load_3d(sDO, &p.in[IM_H1], r * DH, row0, l, h1_full);
                if (!first) {
                    mbar_wait(x_free, xf);
                    xf ^= 1u;
                }
                mbar_expect_tx(x_full, 2 * ATOM);
                tma_load_3d(sX, &p.in[IM_XMID], r * DH, row0, l, x_full);
                tma_load_3d(sX + ATOM, &p.in[IM_XMID], r * DH + 32, row0, l, x_full);
                if (l == 0 && p.first_mlp_only) break;  // an MLP-only layer reads neither Q / K / V nor its own input
                if (!first) {
                    mbar_wait(q_free, qf);
                    qf ^= 1u;
                }
                mbar_expect_tx(qkv_full, 3 * ATOM);
                tma_load_3d(sQ, &p.in[IM_QKV], r * DH, row0, l, qkv_full);
                tma_load_3d(sK, &p.in[IM_QKV], C + r * DH, row0, l, qkv_full);
                tma_load_3d(sV, &p.in[IM_QKV], 2 * C + r * DH, row0, l, qkv_full);
                mbar_wait(x_free, xf);  // x_mid has been read
                xf ^= 1u;
                mbar_expect_tx(x_full, 2 * ATOM);
                const CUtensorMap* mx = (l == 0) ? &p.in[IM_X] : &p.in[IM_XNEXT];
                const int lx = (l == 0) ? 0 : l - 1;
                tma_load_3d(sX, mx, r * DH, row0, lx, x_full);
                tma_load_3d(sX + ATOM, mx, r * DH + 32, row0, lx, x_full);
            }
        }
        __syncwarp();
    } else if (warp == EW_WARPS + 1) {
        // ------------------------------------------------------------ MMA issuer (+ TMA stores of the operand tiles)
        if (lane == 0) {
            const uint32_t idesc_slice = make_idesc(64, false, true);    // A K-major (gathered operand), B = W[k][n]: n contiguous
            const uint32_t idesc_s = make_idesc(128, false, false);
            const uint32_t idesc_dq = make_idesc(DH, false, true);
            const uint32_t idesc_dkv = make_idesc(DH, true, true);
            const uint32_t idesc_part = make_idesc(256, false, true);
            const uint32_t a_addr = smem_u32(sA), q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV);
            const uint32_t do_addr = smem_u32(sDO), p_addr = smem_u32(sP), ds_addr = smem_u32(sDS);
            const uint8_t* own_atom = sA + r * ATOM;
            const int row0 = tile * ROWS;
            int cons = 0;
            uint32_t ew_phase = 0, g_phase = 0, qkv_phase = 0;
            int tr = 0;
            auto wait_ew = [&]() {
                BW_TRACE(0, 1000 + tr);
                mbar_wait(ew_done, ew_phase);
                ew_phase ^= 1u;
                tc_fence_after();
                BW_TRACE(0, 2000 + tr);
                ++tr;
            };
            auto wait_atom = [&](int j) -> int {
                const int a = (r + R - j) % R;
                if (j == 0) BW_TRACE(0, 1000 + tr);
                mbar_wait(&gath_full[a], g_phase);
                tc_fence_after();
                if (j == 0) BW_TRACE(0, 2000 + tr);
                if (j == R - 1) {
                    g_phase ^= 1u;
                    BW_TRACE(0, 3000 + tr);
                    ++tr;
                }
                return a;
            };
            auto ring_wait = [&]() -> uint32_t {
                const int s = cons % SLOTS;
                mbar_wait(&full[s], (uint32_t)(cons / SLOTS) & 1u);
                tc_fence_after();
                return smem_u32(ring + s * SLOT_BYTES);
            };
            auto ring_release = [&]() {
                umma_commit(&empty[cons % SLOTS]);
                ++cons;
            };
            // D[128 x 64] = gathered operand [128 x 256] . W[:, 64 r .. 64 r + 64), k-block by k-block as the atoms arrive;
            // the CTA's own atom is saved (tensor `which`) as soon as it is complete
            auto gathered_product = [&](uint32_t tmem_d, int which, int l) {
                for (int j = 0; j < R; ++j) {
                    const int a = wait_atom(j);
                    if (j == 0) {
                        lb::tma_store_3d(&p.out[which], own_atom, r * DH, row0, l);
                        lb::tma_store_commit();
                    }
                    const uint32_t b = ring_wait();
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem_d, make_smem_desc(a_addr + a * ATOM + k * 32, 16, 1024), make_smem_desc(b + k * 2048, 8192, 1024),
                                  idesc_slice, (j > 0 || k > 0) ? 1u : 0u);
                    ring_release();
                }
                umma_commit(mma_done);
                lb::tma_store_wait_read();
                lb::mbar_arrive(sa_free);
            };
            for (int l = p.layers - 1; l >= 0; --l) {
                gathered_product(T_G, OM_GY, l);    // dG[:, slice] = g W2[:, slice]
                gathered_product(T_N, OM_GH, l);    // dN2[:, slice] = dH W1[:, slice]
                if (l == 0 && p.first_mlp_only == 1) break;
                gathered_product(T_O, OM_GMID, l);  // dO_r = g_mid Wo[:, head r]
                if (l == 0 && p.first_mlp_only == 2) break;
                wait_ew();                          // dO_r as bf16 in sDO
                mbar_wait(qkv_full, qkv_phase);     // Q_r, K_r, V_r have landed
                qkv_phase ^= 1u;
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k)  // S = Q K^T
                    umma_bf16(T_S, make_smem_desc(q_addr + k * 32, 16, 1024), make_smem_desc(k_addr + k * 32, 16, 1024), idesc_s,
                              k > 0 ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < 4; ++k)  // dP = dO V^T
                    umma_bf16(T_P, make_smem_desc(do_addr + k * 32, 16, 1024), make_smem_desc(v_addr + k * 32, 16, 1024), idesc_s,
                              k > 0 ? 1u : 0u);
                umma_commit(mma_done);
                wait_ew();  // P, dS (bf16; zero outside the row's own sample) in shared memory
#pragma unroll
                for (int t = 0; t < 8; ++t)  // dQ = dS K: 16 keys per step
                    umma_bf16(T_G, make_smem_desc(ds_addr + (t >> 2) * ATOM + (t & 3) * 32, 16, 1024),
                              make_smem_desc(k_addr + t * 2048, 8192, 1024), idesc_dq, t > 0 ? 1u : 0u);
#pragma unroll
                for (int t = 0; t < 8; ++t)  // dK = dS^T Q: 16 queries per step, A read MN-major (two 64-key atoms 16 KB apart)
                    umma_bf16(T_N, make_smem_desc(ds_addr + t * 2048, ATOM, 1024), make_smem_desc(q_addr + t * 2048, 8192, 1024), idesc_dkv,
                              t > 0 ? 1u : 0u);
#pragma unroll
                for (int t = 0; t < 8; ++t)  // dV = P^T dO
                    umma_bf16(T_O, make_smem_desc(p_addr + t * 2048, ATOM, 1024), make_smem_desc(do_addr + t * 2048, 8192, 1024), idesc_dkv,
                              t > 0 ? 1u : 0u);
                umma_commit(mma_done);
                umma_commit(do_free);
                umma_commit(sa_w_free);
                wait_ew();  // dQ | dK | dV as bf16 in sQ | sK | sV
                lb::tma_store_3d(&p.out[OM_GQKV], sQ, r * DH, row0, l);
                lb::tma_store_3d(&p.out[OM_GQKV], sK, C + r * DH, row0, l);
                lb::tma_store_3d(&p.out[OM_GQKV], sV, 2 * C + r * DH, row0, l);
                lb::tma_store_commit();
                for (int kk = 0; kk < 6; ++kk) {  // partial of dN1 (all 256 columns) = [dQ | dK | dV] Wqkv[rows of head r, :]
                    const uint32_t b = ring_wait();
                    umma_bf16(T_S, make_smem_desc(q_addr + (kk >> 2) * ATOM + (kk & 3) * 32, 16, 1024), make_smem_desc(b, 2048, 1024), idesc_part,
                              kk > 0 ? 1u : 0u);
                    ring_release();
                }
                mbar_wait(w2_full, (uint32_t)(p.layers - 1 - l) & 1u);
                tc_fence_after();
#pragma unroll
                for (int kk = 6; kk < 12; ++kk)
                    umma_bf16(T_S, make_smem_desc(q_addr + (kk >> 2) * ATOM + (kk & 3) * 32, 16, 1024),
                              make_smem_desc(a_addr + (kk - 6) * SLOT_BYTES, 2048, 1024), idesc_part, 1u);
                umma_commit(mma_done);
                lb::tma_store_wait_read();
                lb::mbar_arrive(qs_free);
                for (int j = 0; j < R; ++j) wait_atom(j);  // the reduce-scatter (consumed by the element-wise warps): keeps the phase in step
            }
            lb::tma_store_wait_all();
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------ element-wise warps
        const int q = warp & 3, cg = warp >> 2;
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const int lcol = cg * 32;
        const int gcol = r * DH + lcol;
        const int64_t grow = (int64_t)tile * ROWS + row;
        const bool row_ok = grow < p.total_rows;
        const int n = p.n;
        const int64_t sample = grow / n;
        const int qi = (int)(grow - sample * n);
        const int kbase = (row / n) * n;
        uint8_t* own_atom = sA + r * ATOM;
        uint32_t mma_phase = 0, stat_phase = 0, free_phase = 0, sa_phase = 0, h1_phase = 0, x_phase = 0, qs_phase = 0;
        bool first_gather = true;
        uint32_t peer_atom[R - 1], peer_bar[R - 1];  // the atom this CTA fills and its gath_full barrier in the three peers
#pragma unroll
        for (int d = 1; d < R; ++d) {
            const uint32_t peer = (uint32_t)((r + d) % R);
            peer_atom[d - 1] = mapa(smem_u32(own_atom), peer);
            peer_bar[d - 1] = mapa(smem_u32(&gath_full[r]), peer);
        }

        int tr = 0;
        auto wait_mma = [&]() {
            if (threadIdx.x == 0) BW_TRACE(1, 1000 + tr);
            mbar_wait(mma_done, mma_phase);
            mma_phase ^= 1u;
            tc_fence_after();
            if (threadIdx.x == 0) BW_TRACE(1, 2000 + tr);
            ++tr;
        };
        auto free_arrive = [&]() {  // this CTA is done with the current contents of sA: tell every CTA of the cluster
            if (threadIdx.x == 0) {
                const uint32_t a = smem_u32(gath_free);
#pragma unroll
                for (int d = 0; d < R; ++d) mbar_arrive_peer(mapa(a, (uint32_t)d));
            }
        };
        auto signal = [&]() {
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) lb::mbar_arrive(ew_done);
        };
        // every CTA is done with the previous contents of sA (and, new_layer: the copies out of sQ / sK / sV have landed)
        auto wait_peers_free = [&](bool new_layer) {
            if (!first_gather) {
                mbar_wait(gath_free, free_phase);
                free_phase ^= 1u;
                if (new_layer && threadIdx.x == 0) lb::mbar_arrive(q_free);
            }
            if (threadIdx.x == 0) {
#pragma unroll
                for (int d = 1; d < R; ++d) mbar_expect_tx(&gath_full[(r + d) % R], ATOM);  // arrive + the 16 KB that peer sends
            }
        };
        // the two warps of a TMEM lane quadrant have written their 32 rows (4 KB) of the tiles to send: one thread copies them
        // into the three peers' atom number r and arrives on the local gath_full[r]; src[d] = the tile for peer (r + 1 + d) % 4
        auto send_quadrant = [&](const uint8_t* s0, const uint8_t* s1, const uint8_t* s2) {
            if (threadIdx.x == 0) BW_TRACE(1, 5000);
            tc_fence_before();
            fence_proxy_async();
            asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
            if (cg == 0 && lane == 0) {
                const uint32_t qoff = (uint32_t)q * 4096u;
                bulk_copy_to_peer(peer_atom[0] + qoff, smem_u32(s0) + qoff, 4096u, peer_bar[0]);
                bulk_copy_to_peer(peer_atom[1] + qoff, smem_u32(s1) + qoff, 4096u, peer_bar[1]);
                bulk_copy_to_peer(peer_atom[2] + qoff, smem_u32(s2) + qoff, 4096u, peer_bar[2]);
                lb::mbar_arrive(&gath_full[r]);
            }
            first_gather = false;
        };
        // 32 values of this thread's row -> bf16 -> the CTA's atom of the next gathered operand (all-gather)
        auto send_operand = [&](const float (&v)[32], bool wait_store, bool new_layer) {
            if (wait_store) {  // the saved copy of the previous operand has left the CTA's atom
                mbar_wait(sa_free, sa_phase);
                sa_phase ^= 1u;
            }
            wait_peers_free(new_layer);
            lb::store_half_row_smem(own_atom, row, cg, v);
            send_quadrant(own_atom, own_atom, own_atom);
        };
        // two row sums over all 256 channels: partial sums of this thread's 32 columns to all four CTAs, then the total
        auto send_sums = [&](float a, float b) {
            const uint32_t mine = smem_u32(&sStat[(r * 2 + cg) * ROWS + row]), sbar = smem_u32(stat_full);
            if (threadIdx.x == 0) mbar_expect_tx(stat_full, R * EW_THREADS * 8);
#pragma unroll
            for (int d = 0; d < R; ++d) st_async_f32x2(mapa(mine, (uint32_t)d), a, b, mapa(sbar, (uint32_t)d));
        };
        auto wait_sums = [&](float& a, float& b) {
            if (threadIdx.x == 0) BW_TRACE(1, 4000);
            mbar_wait(stat_full, stat_phase);
            stat_phase ^= 1u;
            if (threadIdx.x == 0) BW_TRACE(1, 4001);
            a = 0.f; b = 0.f;
#pragma unroll
            for (int g = 0; g < 2 * R; ++g) {  // same order in every thread of every CTA
                const float2 o = sStat[g * ROWS + row];
                a += o.x; b += o.y;
            }
        };
        // LayerNorm backward of this thread's 32 columns: dn = gradient of the normalised output, the LayerNorm input (fp32)
        // in sX, ms = (mean, rstd) of the row, gm = the affine scale (both loaded before the wait for the product: they
        // come from L2); adds the result to the gradient stream in T_D and returns the new stream values in v.  dn also
        // leaves as bf16 (dgamma / dbeta are formed from it by a kernel of their own): staged in the warp's own rows of
        // sX once it has read them (thread = row is the wrong shape for global stores), stored by TMA; the buffer is
        // released to the loader one phase later, when that store has read it (release_x).
        auto ln_prefetch = [&](const float* gamma, const float* stats, float (&gm)[32], float2& ms) {
            ms = row_ok ? __ldg(reinterpret_cast<const float2*>(stats) + grow) : make_float2(0.f, 0.f);
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(gamma + gcol + d));
                gm[d] = t.x; gm[d + 1] = t.y; gm[d + 2] = t.z; gm[d + 3] = t.w;
            }
        };
        // dgamma / dbeta: column sums over the warp's 32 rows by a transposing butterfly (31 shuffles: after the step with
        // distance d a lane keeps the half of its columns selected by its bit d), lane j then adds column j's sum
        auto col_sums_to = [&](float (&a)[32], float* dst) {
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                const bool up = (lane & off) != 0;
#pragma unroll
                for (int k = 0; k < off; ++k) {
                    const float send = up ? a[k] : a[k + off];
                    const float keep = up ? a[k + off] : a[k];
                    a[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                }
            }
            atomicAdd(dst + gcol + lane, a[0]);
        };
        auto ln_backward = [&](float (&dn)[32], const float (&gm)[32], const float2 ms, float* dgamma, float* dbeta, float (&v)[32]) {
            mbar_wait(x_full, x_phase);
            x_phase ^= 1u;
            float xh[32];
            uint8_t* mine = sX + cg * ATOM + q * 4096;  // this warp's 32 rows of its 32-column box
            lb::load_row_f32_smem(sX + cg * ATOM, row, xh);
            __syncwarp();
            if (lane == 0) lb::mbar_arrive(x_free);
#pragma unroll
            for (int i = 0; i < 32; ++i) xh[i] = (xh[i] - ms.x) * ms.y;
            {
                float t[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] = dn[i] * xh[i];
                col_sums_to(t, dgamma);
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] = dn[i];
                col_sums_to(t, dbeta);
            }
            float s1 = 0.f, s2 = 0.f;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                dn[i] *= gm[i];
                s1 += dn[i];
                s2 = fmaf(dn[i], xh[i], s2);
            }
            send_sums(s1, s2);
            lb::ld_chunk(T_D + lane_off + (uint32_t)lcol, v);
            float S1, S2;
            wait_sums(S1, S2);
            S1 *= (1.0f / C);
            S2 *= (1.0f / C);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = fmaf(ms.y, dn[i] - S1 - xh[i] * S2, v[i]);
        };

        // ---- gradient of the block's output: global -> tensor memory
        float v[32];
        if (row_ok) {
            const float4* src = reinterpret_cast<const float4*>(p.dy + grow * C + gcol);
#pragma unroll
            for (int d = 0; d < 8; ++d) {
                const float4 t = __ldg(src + d);
                v[4 * d] = t.x; v[4 * d + 1] = t.y; v[4 * d + 2] = t.z; v[4 * d + 3] = t.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
        }
        lb::st_chunk(T_D + lane_off + (uint32_t)lcol, v);
        tmem_st_wait();

        for (int l = p.layers - 1; l >= 0; --l) {
            const LayerPtrs& L = p.layer[l];
            const float lse2 = row_ok ? __ldg(L.lse + (sample * H + r) * n + qi) * p.log2e : 0.f;

            // ---- g (bf16) -> gather 1
            send_operand(v, false, true);

            // ---- dH = dG * gelu'(h) -> gather 2
            wait_mma();
            free_arrive();
            {
                float t[32], h1[32];
                lb::ld_chunk(T_G + lane_off + (uint32_t)lcol, t);
                mbar_wait(h1_full, h1_phase);
                h1_phase ^= 1u;
                lb::load_half_row_smem(sDO, row, cg, h1);
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] *= h1[i];
                send_operand(t, true, false);
            }

            // ---- g_mid = g + LayerNorm2-backward(dN2) -> tensor memory, gather 3
            float gm[32];
            float2 ms;
            ln_prefetch(L.ln2_w, L.mlp_stats, gm, ms);
            wait_mma();
            free_arrive();
            {
                float dn[32];
                lb::ld_chunk(T_N + lane_off + (uint32_t)lcol, dn);
                ln_backward(dn, gm, ms, L.dln2_w, L.dln2_b, v);
                if (l == 0 && p.first_mlp_only) {  // a first layer without attention of its own: this is the gradient of the block's input
                    if (row_ok) lb::store32_f32_global(p.dx + grow * C + gcol, v);
                    if (p.first_mlp_only == 1) break;
                    // mode 2: dO = g_mid Wo (the CTA's 64 columns = head r's) for the attention backward of the layer in front
                    send_operand(v, true, false);
                    wait_mma();
                    float t[32];
                    lb::ld_chunk(T_O + lane_off + (uint32_t)lcol, t);
                    if (row_ok) lb::store32_bf16_global(p.pre_do16 + grow * C + gcol, t);
                    break;
                }
                lb::st_chunk(T_D + lane_off + (uint32_t)lcol, v);
                tmem_st_wait();
                send_operand(v, true, false);
            }

            // ---- dO_r -> bf16 operand tile (over gelu')
            wait_mma();
            {
                float t[32];
                lb::ld_chunk(T_O + lane_off + (uint32_t)lcol, t);
                lb::store_half_row_smem(sDO, row, cg, t);
            }
            signal();

            // ---- P = exp(S / 8 - lse), dS = P (dP - sum_k P dP) / 8 over the row's own sample; 32-key chunks cg and cg + 2
            wait_mma();
            mbar_wait(sa_free, sa_phase);  // the saved copy of g_mid has left sA: P and dS go there
            sa_phase ^= 1u;
            {
                float part = 0.f;
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int key0 = (cg + 2 * cc) * 32;
                    if (key0 >= kbase && key0 < kbase + n) {  // warp-uniform
                        float s[32], dp[32];
                        lb::ld_chunk(T_S + lane_off + (uint32_t)key0, s);
                        lb::ld_chunk(T_P + lane_off + (uint32_t)key0, dp);
#pragma unroll
                        for (int i = 0; i < 32; ++i) part = fmaf(lb::ex2(fmaf(s[i], p.scale_log2e, -lse2)), dp[i], part);
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
                    if (key0 >= kbase && key0 < kbase + n) {
                        lb::ld_chunk(T_S + lane_off + (uint32_t)key0, s);
                        lb::ld_chunk(T_P + lane_off + (uint32_t)key0, dp);
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            s[i] = lb::ex2(fmaf(s[i], p.scale_log2e, -lse2));
                            dp[i] = s[i] * (dp[i] - delta) * p.scale;
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
            }
            signal();

            // ---- dQ | dK | dV -> bf16 tiles over Q | K | V: operand of the dN1 partial, saved for dWqkv
            wait_mma();
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const int chunk = cg * 3 + j;  // chunks 0, 1 = dQ (T_G), 2, 3 = dK (T_N), 4, 5 = dV (T_O): contiguous columns
                float t[32];
                lb::ld_chunk(T_G + lane_off + (uint32_t)(chunk * 32), t);
                lb::store_half_row_smem(sQ + (chunk >> 1) * ATOM, row, chunk & 1, t);
            }
            signal();

            // ---- reduce-scatter of the four dN1 partials, LayerNorm1-backward -> g for the next (earlier) layer
            ln_prefetch(L.ln1_w, L.sa_stats, gm, ms);
            wait_mma();
            free_arrive();  // sA (P | dS, then the second half of the Wqkv rows) is free from here on: gather 3 counts as consumed only now
            mbar_wait(qs_free, qs_phase);  // the saved copies of dQ | dK | dV have left their tiles: staging for the foreign slices
            qs_phase ^= 1u;
            wait_peers_free(false);
            {
                float dn[32];
#pragma unroll
                for (int d = 0; d < R; ++d) {
                    const int dst = (r + d) % R;  // d = 0: the CTA's own slice stays in registers
                    float t[32];
                    lb::ld_chunk(T_S + lane_off + (uint32_t)(dst * 64 + lcol), t);
                    if (d == 0) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) dn[i] = t[i];
                    } else {
                        lb::store_half_row_smem(sQ + (d - 1) * ATOM, row, cg, t);
                    }
                }
                send_quadrant(sQ, sK, sV);
#pragma unroll
                for (int d = 1; d < R; ++d) {
                    const int src = (r + R - d) % R;  // in the order the peers send
                    if (threadIdx.x == 0 && d == 1) BW_TRACE(1, 6000);
                    mbar_wait(&gath_full[src], 1u);   // (four gathers per layer: the reduce-scatter is always an odd phase)
                    if (threadIdx.x == 0 && d == R - 1) BW_TRACE(1, 6001);
                    float t[32];
                    lb::load_half_row_smem(sA + src * ATOM, row, cg, t);
#pragma unroll
                    for (int i = 0; i < 32; ++i) dn[i] += t[i];
                }
                bar_ew();       // every warp of the CTA has read the received slices:
                free_arrive();  // the reduce-scatter counts as consumed
                ln_backward(dn, gm, ms, L.dln1_w, L.dln1_b, v);
                if (l > 0) {
                    lb::st_chunk(T_D + lane_off + (uint32_t)lcol, v);
                    tmem_st_wait();
                } else if (row_ok) {
                    lb::store32_f32_global(p.dx + grow * C + gcol, v);
                }
            }
        }
        if (lane == 0) lb::tma_store_wait_all();
        __syncwarp();
        tc_fence_before();
    }
    __syncthreads();
    BW_STAMP(2);
    cluster_sync_all();
    if (warp == EW_WARPS + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
    BW_STAMP(3);
}

}  // namespace lbb

#ifdef PIO_LB_TRACE
extern "C" int pio_debug_lbb_cta_times(unsigned long long* out, int ctas) {
    cudaMemcpyFromSymbol(out, lbb::gb_cta_time, sizeof(unsigned long long) * 4 * 1024);
    return ctas;
}

extern "C" int pio_debug_lbb_trace(long long* out, int max_pairs, int who, int reset) {
    int n = 0;
    cudaMemcpyFromSymbol(&n, lbb::gb_trace_n, sizeof(int), who * sizeof(int));
    n = std::min(n, max_pairs);
    if (out && n > 0) {
        cudaMemcpyFromSymbol(out, lbb::gb_trace, (size_t)n * 8, (size_t)(who * 2 + 0) * 8192 * 8);
        cudaMemcpyFromSymbol(out + max_pairs, lbb::gb_trace, (size_t)n * 8, (size_t)(who * 2 + 1) * 8192 * 8);
    }
    if (reset) {
        int z[2] = {0, 0};
        cudaMemcpyToSymbol(lbb::gb_trace_n, z, sizeof(z));
    }
    return n;
}
#endif

// dgamma += sum_rows dn * xhat, dbeta += sum_rows dn for the LayerNorms whose backward ran inside the fused kernel, a
// table of them per launch: dn (bf16, the gradient of the normalised output before the affine scale), x (the LayerNorm
// input), (mean, rstd).  grid (C / 64, row chunks, items), 256 threads = 64 columns x 4 row phases; fp32 atomics into
// the (zero-filled or running) gradients.
struct LnGradTable {
    LnGradItem item[PIO_LN_GRAD_MAX_ITEMS];
};
__global__ void __launch_bounds__(256) ln_param_grads_kernel(const __grid_constant__ LnGradTable tab, int64_t rows, int c,
                                                             int64_t rows_per_chunk) {
    // thread = 4 consecutive columns x one of 16 row phases: a warp reads two rows of the block's 64 columns per step
    // (128 bytes of dn, 256 of x per row), four steps in flight
    __shared__ float4 s_g[16][16], s_b[16][16];
    pdl_trigger();
    pdl_wait();
    const LnGradItem& it = tab.item[blockIdx.z];
    const __nv_bfloat16* __restrict__ dn = (const __nv_bfloat16*)it.dn16;
    const float* __restrict__ x = it.x;
    const float2* __restrict__ stats = reinterpret_cast<const float2*>(it.stats);
    const int cq = threadIdx.x & 15, rp = threadIdx.x >> 4;
    const int col = blockIdx.x * 64 + cq * 4;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    float ag[4] = {0.f, 0.f, 0.f, 0.f}, ab[4] = {0.f, 0.f, 0.f, 0.f};
    if (col < c) {
        for (int64_t rb = r0 + rp; rb < r1; rb += 64) {
            uint2 d[4];
            float4 xv[4];
            float2 ms[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t r = rb + 16 * u;
                if (r < r1) {
                    d[u] = *reinterpret_cast<const uint2*>(dn + r * c + col);
                    xv[u] = *reinterpret_cast<const float4*>(x + r * c + col);
                    ms[u] = __ldg(stats + r);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (rb + 16 * u < r1) {
                    const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&d[u].x));
                    const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&d[u].y));
                    const float dv[4] = {lo.x, lo.y, hi.x, hi.y};
                    const float xs[4] = {xv[u].x, xv[u].y, xv[u].z, xv[u].w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        ag[j] = fmaf(dv[j], (xs[j] - ms[u].x) * ms[u].y, ag[j]);
                        ab[j] += dv[j];
                    }
                }
            }
        }
    }
    s_g[rp][cq] = make_float4(ag[0], ag[1], ag[2], ag[3]);
    s_b[rp][cq] = make_float4(ab[0], ab[1], ab[2], ab[3]);
    __syncthreads();
    if (threadIdx.x < 128) {  // 64 columns x {dgamma, dbeta}
        const int cc = threadIdx.x & 63;
        const bool gamma = threadIdx.x < 64;
        const float* src = reinterpret_cast<const float*>(gamma ? &s_g[0][0] : &s_b[0][0]);
        float t = 0.f;
#pragma unroll
        for (int ph = 0; ph < 16; ++ph) t += src[ph * 64 + cc];
        const int gc = blockIdx.x * 64 + cc;
        if (gc < c) atomicAdd((gamma ? it.dgamma : it.dbeta) + gc, t);
    }
}

int ln_param_grads(const LnGradItem* items, int count, int64_t rows, int c, cudaStream_t st) {
    PIO_CHECK_ARG(items && count >= 1 && count <= PIO_LN_GRAD_MAX_ITEMS && c > 0 && c % 4 == 0 && rows >= 0, "ln_param_grads: bad arguments");
    if (rows == 0) return PIO_OK;
    LnGradTable tab{};
    for (int i = 0; i < count; ++i) {
        PIO_CHECK_ARG(items[i].dn16 && items[i].x && items[i].stats && items[i].dgamma && items[i].dbeta, "ln_param_grads: null pointer in item %d", i);
        tab.item[i] = items[i];
    }
    const int64_t col_blocks = cdiv(c, 64);
    const int64_t chunks = std::max<int64_t>(1, std::min<int64_t>(cdiv(4 * (int64_t)num_sms(), col_blocks * count), cdiv(rows, 64)));
    dim3 grid((unsigned)col_blocks, (unsigned)chunks, (unsigned)count);
    PIO_CUDA(launch_pdl(ln_param_grads_kernel, grid, dim3(256), 0, st, tab, rows, c, cdiv(rows, chunks)));
    count_launch();
    return PIO_OK;
}

// PIO_LATENT_BWD_CLUSTER = 0 / 1 forces the choice (read per call: tests compare the two paths)
bool latent_bwd_cluster_supported(int64_t rows, int n, int ch, int heads, float drop_p) {
    if (getenv("PIO_NO_FUSED_LATENT") != nullptr) return false;
    if (!(ch == lbb::C && heads == lbb::H && (n == 32 || n == 64 || n == 128) && !(drop_p > 0.f))) return false;
    if (const char* e = getenv("PIO_LATENT_BWD_CLUSTER")) return atoi(e) != 0;
    return cdiv(rows, (int64_t)lbb::ROWS) <= 66;
}

int latent_block_bwd_cluster(const float* dy, const float* x, float* dx, const LatentBwdLayer* layers, int nlayers, int batch, int n,
                             cudaStream_t st) {
    PIO_CHECK_ARG(dy && x && dx && layers && nlayers >= 1 && nlayers <= PIO_LATENT_FUSED_MAX_LAYERS, "latent_block_bwd_cluster: bad arguments");
    const int64_t rows = (int64_t)batch * n;
    if (rows == 0) return PIO_OK;
    using namespace lbb;
    Params p{};
    const int first_mlp_only = layers[0].wqkv != nullptr ? 0 : (layers[0].wo != nullptr ? 2 : 1);
    PIO_CHECK_ARG(first_mlp_only != 2 || layers[0].pre_do16 != nullptr, "latent_block_bwd_cluster: an out-projection first layer needs pre_do16");
    p.first_mlp_only = first_mlp_only;
    p.pre_do16 = (bf16*)layers[0].pre_do16;
    for (int l = 0; l < nlayers; ++l) {
        const LatentBwdLayer& s = layers[l];
        PIO_CHECK_ARG(s.wqkv != nullptr || l == 0, "latent_block_bwd_cluster: only the first layer may be MLP-only");
        PIO_TRY(make_tmap_bf16(&p.w[l][0], s.w2, C, C, C, 64));
        PIO_TRY(make_tmap_bf16(&p.w[l][1], s.w1, C, C, C, 64));
        PIO_TRY(make_tmap_bf16(&p.w[l][2], s.wo ? s.wo : s.w1, C, C, C, 64));
        PIO_TRY(make_tmap_bf16(&p.w[l][3], s.wqkv ? s.wqkv : s.w1, s.wqkv ? 3 * C : C, C, C, 16));
        LayerPtrs& d = p.layer[l];
        d.ln2_w = s.ln2_w; d.ln1_w = s.ln1_w; d.mlp_stats = s.mlp_stats; d.sa_stats = s.sa_stats; d.lse = s.lse;
        d.dln2_w = s.dln2_w; d.dln2_b = s.dln2_b; d.dln1_w = s.dln1_w; d.dln1_b = s.dln1_b;
        PIO_CHECK_ARG(d.dln2_w && d.dln2_b && (s.wqkv == nullptr || (d.dln1_w && d.dln1_b)), "latent_block_bwd_cluster: LayerNorm gradient destinations missing");
    }
    const LatentBwdLayer& f0 = layers[0];
    const LatentBwdLayer& f1 = layers[nlayers > 1 ? 1 : 0];
    auto stride_of = [&](const void* a0, const void* a1) -> int64_t {
        return nlayers > 1 ? (int64_t)((const char*)a1 - (const char*)a0) : (int64_t)256;
    };
    for (int l = 2; l < nlayers; ++l) {  // saved slabs and gradient buffers must be equally spaced over the layers
        PIO_CHECK_ARG((const char*)layers[l].xmid - (const char*)layers[l - 1].xmid == (const char*)f1.xmid - (const char*)f0.xmid &&
                          (const char*)layers[l].gy16 - (const char*)layers[l - 1].gy16 == (const char*)f1.gy16 - (const char*)f0.gy16,
                      "latent_block_bwd_cluster: per-layer buffers are not equally spaced");
    }
    PIO_TRY(make_out_map(&p.in[IM_H1], false, f0.mlp_h1, C, rows, nlayers, C * 2, stride_of(f0.mlp_h1, f1.mlp_h1), 64, ROWS));
    PIO_TRY(make_out_map(&p.in[IM_XMID], true, f0.xmid, C, rows, nlayers, C * 4, stride_of(f0.xmid, f1.xmid), 32, ROWS));
    PIO_TRY(make_out_map(&p.in[IM_XNEXT], true, f0.xnext, C, rows, std::max(nlayers - 1, 1), C * 4, stride_of(f0.xmid, f1.xmid), 32, ROWS));
    PIO_TRY(make_out_map(&p.in[IM_X], true, x, C, rows, 1, C * 4, 256, 32, ROWS));
    PIO_TRY(make_out_map(&p.in[IM_QKV], false, f0.qkv, 3 * C, rows, nlayers, 3 * C * 2, stride_of(f0.qkv, f1.qkv), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_GY], false, f0.gy16, C, rows, nlayers, C * 2, stride_of(f0.gy16, f1.gy16), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_GH], false, f0.gh16, C, rows, nlayers, C * 2, stride_of(f0.gh16, f1.gh16), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_GMID], false, f0.gmid16, C, rows, nlayers, C * 2, stride_of(f0.gmid16, f1.gmid16), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_GQKV], false, f0.gqkv16, 3 * C, rows, nlayers, 3 * C * 2, stride_of(f0.gqkv16, f1.gqkv16), 64, ROWS));
    p.dy = dy;
    p.dx = dx;
    p.layers = nlayers;
    p.n = n;
    p.total_rows = rows;
    p.scale = 1.0f / sqrtf((float)DH);
    p.log2e = 1.44269504088896340736f;
    p.scale_log2e = p.scale * p.log2e;
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(latent_block_bwd_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    const unsigned grid = (unsigned)(cdiv(rows, (int64_t)ROWS) * R);
    PIO_CUDA(launch_pdl(latent_block_bwd_cluster_kernel, dim3(grid), dim3(THREADS), SMEM_BYTES, st, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
