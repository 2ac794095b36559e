// Fused latent self-attention block, forward, one CLUSTER of four CTAs per 128-row tile (PIO_MODE_BF16; C = 256
// channels, 4 heads of 64).
//
// Reference: perceiver/model.py:36-44 (self_attention_layer / self_attention_block), :102-116 (SelfAttention),
// :20-26 (mlp), :47-56 (Residual).  Same arithmetic and the same saved-for-backward layout as
// latent_block_fwd_kernel (pio_latent_tc.cu), which runs a tile on ONE SM and is bound by that SM's instruction
// issue (~80 K warp instructions of LayerNorm / softmax / GELU per layer) and TMA throughput (1.5 MB per layer).
// Here the tile's COLUMNS are split over the four CTAs of a cluster, so each SM does a quarter of all of it:
//
//   * CTA r of the cluster owns head r and channels [64 r, 64 r + 64): its 64 columns of the fp32 residual stream
//     live in its tensor memory for the whole kernel, and it loads only the weight rows that produce its columns
//     (Wq_r, Wk_r, Wv_r; 64 rows each of Wo, W1, W2: 192 KB per layer instead of 768 KB);
//   * every product contracts over all 256 channels, so its A operand (LayerNorm output, attention output, GELU
//     output: 128 rows x 256 bf16 = four 16 KB swizzle atoms) is ALL-GATHERED: each CTA writes the atom it
//     produced into its own shared memory and copies it into the three other CTAs' (cp.async.bulk shared::cta ->
//     shared::cluster, 16 KB each, completing on the receiver's mbarrier).  Four gathers per layer;
//   * the out-projection and the second MLP product accumulate into the CTA's residual columns (D += A B), the
//     attention of head r (scores, softmax, P V with P through tensor memory) is local to CTA r;
//   * LayerNorm statistics: every warp sends its (mean, M2) over 32 columns of each row to all four CTAs
//     (st.shared::cluster, 8 bytes per row), every thread merges the eight partials in the same order (Chan), so
//     all four CTAs normalise with bit-identical (mean, rstd).
//
// Cross-CTA protocol (all mbarriers; no cluster-wide barrier inside the layer loop):
//   gath_full[j] (j = r: count 4, one arrival per TMEM lane quadrant; j != r: count 1 + 16 KB of transaction
//       bytes): atom j of the operand is complete in this CTA -> the MMA thread may issue that k-block's products.
//   gath_free (count 4): every CTA of the cluster, itself included, has seen the products that read the previous
//       gather complete (thread 0 of each CTA arrives remotely after its mma_done wait) -> thread 0 may overwrite
//       the peers' atoms with the next gather.
//   stat_full (count 1 + 8 KB of transaction bytes): all LayerNorm partials of this round have arrived (st.async).
// A peer can never run a whole phase ahead on any of them: each gather needs every CTA's previous product.
//
// Warp roles: warps 0..7 element-wise (TMEM lane quadrant = warp & 3, column half = warp >> 2: a thread owns 32 of
// its row's 64 local columns), warp 8 = TMA producer (weights, 4 x 24 KB ring), warp 9 = TMEM allocator + MMA
// issuer + TMA stores of the bf16 saved state straight out of the operand tiles.
#include <cuda.h>
#include <stdlib.h>

#include <atomic>

#include "pio_latent_common.cuh"

namespace pio {

int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

namespace lbc {

using lb::bf16;
using lb::Stat;

constexpr int R = 4;  // CTAs per cluster = heads = column slices
constexpr int C = 256;
constexpr int DH = 64;
constexpr int H = 4;
constexpr int ROWS = 128;
constexpr int EW_WARPS = 8;
constexpr int EW_THREADS = EW_WARPS * 32;
constexpr int THREADS = EW_THREADS + 64;
constexpr int STAGES = 4;
constexpr int STAGE_BYTES = 24576;
constexpr int QKV_STAGE_BYTES = 3 * 64 * 128;  // 64 rows each of Wq_r, Wk_r, Wv_r x one 64-column k-block: a whole stage
constexpr int HALF_SLICE_BYTES = 2 * 64 * 128;  // two 64 x 64 k-blocks of the CTA's rows of Wo / W1 / W2
constexpr int ATOM = ROWS * 128;               // 16 KB: 128 rows x 64 bf16, 128-byte swizzle
constexpr int TMEM_COLS = 512;
// this CTA's slice of the layer's vectors, 64 floats each
enum Par : int { PAR_LN1_W = 0, PAR_LN1_B, PAR_BQ, PAR_BK, PAR_BV, PAR_BO, PAR_LN2_W, PAR_LN2_B, PAR_B1, PAR_B2, PAR_COUNT };
constexpr int PAR_FLOATS = PAR_COUNT * 64;

constexpr size_t SMEM_BYTES = 4 * ATOM + 3 * ATOM + STAGES * STAGE_BYTES + 8 * ROWS * sizeof(float2) + ROWS * sizeof(float) +
                              2 * PAR_FLOATS * sizeof(float) + 24 * sizeof(uint64_t) + 1024;
static_assert(SMEM_BYTES <= 232448, "cluster latent block kernel: more than 227 KB of shared memory");

struct LayerPtrs {
    const float* par[PAR_COUNT];  // full-length vectors (ln1_w, ln1_b, bq, bk, bv, bo, ln2_w, ln2_b, b1, b2)
    float* sa_stats;
    float* mlp_stats;
    float* lse;
};

enum OutMap : int { OM_SA_H0 = 0, OM_QKV, OM_O16, OM_MLP_H0, OM_MLP_G, OM_MLP_H1, OM_XMID, OM_XNEXT, OM_Y, OM_COUNT };

struct Params {
    CUtensorMap maps[PIO_LATENT_FUSED_MAX_LAYERS][4];  // Wqkv (768 x 256), Wo, W1, W2 (256 x 256): boxes of 64 rows x 64 columns
    CUtensorMap out[OM_COUNT];
    LayerPtrs layer[PIO_LATENT_FUSED_MAX_LAYERS];
    const float* x;
    int layers;
    int n;  // latents per sample: 32, 64 or 128
    int64_t total_rows;
    float scale_log2e;
    int first_mlp_only;  // layer 0 has no attention half of its own: 1 = only the MLP of the cross-attention layer in front of the block,
                         // 2 = that layer's out-projection + residual (x += O Wo^T + bo, O given) and its MLP
    CUtensorMap o_map;   // mode 2: the attention output O (rows x 256, bf16) as (columns, rows, 1)
};

using namespace lb::cl;

__device__ __forceinline__ void bar_ew() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// Phase timeline for tuning (only in builds with -DPIO_LB_TRACE, tools/latent_trace.py --cluster): CTA 0's MMA thread
// (who 0) and thread 0 of its element-wise warps (who 1) log (tag, clock64) pairs; tag = kind * 1000 + running index.
#ifdef PIO_LB_TRACE
__device__ long long gc_trace[2][2][8192];
__device__ int gc_trace_n[2];
#define CL_TRACE(who, tag)                                               \
    do {                                                                 \
        if (blockIdx.x == 0) {                                           \
            const int _i = gc_trace_n[who];                              \
            if (_i < 8192) {                                             \
                gc_trace[who][0][_i] = (tag);                            \
                gc_trace[who][1][_i] = clock64();                        \
                gc_trace_n[who] = _i + 1;                                \
            }                                                            \
        }                                                                \
    } while (0)
#else
#define CL_TRACE(who, tag) do { } while (0)
#endif

__global__ void __cluster_dims__(R, 1, 1) __launch_bounds__(THREADS, 1) latent_block_fwd_cluster_kernel(const __grid_constant__ Params p) {
    extern __shared__ uint8_t smem_raw[];
    // (the dynamic shared-memory window starts at the same offset in every CTA of the cluster, so the rounded-up
    // addresses below are the same in all four and mapa() of a local address names the same object in a peer)
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                  // gathered A operand: atom j = channels [64 j, 64 j + 64), produced by CTA j
    uint8_t* sQ = sA + 4 * ATOM;         // Q_r | K_r | V_r tiles; sQ and sK double as staging of the fp32 residual copies,
    uint8_t* sK = sQ + ATOM;             // sV as the tile GELU' leaves from
    uint8_t* sV = sK + ATOM;
    uint8_t* ring = sV + ATOM;           // [STAGES][24 KB] weight tiles
    float2* sStat = (float2*)(ring + STAGES * STAGE_BYTES);  // [8 partials = (CTA, column half)][128 rows] (mean, M2) over 32 columns
    float* sInv = (float*)(sStat + 8 * ROWS);                // 1 / softmax denominator of each row
    float* sPar = sInv + ROWS;
    uint64_t* bars = (uint64_t*)(sPar + 2 * PAR_FLOATS);
    uint64_t* full = bars;               // [STAGES]
    uint64_t* empty = bars + STAGES;     // [STAGES]
    uint64_t* mma_done = bars + 2 * STAGES;
    uint64_t* ew_done = bars + 2 * STAGES + 1;
    uint64_t* gath_full = bars + 2 * STAGES + 2;  // [R]: atom j of sA holds CTA j's part of the new operand
    uint64_t* gath_free = bars + 2 * STAGES + 6;
    uint64_t* stat_full = bars + 2 * STAGES + 7;
    uint64_t* sa_free = bars + 2 * STAGES + 8;   // the TMA store out of the CTA's own atom of sA (and of sV) has read it
    uint64_t* qkv_free = bars + 2 * STAGES + 9;  // the TMA stores out of sQ / sK / sV have read them
    uint64_t* o_full = bars + 2 * STAGES + 10;   // mode 2: the attention output tile has landed in sA
    uint32_t* tmem_slot = (uint32_t*)(bars + 2 * STAGES + 11);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = (int)cluster_rank();
    const int tile = blockIdx.x / R;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(mma_done, 1);
        mbar_init(ew_done, EW_WARPS);
        for (int j = 0; j < R; ++j) mbar_init(&gath_full[j], j == r ? 4 : 1);
        mbar_init(gath_free, R);
        mbar_init(stat_full, 1);
        mbar_init(sa_free, 1);
        mbar_init(qkv_free, 1);
        mbar_init(o_full, 1);
        fence_barrier_init();
    }
    if (warp == EW_WARPS + 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();  // every CTA's barriers exist before anyone arrives on or copies into a peer
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t T_RES = tmem_base;        // the CTA's 64 residual columns
    const uint32_t T_QKV = tmem_base + 64;   // [Q_r | K_r | V_r] (192 columns); the CTA's 64 MLP hidden columns later
    const uint32_t T_S = tmem_base + 256;    // scores, 128 keys
    const uint32_t T_P = tmem_base + 384;    // probabilities, bf16 pairs (64 columns)
    const uint32_t T_O = tmem_base + 448;    // O_r, 64 columns
    pdl_wait();  // everything below reads what earlier kernels wrote (activations, bf16 weight shadows)

    if (warp == EW_WARPS) {
        // ------------------------------------------------------------ TMA producer: this CTA's weight rows
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
                if (l == 0) {
                    prefetch_tmap(mq); prefetch_tmap(&p.maps[l][1]); prefetch_tmap(&p.maps[l][2]); prefetch_tmap(&p.maps[l][3]);
                }
                const bool mlp_only = (l == 0 && p.first_mlp_only);
                if (mlp_only && p.first_mlp_only == 2) {  // the attention output of the cross-attention layer: all four atoms from global memory
                    prefetch_tmap(&p.o_map);
                    mbar_expect_tx(o_full, 4 * ATOM);
#pragma unroll
                    for (int a = 0; a < 4; ++a) lb::tma_load_3d(sA + a * ATOM, &p.o_map, a * 64, tile * ROWS, 0, o_full);
                }
                // k-blocks in the order the atoms of the gathered operand arrive: the CTA's own, then r - 1, r - 2, r - 3
                for (int i = 0; i < 4 && !mlp_only; ++i) {
                    const int kb = (r + R - i) % R;
                    uint8_t* dst = acquire(QKV_STAGE_BYTES);
                    uint64_t* bar = &full[prod % STAGES];
#pragma unroll
                    for (int j = 0; j < 3; ++j) tma_load_2d(dst + j * 8192, mq, kb * 64, j * C + r * DH, bar);
                    ++prod;
                }
                for (int w = (mlp_only && p.first_mlp_only == 1) ? 2 : 1; w < 4; ++w)  // rows [64 r, 64 r + 64) of Wo, W1, W2: four 64 x 64 k-blocks each, two per stage
                    for (int hf = 0; hf < 2; ++hf) {
                        uint8_t* dst = acquire(HALF_SLICE_BYTES);
                        uint64_t* bar = &full[prod % STAGES];
#pragma unroll
                        for (int kb = 0; kb < 2; ++kb) tma_load_2d(dst + kb * 8192, &p.maps[l][w], ((r + R - (hf * 2 + kb)) % R) * 64, r * DH, bar);
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
            const uint32_t idesc_64 = make_idesc(64, false, false);
            const uint32_t a_addr = smem_u32(sA), q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV);
            const uint8_t* own_atom = sA + r * ATOM;
            const int row0 = tile * ROWS;
            int cons = 0;
            uint32_t ew_phase = 0, g_phase = 0;
            int tr = 0;
            auto wait_ew = [&]() {
                CL_TRACE(0, 1000 + tr);
                mbar_wait(ew_done, ew_phase);
                ew_phase ^= 1u;
                tc_fence_after();
                CL_TRACE(0, 2000 + tr);
                ++tr;
            };
            // atom (r - j) mod 4 of the gathered operand has arrived (j = 0: the CTA's own, then in the order the peers send)
            auto wait_atom = [&](int j) -> int {
                const int a = (r + R - j) % R;
                if (j == 0) CL_TRACE(0, 1000 + tr);
                mbar_wait(&gath_full[a], g_phase);
                tc_fence_after();
                if (j == 0) CL_TRACE(0, 2000 + tr);
                if (j == R - 1) {
                    g_phase ^= 1u;
                    CL_TRACE(0, 3000 + tr);
                    ++tr;
                }
                return a;
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
            // the saved copy of the operand leaves the CTA's own atom by TMA; the element-wise warps may overwrite the atom
            // once the store has read it (off the critical path: they first wait for the product just committed)
            auto release_own_atom = [&]() {
                lb::tma_store_wait_read();
                lb::mbar_arrive(sa_free);
            };
            auto store_tile = [&](int which, const uint8_t* tile_ptr, int col0, int l) {  // one 128-row x 64-column bf16 atom
                lb::tma_store_3d(&p.out[which], tile_ptr, col0, row0, l);
            };
            // D[128 x 64] (+)= sA[128 x 256] W[64 r .. 64 r + 64, :]^T, k-block by k-block as the atoms of sA arrive (the
            // producer loads the weight k-blocks in the same order, two per ring stage); `which` = where the CTA's own
            // atom is saved, stored as soon as it is complete
            auto slice_product = [&](uint32_t tmem_d, bool accumulate, int which, int l, bool with_h1) {
                for (int hf = 0; hf < 2; ++hf) {
                    uint32_t b = 0;
#pragma unroll
                    for (int kb = 0; kb < 2; ++kb) {
                        const int j = hf * 2 + kb;
                        const int a = wait_atom(j);
                        if (j == 0) {
                            store_tile(which, own_atom, r * DH, l);
                            if (with_h1) store_tile(OM_MLP_H1, sV, r * DH, l);
                            lb::tma_store_commit();
                        }
                        if (kb == 0) b = ring_wait();
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16(tmem_d, make_smem_desc(a_addr + a * ATOM + k * 32, 16, 1024),
                                      make_smem_desc(b + kb * 8192 + k * 32, 16, 1024), idesc_64, (accumulate || j > 0 || k > 0) ? 1u : 0u);
                    }
                    ring_release();
                }
                umma_commit(mma_done);
                release_own_atom();
            };
            for (int l = 0; l < p.layers; ++l) {
                if (!(l == 0 && p.first_mlp_only)) {
                for (int j = 0; j < R; ++j) {  // [Q_r | K_r | V_r] = LN(x) [Wq_r; Wk_r; Wv_r]^T
                    const int a = wait_atom(j);
                    if (j == 0) {
                        store_tile(OM_SA_H0, own_atom, r * DH, l);
                        lb::tma_store_commit();
                    }
                    const uint32_t b = ring_wait();
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(T_QKV, make_smem_desc(a_addr + a * ATOM + k * 32, 16, 1024), make_smem_desc(b + k * 32, 16, 1024),
                                  idesc_qkv, (j > 0 || k > 0) ? 1u : 0u);
                    ring_release();
                }
                umma_commit(mma_done);
                release_own_atom();
                wait_ew();  // Q_r, K_r, V_r as bf16 in shared memory
                store_tile(OM_QKV, sQ, r * DH, l);
                store_tile(OM_QKV, sK, C + r * DH, l);
                store_tile(OM_QKV, sV, 2 * C + r * DH, l);
                lb::tma_store_commit();
#pragma unroll
                for (int k = 0; k < 4; ++k)  // S = Q_r K_r^T for all 128 x 128 (row, key) pairs of the tile
                    umma_bf16(T_S, make_smem_desc(q_addr + k * 32, 16, 1024), make_smem_desc(k_addr + k * 32, 16, 1024), idesc_s,
                              k > 0 ? 1u : 0u);
                umma_commit(mma_done);
                wait_ew();  // P (bf16 pairs, zero outside the row's own sample) in tensor memory
#pragma unroll
                for (int t = 0; t < 8; ++t)  // O_r = P V_r
                    umma_bf16_ts(T_O, T_P + (uint32_t)t * 8, make_smem_desc(v_addr + t * 2048, ROWS * 128, 1024), idesc_pv, t > 0 ? 1u : 0u);
                umma_commit(mma_done);
                lb::tma_store_wait_read();  // Q, K, V have left their tiles (staging / GELU' use them next)
                lb::mbar_arrive(qkv_free);

                slice_product(T_RES, true, OM_O16, l, false);     // x[:, slice] += O Wo[slice, :]^T
                } else if (p.first_mlp_only == 2) {
                    // x[:, slice] += O Wo[slice, :]^T with O loaded by TMA (no gather); k-blocks in the order the producer loads Wo's
                    mbar_wait(o_full, 0);
                    wait_ew();  // the element-wise warps have put x into the residual columns
                    for (int hf = 0; hf < 2; ++hf) {
                        const uint32_t b = ring_wait();
#pragma unroll
                        for (int kb = 0; kb < 2; ++kb) {
                            const int a = (r + R - (hf * 2 + kb)) % R;
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma_bf16(T_RES, make_smem_desc(a_addr + a * ATOM + k * 32, 16, 1024),
                                          make_smem_desc(b + kb * 8192 + k * 32, 16, 1024), idesc_64, 1u);
                        }
                        ring_release();
                    }
                    umma_commit(mma_done);
                    lb::mbar_arrive(sa_free);  // (nothing is stored out of sA here)
                }
                slice_product(T_QKV, false, OM_MLP_H0, l, false);  // hidden[:, slice] = LN2(x) W1[slice, :]^T
                slice_product(T_RES, true, OM_MLP_G, l, true);     // x[:, slice] += G W2[slice, :]^T
            }
            lb::tma_store_wait_all();
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------ element-wise warps
        const int q = warp & 3, cg = warp >> 2;  // TMEM lane quadrant, which 32 of the CTA's 64 columns
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const int lcol = cg * 32;        // first column inside the CTA's slice
        const int gcol = r * DH + lcol;  // first channel
        const int64_t grow = (int64_t)tile * ROWS + row;
        const bool row_ok = grow < p.total_rows;
        const int n = p.n;
        const int64_t sample = grow / n;
        const int qi = (int)(grow - sample * n);
        const int kbase = (row / n) * n;  // first key column of this row's sample (warp-uniform: n >= 32)
        uint8_t* own_atom = sA + r * ATOM;
        uint8_t* slice = sQ + warp * 4096;  // 32 rows x 128 bytes of fp32 staging, private to the warp
        const int slice_row0 = tile * ROWS + q * 32;
        uint32_t mma_phase = 0, stat_phase = 0, free_phase = 0, sa_phase = 0, qkv_phase = 0;
        bool first_gather = true;
        uint32_t peer_atom[R - 1], peer_bar[R - 1];  // the CTA's atom and its gath_full barrier in the three peers
#pragma unroll
        for (int d = 1; d < R; ++d) {
            const uint32_t peer = (uint32_t)((r + d) % R);
            peer_atom[d - 1] = mapa(smem_u32(own_atom), peer);
            peer_bar[d - 1] = mapa(smem_u32(&gath_full[r]), peer);  // (a peer counts this CTA's bytes on its barrier number r)
        }

        int tr = 0;
        auto wait_mma = [&]() {
            if (threadIdx.x == 0) CL_TRACE(1, 1000 + tr);
            mbar_wait(mma_done, mma_phase);
            mma_phase ^= 1u;
            tc_fence_after();
            if (threadIdx.x == 0) CL_TRACE(1, 2000 + tr);
            ++tr;
        };
        // this CTA has seen the products that read the current gather complete: tell every CTA of the cluster
        auto free_arrive = [&]() {
            if (threadIdx.x == 0) {
                const uint32_t a = smem_u32(gath_free);
#pragma unroll
                for (int d = 0; d < R; ++d) mbar_arrive_peer(mapa(a, (uint32_t)d));
            }
        };
        auto signal = [&]() {  // everything this warp wrote (shared memory, tensor memory) is visible to the MMA thread
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) lb::mbar_arrive(ew_done);
        };
        // 32 values of this thread's row -> bf16 -> the CTA's atom of the next A operand in its own shared memory; the two
        // warps of a TMEM lane quadrant then hold 32 complete rows (4 KB), which one of their threads copies into the
        // three peers (cp.async.bulk shared::cta -> shared::cluster; the peers count the bytes on their gath_full[r]) before
        // it arrives on the local gath_full[r].  (Sending every thread's 64 bytes with st.async as they are produced was
        // slower: 25.0 instead of 18.7 us per layer.  Either way the 48 KB a CTA sends per gather leave at the ~20 bytes
        // per clock of the SM-to-SM network, which is the floor of this design: four gathers = ~9 K cycles per layer.)
        auto send_operand = [&](const float (&v)[32]) {
            if (!first_gather) {
                mbar_wait(sa_free, sa_phase);  // the saved copy of the previous operand has left the CTA's atom
                sa_phase ^= 1u;
                mbar_wait(gath_free, free_phase);  // every CTA of the cluster is done reading the previous operand
                free_phase ^= 1u;
            }
            if (threadIdx.x == 0) {
                CL_TRACE(1, 5000);
#pragma unroll
                for (int d = 1; d < R; ++d) mbar_expect_tx(&gath_full[(r + d) % R], ATOM);  // arrive + the 16 KB that peer sends
            }
            lb::store_half_row_smem(own_atom, row, cg, v);
            if (lane == 0) lb::tma_store_wait_read();  // (fp32 copies staged by this warp have left shared memory)
            tc_fence_before();
            fence_proxy_async();
            asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");  // the quadrant's two warps
            if (cg == 0 && lane == 0) {
                const uint32_t src = smem_u32(own_atom) + (uint32_t)q * 4096u;
#pragma unroll
                for (int d = 0; d < R - 1; ++d) bulk_copy_to_peer(peer_atom[d] + (uint32_t)q * 4096u, src, 4096u, peer_bar[d]);
                lb::mbar_arrive(&gath_full[r]);
            }
            if (threadIdx.x == 0) CL_TRACE(1, 5001);
            first_gather = false;
        };
        // 32 fp32 per row -> the warp's slice -> TMA store; an earlier store out of the slice must have read it
        auto stage_store_f32 = [&](int which, int col0, int l, const float* v) {
            if (lane == 0) lb::tma_store_wait_read();
            __syncwarp();
            uint8_t* prow = slice + lane * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(prow + ((j ^ (lane & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                lb::tma_store_3d(&p.out[which], slice, col0, slice_row0, l);
                lb::tma_store_commit();
            }
        };
        // LayerNorm over the cluster: every thread sends the (mean, M2) of its 32 columns to all four CTAs ...
        auto send_stats = [&](const Stat& st) {
            const uint32_t mine = smem_u32(&sStat[(r * 2 + cg) * ROWS + row]), sbar = smem_u32(stat_full);
            if (threadIdx.x == 0) mbar_expect_tx(stat_full, R * EW_THREADS * 8);  // 8 bytes from every element-wise thread of the cluster
#pragma unroll
            for (int d = 0; d < R; ++d) st_async_f32x2(mapa(mine, (uint32_t)d), st.mean, st.m2, mapa(sbar, (uint32_t)d));
        };
        // ... and, once all eight partials of its row are there, normalises its 32 values and sends them on as bf16
        auto ln_finish = [&](float (&v)[32], const float* gamma, const float* beta, float* stats_out) {
            if (threadIdx.x == 0) CL_TRACE(1, 4000);
            mbar_wait(stat_full, stat_phase);
            stat_phase ^= 1u;
            if (threadIdx.x == 0) CL_TRACE(1, 4001);
            Stat t{0.f, 0.f, 0.f};
#pragma unroll
            for (int g = 0; g < 2 * R; ++g) {  // same order in every thread of every CTA: all agree bit for bit
                const float2 o = sStat[g * ROWS + row];
                if (g == 0) {
                    t.n = 32.f; t.mean = o.x; t.m2 = o.y;
                } else {
                    lb::stat_merge(t, 32.f, o.x, o.y);
                }
            }
            const float mean = t.mean;
            const float rstd = rsqrtf(t.m2 * (1.0f / C) + LN_EPS);
            if (r == 0 && cg == 0 && row_ok) *reinterpret_cast<float2*>(stats_out + 2 * grow) = make_float2(mean, rstd);
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 g = *reinterpret_cast<const float4*>(gamma + lcol + d);
                const float4 b = *reinterpret_cast<const float4*>(beta + lcol + d);
                v[d] = fmaf((v[d] - mean) * rstd, g.x, b.x);
                v[d + 1] = fmaf((v[d + 1] - mean) * rstd, g.y, b.y);
                v[d + 2] = fmaf((v[d + 2] - mean) * rstd, g.z, b.z);
                v[d + 3] = fmaf((v[d + 3] - mean) * rstd, g.w, b.w);
            }
            send_operand(v);
        };

        // this CTA's 64-float slices of the layer's ten vectors, double-buffered in shared memory: layer l + 1's are
        // loaded into registers at the start of layer l and written to the other buffer in its Q/K/V phase (every
        // thread has left layer l - 1 by then: the gather it waited for needed all of them)
        float pre[3];
        auto prefetch_params = [&](int l) {
            const LayerPtrs& Lp = p.layer[l];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int e = k * EW_THREADS + (int)threadIdx.x;
                if (e < PAR_FLOATS) pre[k] = __ldg(Lp.par[e >> 6] + r * DH + (e & 63));
            }
        };
        auto commit_params = [&](int l) {
            float* dst = sPar + (l & 1) * PAR_FLOATS;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int e = k * EW_THREADS + (int)threadIdx.x;
                if (e < PAR_FLOATS) dst[e] = pre[k];
            }
        };
        prefetch_params(0);

        // ---- layer 0 input: global -> tensor memory, statistics on the way
        float v[32];
        Stat st{0.f, 0.f, 0.f};
        if (row_ok) {
            const float4* src = reinterpret_cast<const float4*>(p.x + grow * C + gcol);
#pragma unroll
            for (int d = 0; d < 8; ++d) {
                const float4 t = src[d];
                v[4 * d] = t.x; v[4 * d + 1] = t.y; v[4 * d + 2] = t.z; v[4 * d + 3] = t.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
        }
        if (p.first_mlp_only != 2) {  // (mode 2: the first LayerNorm sees x + O Wo^T + bo, formed below)
            lb::stat_chunk(st, v);
            send_stats(st);
        } else {
            first_gather = false;  // sA holds the attention output: the first gather waits until every CTA's out-projection has read it
        }
        lb::st_chunk(T_RES + lane_off + (uint32_t)lcol, v);
        tmem_st_wait();
        commit_params(0);
        bar_ew();  // the parameters are in shared memory before anyone reads them
        if (p.first_mlp_only == 2) signal();  // the out-projection accumulates into the residual columns just written

        for (int l = 0; l < p.layers; ++l) {
            const LayerPtrs& L = p.layer[l];
            const float* par = sPar + (l & 1) * PAR_FLOATS;
            const bool last = (l + 1 == p.layers);
            if (!last) prefetch_params(l + 1);
            const bool mlp_only = (l == 0 && p.first_mlp_only);
            if (mlp_only && p.first_mlp_only == 1) {
                // the MLP of the cross-attention layer in front of the block: x is the attention half's output already
                stage_store_f32(OM_XMID, gcol, l, v);  // (the backward pass reads the layer's x_mid)
            } else {
            if (!mlp_only) {
            ln_finish(v, par + PAR_LN1_W * 64, par + PAR_LN1_B * 64, L.sa_stats);

            // ---- [Q_r | K_r | V_r] + bias -> bf16 operand tiles: six 32-column chunks, three per column half
            wait_mma();
            free_arrive();
            if (!last) commit_params(l + 1);
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const int chunk = cg * 3 + j;
                float t[32];
                lb::ld_chunk(T_QKV + lane_off + (uint32_t)(chunk * 32), t);
                lb::add_vec32(t, par + PAR_BQ * 64 + chunk * 32);
                lb::store_half_row_smem(sQ + (chunk >> 1) * ATOM, row, chunk & 1, t);
            }
            signal();

            // ---- softmax over the row's own sample (column half 0 warps)
            wait_mma();
            if (cg == 0) {
                float mx = -INFINITY;
                for (int c = 0; c < (n >> 5); ++c) {
                    float t[32];
                    lb::ld_chunk(T_S + lane_off + (uint32_t)(kbase + c * 32), t);
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
                            lb::ld_chunk(T_S + lane_off + (uint32_t)key0, t);
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                t[i] = lb::ex2(fmaf(t[i], p.scale_log2e, -m2));
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
                    tmem_st32(T_P + (uint32_t)(g * 32) + lane_off, pw);
                }
                tmem_st_wait();
                sInv[row] = 1.0f / lsum;
                if (row_ok) L.lse[(sample * H + r) * n + qi] = (m2 + log2f(lsum)) * 0.69314718055994530942f;
            }
            signal();

            // ---- O_r / l -> the CTA's atom of the out-projection's A operand
            wait_mma();
            {
                const float inv_l = sInv[row];  // (written before the arrive that let the P V product start)
                float t[32];
                lb::ld_chunk(T_O + lane_off + (uint32_t)lcol, t);
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] *= inv_l;
                send_operand(t);
            }

            }  // (mode 2 starts here: the out-projection of the given attention output has accumulated into the residual columns)
            // ---- x_mid = x + attention + bo (back into the residual columns, fp32 copy out), LayerNorm 2
            wait_mma();
            free_arrive();
            st = Stat{0.f, 0.f, 0.f};
            lb::ld_chunk(T_RES + lane_off + (uint32_t)lcol, v);
            lb::add_vec32(v, par + PAR_BO * 64 + lcol);
            lb::stat_chunk(st, v);
            send_stats(st);  // first: the peers wait for these
            lb::st_chunk(T_RES + lane_off + (uint32_t)lcol, v);
            if (!mlp_only) {
                mbar_wait(qkv_free, qkv_phase);  // Q, K, V have left their tiles: sQ / sK are staging space, sV takes GELU'
                qkv_phase ^= 1u;
            }
            stage_store_f32(OM_XMID, gcol, l, v);
            tmem_st_wait();
            }
            ln_finish(v, par + PAR_LN2_W * 64, par + PAR_LN2_B * 64, L.mlp_stats);

            // ---- GELU(hidden + b1) -> the CTA's atom of sA, GELU' -> sV; both leave by TMA
            wait_mma();
            free_arrive();
            if (mlp_only && !last) commit_params(l + 1);
            {
                float t[32], dg[32];
                lb::ld_chunk(T_QKV + lane_off + (uint32_t)lcol, t);
                lb::add_vec32(t, par + PAR_B1 * 64 + lcol);
#pragma unroll
                for (int i = 0; i < 32; ++i) gelu_erf_both(t[i], t[i], dg[i]);
                lb::store_half_row_smem(sV, row, cg, dg);
                send_operand(t);
            }

            // ---- x_next = x_mid + mlp + b2: fp32 copy out; stays in the residual columns (and registers) for the next layer
            wait_mma();
            free_arrive();
            st = Stat{0.f, 0.f, 0.f};
            lb::ld_chunk(T_RES + lane_off + (uint32_t)lcol, v);
            lb::add_vec32(v, par + PAR_B2 * 64 + lcol);
            if (!last) {
                lb::stat_chunk(st, v);
                send_stats(st);
                lb::st_chunk(T_RES + lane_off + (uint32_t)lcol, v);
            }
            stage_store_f32(last ? OM_Y : OM_XNEXT, gcol, last ? 0 : l, v);
            if (!last) tmem_st_wait();
        }
        if (lane == 0) lb::tma_store_wait_all();
        __syncwarp();
        tc_fence_before();
    }
    __syncthreads();
    cluster_sync_all();  // no CTA leaves while a peer may still arrive on its barriers or copy into its shared memory
    if (warp == EW_WARPS + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace lbc

#ifdef PIO_LB_TRACE
extern "C" int pio_debug_lbc_trace(long long* out, int max_pairs, int who, int reset) {
    int n = 0;
    cudaMemcpyFromSymbol(&n, lbc::gc_trace_n, sizeof(int), who * sizeof(int));
    n = std::min(n, max_pairs);
    if (out && n > 0) {
        cudaMemcpyFromSymbol(out, lbc::gc_trace, (size_t)n * 8, (size_t)(who * 2 + 0) * 8192 * 8);
        cudaMemcpyFromSymbol(out + max_pairs, lbc::gc_trace, (size_t)n * 8, (size_t)(who * 2 + 1) * 8192 * 8);
    }
    if (reset) {
        int z[2] = {0, 0};
        cudaMemcpyToSymbol(lbc::gc_trace_n, z, sizeof(z));
    }
    return n;
}
#endif

// Which fused kernel runs a block of `rows` latent rows.  The cluster kernel spends four SMs per 128-row tile and
// takes about the same time per layer for any number of tiles up to one wave (the GPU places 33 clusters of four:
// same-GPC constraint), so it is the choice while the tiles fit two waves; beyond that the one-CTA-per-tile kernel,
// whose 148 tiles per wave cost one layer time, catches up.  PIO_LATENT_CLUSTER = 0 / 1 forces the choice.
bool latent_cluster_preferred(int64_t rows) {
    if (const char* e = getenv("PIO_LATENT_CLUSTER")) return atoi(e) != 0;
    return cdiv(rows, (int64_t)lbc::ROWS) <= 66;
}

int latent_block_fwd_cluster(const float* x, const LatentFusedLayer* layers, int nlayers, int batch, int n, cudaStream_t st) {
    PIO_CHECK_ARG(x && layers && nlayers >= 1 && nlayers <= PIO_LATENT_FUSED_MAX_LAYERS, "latent_block_fwd_cluster: bad arguments");
    const int64_t rows = (int64_t)batch * n;
    if (rows == 0) return PIO_OK;
    using namespace lbc;
    Params p{};
    // a first layer without an attention half of its own (see latent_block_fwd in pio_ops_bf16.cu): 1 = MLP only, 2 = the
    // out-projection of a given attention output (pre_o16) + residual, then the MLP
    const int first_mlp_only = layers[0].wqkv != nullptr ? 0 : (layers[0].wo != nullptr ? 2 : 1);
    PIO_CHECK_ARG(first_mlp_only != 2 || (layers[0].pre_o16 != nullptr && layers[0].bo != nullptr),
                  "latent_block_fwd_cluster: an out-projection first layer needs the attention output and the bias");
    for (int l = 0; l < nlayers; ++l) {
        const LatentFusedLayer& s = layers[l];
        PIO_CHECK_ARG(s.wqkv != nullptr || l == 0, "latent_block_fwd_cluster: only the first layer may be MLP-only");
        PIO_TRY(make_tmap_bf16(&p.maps[l][0], s.wqkv ? s.wqkv : s.w1, s.wqkv ? 3 * C : C, C, C, 64));
        PIO_TRY(make_tmap_bf16(&p.maps[l][1], s.wo ? s.wo : s.w1, C, C, C, 64));
        PIO_TRY(make_tmap_bf16(&p.maps[l][2], s.w1, C, C, C, 64));
        PIO_TRY(make_tmap_bf16(&p.maps[l][3], s.w2, C, C, C, 64));
        LayerPtrs& d = p.layer[l];
        d.par[PAR_LN1_W] = s.ln1_w; d.par[PAR_LN1_B] = s.ln1_b;
        d.par[PAR_BQ] = s.bqkv; d.par[PAR_BK] = s.bqkv + C; d.par[PAR_BV] = s.bqkv + 2 * C;
        d.par[PAR_BO] = s.bo; d.par[PAR_LN2_W] = s.ln2_w; d.par[PAR_LN2_B] = s.ln2_b;
        d.par[PAR_B1] = s.b1; d.par[PAR_B2] = s.b2;
        if (l == 0 && first_mlp_only)
            for (int k = 0; k < PAR_LN2_W; ++k)
                if (!(k == PAR_BO && first_mlp_only == 2)) d.par[k] = s.ln2_w;  // (loaded with the rest, never used: any readable 256 floats)
        d.sa_stats = s.sa_stats; d.mlp_stats = s.mlp_stats; d.lse = s.lse;
    }
    const LatentFusedLayer& f0 = layers[0];
    const LatentFusedLayer& f1 = layers[nlayers > 1 ? 1 : 0];
    auto stride_of = [&](const void* a0, const void* a1) -> int64_t {
        return nlayers > 1 ? (int64_t)((const char*)a1 - (const char*)a0) : (int64_t)256;
    };
    for (int l = 2; l < nlayers; ++l)  // the layers' slabs must be equally spaced (they are: BlockLayout::per_layer)
        PIO_CHECK_ARG((const char*)layers[l].xmid - (const char*)layers[l - 1].xmid == (const char*)f1.xmid - (const char*)f0.xmid,
                      "latent_block_fwd_cluster: saved slabs are not equally spaced");
    PIO_TRY(make_out_map(&p.out[OM_SA_H0], false, f0.sa_h0, C, rows, nlayers, C * 2, stride_of(f0.sa_h0, f1.sa_h0), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_QKV], false, f0.qkv, 3 * C, rows, nlayers, 3 * C * 2, stride_of(f0.qkv, f1.qkv), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_O16], false, f0.o16, C, rows, nlayers, C * 2, stride_of(f0.o16, f1.o16), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_MLP_H0], false, f0.mlp_h0, C, rows, nlayers, C * 2, stride_of(f0.mlp_h0, f1.mlp_h0), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_MLP_G], false, f0.mlp_g, C, rows, nlayers, C * 2, stride_of(f0.mlp_g, f1.mlp_g), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_MLP_H1], false, f0.mlp_h1, C, rows, nlayers, C * 2, stride_of(f0.mlp_h1, f1.mlp_h1), 64, ROWS));
    PIO_TRY(make_out_map(&p.out[OM_XMID], true, f0.xmid, C, rows, nlayers, C * 4, stride_of(f0.xmid, f1.xmid), 32, 32));
    if (nlayers > 1)  // (the last layer's output is the OM_Y map; the slabs are spaced like every other saved tensor)
        PIO_TRY(make_out_map(&p.out[OM_XNEXT], true, f0.xout, C, rows, nlayers - 1, C * 4, stride_of(f0.xmid, f1.xmid), 32, 32));
    PIO_TRY(make_out_map(&p.out[OM_Y], true, layers[nlayers - 1].xout, C, rows, 1, C * 4, 256, 32, 32));
    p.x = x;
    p.layers = nlayers;
    p.n = n;
    p.total_rows = rows;
    p.scale_log2e = (1.0f / sqrtf((float)DH)) * 1.44269504088896340736f;
    p.first_mlp_only = first_mlp_only;
    if (first_mlp_only == 2) PIO_TRY(make_out_map(&p.o_map, false, layers[0].pre_o16, C, rows, 1, C * 2, 256, 64, ROWS));
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(latent_block_fwd_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    const unsigned grid = (unsigned)(cdiv(rows, (int64_t)ROWS) * R);
    PIO_CUDA(launch_pdl(latent_block_fwd_cluster_kernel, dim3(grid), dim3(THREADS), SMEM_BYTES, st, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
