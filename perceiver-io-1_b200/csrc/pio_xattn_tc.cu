// Tensor-core backward of the encoder's cross-attention when all queries of a sample fit one tile (the latent array:
// n <= 128 queries against m keys), head size 64.
//
// Reference: CrossAttention of the encoder layers (perceiver/model.py:77-99, applied :186-191) — the n latents of a
// sample attend to its m input positions.  The general kernels (pio_attn_tc.cu: one kernel owns query tiles and streams
// the keys for dQ, a second owns key tiles and streams the queries for dK / dV, both recomputing S and dP) read K and V
// twice and run seven products; with every query of a (sample, head) in ONE tile nothing has to be recomputed or
// exchanged: a CTA owns one (sample, head), keeps Q_h and dO_h (n x 64, zero rows up to 128 by TMA out-of-bounds fill)
// in shared memory and streams the keys 128 at a time through a two-stage ring:
//   * per key tile: S = Q K^T and dP = dO V^T on tcgen05 (N = 128); thread = query row: p = exp(s / 8 - lse),
//     dS = p (dP - delta) / 8 with delta = dO . O formed once per CTA; P and dS leave as bf16 tiles that serve
//     dQ += dS K as the K-major A operand and dK = dS^T Q, dV = P^T dO as MN-major A operands;
//   * dK / dV of the tile are complete (every query is here) and go straight to global memory — or are added to what
//     is there (dkv_add: the second consumer of a shared kv); dQ accumulates in tensor memory across the tiles and
//     leaves once;
//   * the tensor pipe runs in order, so S and dP of tile j + 1 are issued IN FRONT of the gradient products of tile j
//     (as soon as tile j's S and dP sit in registers): the element-wise warps turn them into P and dS while those
//     products run, and per tile only "drain dK / dV, store P / dS" remains between two batches of products.
// Five products instead of seven, K and V read once, no delta pass and no second kernel.
#include <cuda.h>

#include <algorithm>
#include <atomic>

#include "pio_latent_common.cuh"

namespace pio {

namespace xat {

using lb::bf16;

constexpr int DH = 64;
constexpr int ROWS = 128;
constexpr int ATOM = ROWS * 128;
constexpr int EW_WARPS = 8;
constexpr int THREADS = EW_WARPS * 32 + 64;
constexpr int MAX_TILES = 64;  // keys per sample <= 8192
constexpr int STAGES = 2;      // key / value ring (three stages measured slower: 51 vs 45 us at the benchmark shape)
constexpr size_t SMEM_BYTES = (6 + 2 * STAGES) * ATOM + 2 * ROWS * sizeof(float) + MAX_TILES * 4 * sizeof(uint32_t) + 16 * sizeof(uint64_t) + 1024;

struct Params {
    const float* lse;     // (batch, heads, n), natural log
    const uint8_t* pad;   // (batch, m) or null
    const bf16* o16;      // forward output (batch n, e) bf16, or null: o32
    const float* o32;
    const bf16* do16;     // (batch n, e)
    bf16* dq16;           // (batch n, e)
    bf16* dk16;           // strides of k
    bf16* dv16;
    int64_t o_rs, o_bs, q_rs, q_bs, k_rs, k_bs, v_rs, v_bs;
    int n, m, heads;
    int dkv_add;
    float scale, scale_log2e, log2e;
};

// Phase timestamps of CTA (0, 0) for tuning (only in builds with -DPIO_LB_TRACE, tools/xattn_trace.py): slot = 8 * tile + k,
// k = 0..4 the element-wise warp 0 (S ready, S in registers, products of the tile before done, dK / dV drained, P / dS
// stored), k = 5..7 the MMA thread (S of the next tile issued, P / dS seen, products issued); the last 8 slots: kernel
// entry, prologue done, kernel end.
#ifdef PIO_LB_TRACE
__device__ long long g_xat_trace[8 * (MAX_TILES + 1)];
#define XT(slot)                                                                          \
    do {                                                                                  \
        if (blockIdx.x == 0 && blockIdx.y == 0) g_xat_trace[slot] = (long long)clock64(); \
    } while (0)
#else
#define XT(slot) \
    do {         \
    } while (0)
#endif

__device__ __forceinline__ void add32_bf16_global(const bf16* src, float (&v)[32]) {
#pragma unroll
    for (int d = 0; d < 32; d += 8) {
        const uint4 w = *reinterpret_cast<const uint4*>(src + d);
        const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int e2 = 0; e2 < 4; ++e2) {
            const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ws[e2]));
            v[d + 2 * e2] += f.x;
            v[d + 2 * e2 + 1] += f.y;
        }
    }
}

// 32 bf16 (16 packed pairs) into row `row`, 64-byte half `hh` of a 128-row x 128-byte swizzled (SW128, K-major) atom
__device__ __forceinline__ void store_half_row_packed(uint8_t* atom, int row, int hh, const uint32_t (&w)[16]) {
    uint8_t* prow = atom + row * 128;
#pragma unroll
    for (int piece = 0; piece < 4; ++piece) {
        const int ci = hh * 4 + piece;
        *reinterpret_cast<uint4*>(prow + ((ci ^ (row & 7)) << 4)) = make_uint4(w[piece * 4], w[piece * 4 + 1], w[piece * 4 + 2], w[piece * 4 + 3]);
    }
}

__global__ void __launch_bounds__(THREADS, 1)
    xattn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                        const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sDO = sQ + ATOM;
    uint8_t* sK = sDO + ATOM;            // [STAGES]
    uint8_t* sV = sK + STAGES * ATOM;    // [STAGES]
    uint8_t* sP = sV + STAGES * ATOM;    // [128 queries][128 keys] bf16 as two 64-key atoms
    uint8_t* sDS = sP + 2 * ATOM;
    float* sDelta = (float*)(sDS + 2 * ATOM);          // [2 column halves][128 rows]
    uint32_t* sMask = (uint32_t*)(sDelta + 2 * ROWS);  // [tile][4 chunks of 32 keys]: bit = key takes part
    uint64_t* bars = (uint64_t*)(sMask + MAX_TILES * 4);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;   // [STAGES]
    uint64_t* kv_free = bars + 4;   // [STAGES]
    uint64_t* s_done = bars + 7;    // S and dP of a tile are in tensor memory
    uint64_t* g_done = bars + 8;    // dQ, dK, dV products of a tile have run
    uint64_t* ew_p = bars + 9;      // P, dS of a tile written, dK, dV of the tile before it drained
    uint64_t* s_free = bars + 10;   // S, dP of a tile read into registers
    uint32_t* tmem_slot = (uint32_t*)(bars + 11);

    pdl_trigger();
    if (threadIdx.x == 0) XT(8 * MAX_TILES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, b = blockIdx.y;
    const int n = p.n, m = p.m;
    const int tiles = (m + ROWS - 1) / ROWS;
    const int qsteps = (n + 15) / 16;  // 16-query steps of the dK / dV products (query rows past them are never read)

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmDO);
        prefetch_tmap(&tmK);
        prefetch_tmap(&tmV);
        mbar_init(q_full, 1);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(kv_full + i, 1);
            mbar_init(kv_free + i, 1);
        }
        mbar_init(s_done, 1);
        mbar_init(g_done, 1);
        mbar_init(ew_p, EW_WARPS);
        mbar_init(s_free, EW_WARPS);
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
    pdl_wait();

    if (warp == EW_WARPS) {
        if (lane == 0) {
            mbar_expect_tx(q_full, 2 * ATOM);
            lb::tma_load_3d(sQ, &tmQ, h * DH, 0, b, q_full);
            lb::tma_load_3d(sDO, &tmDO, h * DH, 0, b, q_full);
            for (int j = 0; j < tiles; ++j) {
                const int st = j % STAGES;
                if (j >= STAGES) mbar_wait(kv_free + st, (uint32_t)(j / STAGES - 1) & 1u);
                mbar_expect_tx(kv_full + st, 2 * ATOM);
                lb::tma_load_3d(sK + st * ATOM, &tmK, h * DH, j * ROWS, b, kv_full + st);
                lb::tma_load_3d(sV + st * ATOM, &tmV, h * DH, j * ROWS, b, kv_full + st);
            }
        }
        __syncwarp();
    } else if (warp == EW_WARPS + 1) {
        if (lane == 0) {
            const uint32_t idesc_s = make_idesc(128, false, false);
            const uint32_t idesc_dq = make_idesc(DH, false, true);
            const uint32_t idesc_dkv = make_idesc(DH, true, true);
            const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sDO);
            const uint32_t p_addr = smem_u32(sP), ds_addr = smem_u32(sDS);
            auto issue_s = [&](int j) {  // S = Q K^T, dP = dO V^T of tile j
                const uint32_t k_addr = smem_u32(sK + (j % STAGES) * ATOM), v_addr = smem_u32(sV + (j % STAGES) * ATOM);
                mbar_wait(kv_full + j % STAGES, (uint32_t)(j / STAGES) & 1u);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(T_S, make_smem_desc(q_addr + k * 32, 16, 1024), make_smem_desc(k_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(T_P, make_smem_desc(do_addr + k * 32, 16, 1024), make_smem_desc(v_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
                umma_commit(s_done);
            };
            mbar_wait(q_full, 0);
            issue_s(0);
            for (int j = 0; j < tiles; ++j) {
                const uint32_t k_addr = smem_u32(sK + (j % STAGES) * ATOM);
                if (j + 1 < tiles) {  // S, dP of the next tile go in FRONT of this tile's gradient products: they only need
                    mbar_wait(s_free, (uint32_t)j & 1u);  // tile j's S, dP to have been read into registers,
                    issue_s(j + 1);                       // and the element-wise warps work on them while the products run
                }
                XT(8 * j + 5);
                mbar_wait(ew_p, (uint32_t)j & 1u);  // P, dS of tile j written; dK, dV of tile j - 1 drained
                tc_fence_after();
                XT(8 * j + 6);
#pragma unroll
                for (int s = 0; s < 8; ++s)  // dQ += dS K: 16 keys per step
                    umma_bf16(T_DQ, make_smem_desc(ds_addr + (s >> 2) * ATOM + (s & 3) * 32, 16, 1024), make_smem_desc(k_addr + s * 2048, 8192, 1024),
                              idesc_dq, (j > 0 || s > 0) ? 1u : 0u);
                for (int s = 0; s < qsteps; ++s)  // dK = dS^T Q: 16 queries per step, A read MN-major (two 64-key atoms 16 KB apart)
                    umma_bf16(T_DK, make_smem_desc(ds_addr + s * 2048, ATOM, 1024), make_smem_desc(q_addr + s * 2048, 8192, 1024), idesc_dkv,
                              s > 0 ? 1u : 0u);
                for (int s = 0; s < qsteps; ++s)  // dV = P^T dO
                    umma_bf16(T_DV, make_smem_desc(p_addr + s * 2048, ATOM, 1024), make_smem_desc(do_addr + s * 2048, 8192, 1024), idesc_dkv,
                              s > 0 ? 1u : 0u);
                umma_commit(g_done);
                umma_commit(kv_free + j % STAGES);
                XT(8 * j + 7);
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3, cg = warp >> 2;
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const bool row_ok = row < n;
        const bool live = q * 32 < qsteps * 16;  // (warp-uniform) this warp's query rows are read by the dK / dV products
        auto signal = [&](uint64_t* bar) {
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) lb::mbar_arrive(bar);
        };
        // which keys take part (in range, not padded), 32 per word
        for (int idx = warp; idx < tiles * 4; idx += EW_WARPS) {
            const int key = idx * 32 + lane;
            const bool ok = key < m && !(p.pad != nullptr && __ldg(p.pad + (int64_t)b * m + key) != 0);
            const uint32_t word = __ballot_sync(0xffffffffu, ok);
            if (lane == 0) sMask[idx] = word;
        }
        // delta = dO . O over the head's 64 columns (two threads per row)
        float part = 0.f;
        if (row_ok) {
            const int64_t off = (int64_t)b * p.o_bs + (int64_t)row * p.o_rs + h * DH + cg * 32;
            const bf16* dop = p.do16 + off;
#pragma unroll
            for (int d = 0; d < 32; d += 8) {
                const uint4 wd = *reinterpret_cast<const uint4*>(dop + d);
                const uint32_t a[4] = {wd.x, wd.y, wd.z, wd.w};
                float ov[8];
                if (p.o16 != nullptr) {
                    const uint4 wo = *reinterpret_cast<const uint4*>(p.o16 + off + d);
                    const uint32_t c[4] = {wo.x, wo.y, wo.z, wo.w};
#pragma unroll
                    for (int e2 = 0; e2 < 4; ++e2) {
                        const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&c[e2]));
                        ov[2 * e2] = f.x;
                        ov[2 * e2 + 1] = f.y;
                    }
                } else {
                    const float4 f0 = *reinterpret_cast<const float4*>(p.o32 + off + d);
                    const float4 f1 = *reinterpret_cast<const float4*>(p.o32 + off + d + 4);
                    ov[0] = f0.x; ov[1] = f0.y; ov[2] = f0.z; ov[3] = f0.w;
                    ov[4] = f1.x; ov[5] = f1.y; ov[6] = f1.z; ov[7] = f1.w;
                }
#pragma unroll
                for (int e2 = 0; e2 < 4; ++e2) {
                    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&a[e2]));
                    part = fmaf(f.x, ov[2 * e2], part);
                    part = fmaf(f.y, ov[2 * e2 + 1], part);
                }
            }
        }
        sDelta[cg * ROWS + row] = part;
        asm volatile("bar.sync 1, 256;" ::: "memory");  // (the mask words too)
        const float delta = sDelta[row] + sDelta[ROWS + row];
        const float lse2 = row_ok ? __ldg(p.lse + ((int64_t)b * p.heads + h) * n + row) * p.log2e : 0.f;
        if (threadIdx.x == 0) XT(8 * MAX_TILES + 1);

        auto drain = [&](int j) {  // dK_h, dV_h of tile j: rows = keys
            const int key = j * ROWS + row;
            float v[32];
            lb::ld_chunk(T_DK + lane_off + (uint32_t)(cg * 32), v);
            if (key < m) {
                bf16* out = p.dk16 + (int64_t)b * p.k_bs + (int64_t)key * p.k_rs + h * DH + cg * 32;
                if (p.dkv_add) add32_bf16_global(out, v);
                lb::store32_bf16_global(out, v);
            }
            lb::ld_chunk(T_DV + lane_off + (uint32_t)(cg * 32), v);
            if (key < m) {
                bf16* out = p.dv16 + (int64_t)b * p.v_bs + (int64_t)key * p.v_rs + h * DH + cg * 32;
                if (p.dkv_add) add32_bf16_global(out, v);
                lb::store32_bf16_global(out, v);
            }
        };

        for (int j = 0; j < tiles; ++j) {
            uint32_t pw[2][16], dw[2][16];  // this thread's 2 x 32 probabilities and score gradients, bf16 pairs
            mbar_wait(s_done, (uint32_t)j & 1u);
            tc_fence_after();
            if (threadIdx.x == 0) XT(8 * j);
            if (live) {
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int chunk = cg + 2 * cc;
                    const uint32_t word = row_ok ? sMask[j * 4 + chunk] : 0u;
                    float s[32], dp[32];
                    lb::ld_chunk(T_S + lane_off + (uint32_t)(chunk * 32), s);
                    lb::ld_chunk(T_P + lane_off + (uint32_t)(chunk * 32), dp);
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        float pv[2], dv[2];
#pragma unroll
                        for (int u = 0; u < 2; ++u) {
                            const bool ok = (word >> (i + u)) & 1u;
                            pv[u] = ok ? lb::ex2(fmaf(s[i + u], p.scale_log2e, -lse2)) : 0.f;
                            dv[u] = ok ? pv[u] * (dp[i + u] - delta) * p.scale : 0.f;
                        }
                        const __nv_bfloat162 hp = __floats2bfloat162_rn(pv[0], pv[1]), hd = __floats2bfloat162_rn(dv[0], dv[1]);
                        pw[cc][i >> 1] = *reinterpret_cast<const uint32_t*>(&hp);
                        dw[cc][i >> 1] = *reinterpret_cast<const uint32_t*>(&hd);
                    }
                }
            }
            {   // S, dP are in registers: the tensor core may overwrite them with the next tile's
                tc_fence_before();
                __syncwarp();
                if (lane == 0) lb::mbar_arrive(s_free);
            }
            if (threadIdx.x == 0) XT(8 * j + 1);
            if (j > 0) {  // the previous tile's gradient products have run: its dK, dV leave, and sP / sDS may be rewritten
                mbar_wait(g_done, (uint32_t)(j - 1) & 1u);
                tc_fence_after();
                if (threadIdx.x == 0) XT(8 * j + 2);
                drain(j - 1);
                if (threadIdx.x == 0) XT(8 * j + 3);
            }
            if (live) {
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int chunk = cg + 2 * cc;
                    store_half_row_packed(sP + (chunk >> 1) * ATOM, row, chunk & 1, pw[cc]);
                    store_half_row_packed(sDS + (chunk >> 1) * ATOM, row, chunk & 1, dw[cc]);
                }
            }
            signal(ew_p);
            if (threadIdx.x == 0) XT(8 * j + 4);
        }
        mbar_wait(g_done, (uint32_t)(tiles - 1) & 1u);
        tc_fence_after();
        drain(tiles - 1);
        {   // dQ_h, accumulated over all tiles
            float v[32];
            lb::ld_chunk(T_DQ + lane_off + (uint32_t)(cg * 32), v);
            if (row_ok) lb::store32_bf16_global(p.dq16 + (int64_t)b * p.q_bs + (int64_t)row * p.q_rs + h * DH + cg * 32, v);
        }
        tc_fence_before();
        if (threadIdx.x == 0) XT(8 * MAX_TILES + 2);
    }
    __syncthreads();
    if (warp == EW_WARPS + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace xat

#ifdef PIO_LB_TRACE
extern "C" int pio_debug_xat_trace(long long* out, int slots) {
    const int n = std::min(slots, 8 * (xat::MAX_TILES + 1));
    cudaMemcpyFromSymbol(out, xat::g_xat_trace, (size_t)n * 8);
    return n;
}
#endif

static bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

// all queries of a sample in one 128-row tile, head size 64, no attention dropout; rows of q / o / dO / dq contiguous
bool xattn_bwd_tc_supported(const pio_attn_desc* d, const DropCfg* drop) {
    const bool off = getenv("PIO_NO_XATTN_TC") != nullptr;
    if (off || d == nullptr) return false;
    if (drop != nullptr && drop->threshold != 0) return false;
    const int64_t cols = (int64_t)d->heads * d->dh;
    return d->dh == xat::DH && d->n >= 1 && d->n <= xat::ROWS && d->m >= 1 && d->m <= xat::ROWS * xat::MAX_TILES && d->batch >= 1 &&
           d->batch <= 65535 && d->heads <= 65535 && d->q_rs == cols && d->o_rs == cols && d->q_bs == (int64_t)d->n * cols &&
           d->o_bs == (int64_t)d->n * cols && d->k_rs % 8 == 0 && d->v_rs % 8 == 0 && d->k_bs % 8 == 0 && d->v_bs % 8 == 0 &&
           aligned16(d->q) && aligned16(d->k) && aligned16(d->v) && (d->o == nullptr || aligned16(d->o));
}

// q, k, v, dO (and o16 when given, else d->o fp32) as attn_bwd_tc takes them; dq16 / dk16 / dv16 with the strides of q / k / v
int xattn_bwd_tc(const pio_attn_desc* d, const __nv_bfloat16* do16, __nv_bfloat16* dq16, __nv_bfloat16* dk16, __nv_bfloat16* dv16,
                 cudaStream_t st, const __nv_bfloat16* o16, bool dkv_add) {
    using namespace xat;
    PIO_CHECK_ARG(xattn_bwd_tc_supported(d, nullptr) && do16 && dq16 && dk16 && dv16 && (o16 || d->o) && d->lse && aligned16(do16) &&
                      aligned16(dq16) && aligned16(dk16) && aligned16(dv16) && (o16 == nullptr || aligned16(o16)),
                  "xattn_bwd_tc: unsupported arguments");
    const int64_t cols = (int64_t)d->heads * d->dh;
    CUtensorMap tq, tdo, tk, tv;
    PIO_TRY(make_out_map(&tq, false, d->q, cols, d->n, d->batch, d->q_rs * 2, d->q_bs * 2, 64, ROWS));
    PIO_TRY(make_out_map(&tdo, false, do16, cols, d->n, d->batch, d->o_rs * 2, d->o_bs * 2, 64, ROWS));
    PIO_TRY(make_out_map(&tk, false, d->k, cols, d->m, d->batch, d->k_rs * 2, d->k_bs * 2, 64, ROWS));
    PIO_TRY(make_out_map(&tv, false, d->v, cols, d->m, d->batch, d->v_rs * 2, d->v_bs * 2, 64, ROWS));
    Params p{};
    p.lse = d->lse; p.pad = d->pad; p.o16 = (const bf16*)o16; p.o32 = d->o; p.do16 = (const bf16*)do16;
    p.dq16 = (bf16*)dq16; p.dk16 = (bf16*)dk16; p.dv16 = (bf16*)dv16;
    p.o_rs = d->o_rs; p.o_bs = d->o_bs; p.q_rs = d->q_rs; p.q_bs = d->q_bs;
    p.k_rs = d->k_rs; p.k_bs = d->k_bs; p.v_rs = d->v_rs; p.v_bs = d->v_bs;
    p.n = d->n; p.m = d->m; p.heads = d->heads; p.dkv_add = dkv_add ? 1 : 0;
    p.scale = d->scale;
    p.log2e = 1.44269504088896340736f;
    p.scale_log2e = d->scale * p.log2e;
    int dev = 0;
    PIO_CUDA(cudaGetDevice(&dev));
    static std::atomic<bool> configured[64];
    if (dev >= 0 && dev < 64 && !configured[dev].load()) {
        PIO_CUDA(cudaFuncSetAttribute(xattn_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    PIO_CUDA(launch_pdl(xattn_bwd_tc_kernel, dim3((unsigned)d->heads, (unsigned)d->batch), dim3(THREADS), SMEM_BYTES, st, tq, tdo, tk, tv, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
