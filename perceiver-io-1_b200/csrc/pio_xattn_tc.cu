// Tensor-core backward of the encoder's cross-attention when all queries of a sample fit one tile (the latent array:
// n <= 128 queries against m keys), head size 64.
//
// Reference: CrossAttention of the encoder layers (perceiver/model.py:77-99, applied :186-191) — the n latents of a
// sample attend to its m input positions.  The general kernels (pio_attn_tc.cu: one kernel owns query tiles and streams
// the keys for dQ, a second owns key tiles and streams the queries for dK / dV, both recomputing S and dP) pad the
// queries to the 128-row MMA tile, read K and V twice and run seven products.  With every query of a (sample, head) in
// ONE tile nothing has to be recomputed or exchanged, and the tile can be TRANSPOSED: the 128 keys of a tile sit on the
// MMA's M axis (the tensor-memory lanes, all of them real rows) and the queries on N (64 or 128 columns), so the
// element-wise work is spread over all 256 threads whatever n is.  A CTA owns one (sample, head), keeps Q_h and dO_h
// in shared memory and streams the keys 128 at a time through a two-stage ring:
//   * per key tile: S^T = K Q^T and dP^T = V dO^T on tcgen05 (N = 64 / 128); thread = key: p = exp(s / 8 - lse_q),
//     dS = p (dP - delta_q) / 8 with delta = dO . O formed once per CTA (per-query values come from shared memory as
//     broadcast reads); P^T and dS^T leave as bf16 [key][query] tiles: the K-major A operand of dK = dS^T Q and
//     dV = P^T dO, and the MN-major A operand of dQ += dS K;
//   * dK / dV of the tile are complete (every query is here) and go straight to global memory — or are added to what
//     is there (dkv_add: the second consumer of a shared kv); dQ accumulates in tensor memory across the tiles and
//     leaves once;
//   * the tensor pipe runs in order, so S^T and dP^T of tile j + 1 are issued IN FRONT of the gradient products of
//     tile j (as soon as tile j's sit in registers): the element-wise warps turn them into P^T and dS^T while those
//     products run, and per tile only "drain dK / dV, store P^T / dS^T" remains between two batches of products.
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
constexpr size_t SMEM_BYTES = 12 * ATOM + 4 * ROWS * sizeof(float) + MAX_TILES * ROWS + 16 * sizeof(uint64_t) + 1024;

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

// 32 bf16 (16 packed pairs) into row `row`, 64-byte half `hh` of a 128-row x 128-byte swizzled (SW128, K-major) atom
__device__ __forceinline__ void store_half_row_packed(uint8_t* atom, int row, int hh, const uint32_t (&w)[16]) {
    uint8_t* prow = atom + row * 128;
#pragma unroll
    for (int piece = 0; piece < 4; ++piece) {
        const int ci = hh * 4 + piece;
        *reinterpret_cast<uint4*>(prow + ((ci ^ (row & 7)) << 4)) = make_uint4(w[piece * 4], w[piece * 4 + 1], w[piece * 4 + 2], w[piece * 4 + 3]);
    }
}

// mbarrier wait with a watchdog: every wait in this kernel is for work that takes microseconds, so two seconds of spinning
// can only be a protocol fault — trap (the launch fails with an error) instead of hanging the GPU
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try(bar, parity)) return;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (!mbar_try(bar, parity)) {
        if ((++spins & 1023u) == 0 && clock64() - t0 > 4000000000ll) __trap();
    }
}

__device__ __forceinline__ void pack32(const float (&v)[32], uint32_t (&w)[16]) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&h2);
    }
}
// 8 bf16 + 8 bf16, summed in fp32
__device__ __forceinline__ uint4 add8_bf16(const uint4 a, const uint4 b) {
    const uint32_t x[4] = {a.x, a.y, a.z, a.w}, y[4] = {b.x, b.y, b.z, b.w};
    uint32_t r[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&x[i]));
        const float2 g = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&y[i]));
        const __nv_bfloat162 h2 = __floats2bfloat162_rn(f.x + g.x, f.y + g.y);
        r[i] = *reinterpret_cast<const uint32_t*>(&h2);
    }
    return make_uint4(r[0], r[1], r[2], r[3]);
}

// NQ = 64 or 128: the query axis of the score tiles (n rounded up; query rows past n are zero-filled by TMA)
template <int NQ>
__global__ void __launch_bounds__(THREADS, 1)
    xattn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                        const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const Params p) {
    constexpr int CH = NQ / 64;  // 32-query chunks per thread
    // key / value ring: with two stages the load of tile j + 1 is issued when the products of tile j - 1 end and its
    // ~1.4 us of HBM latency showed in front of the products of tile j; the third stage fits where the second query atoms
    // of P^T / dS^T are not needed
    constexpr int STAGES = NQ == 64 ? 3 : 2;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;                  // [queries][64] (rows >= n zero)
    uint8_t* sDO = sQ + ATOM;
    uint8_t* sK = sDO + ATOM;            // [STAGES][128 keys][64]
    uint8_t* sV = sK + STAGES * ATOM;    // [STAGES]
    uint8_t* sPT = sV + STAGES * ATOM;   // P^T: [128 keys][queries] bf16, one 64-query atom or two
    uint8_t* sDST = sPT + CH * ATOM;     // dS^T; the dQ product reads TWO atoms (M = 128 queries): for NQ = 64 the second is
    uint8_t* sStK = sDST + CH * ATOM;    // whatever the staging tile holds — rows of dQ past n, which nobody stores
    uint8_t* sStV = sStK + ATOM;         // dK, dV of a tile on their way out: [128 keys][64] bf16
    float* sPart = (float*)(sStV + ATOM);  // [2 column halves][128 queries]: partial dO . O
    float* sDelta = sPart + 2 * ROWS;          // [128 queries]
    float* sLse2 = sDelta + ROWS;              // [128 queries]: lse * log2(e); +inf past n (their probabilities come out 0)
    uint8_t* sOk = (uint8_t*)(sLse2 + ROWS);   // [tiles * 128]: the key takes part (in range, not padded)
    uint64_t* bars = (uint64_t*)(sOk + MAX_TILES * ROWS);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;   // [STAGES]
    uint64_t* kv_free = bars + 4;   // [STAGES]
    uint64_t* s_done = bars + 7;    // S^T and dP^T of a tile are in tensor memory
    uint64_t* g_done = bars + 8;    // dQ, dK, dV products of a tile have run
    uint64_t* ew_p = bars + 9;      // P^T, dS^T of a tile written, dK, dV of the tile before it drained
    uint64_t* s_free = bars + 10;   // S^T, dP^T of a tile read into registers
    uint32_t* tmem_slot = (uint32_t*)(bars + 11);

    pdl_trigger();
    if (threadIdx.x == 0) XT(8 * MAX_TILES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, b = blockIdx.y;
    const int n = p.n, m = p.m;
    const int tiles = (m + ROWS - 1) / ROWS;
    const int qsteps = (n + 15) / 16;  // 16-query k-steps of the dK / dV products (P^T, dS^T columns past them are never read)

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
    const uint32_t T_S = tmem_base;         // S^T: lanes = keys, NQ query columns
    const uint32_t T_P = tmem_base + 128;   // dP^T
    const uint32_t T_DQ = tmem_base + 256;  // dQ (lanes = queries) | dK | dV (lanes = keys): 64 columns each
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
                if (j >= STAGES) wait(kv_free + st, (uint32_t)(j / STAGES - 1) & 1u);
                mbar_expect_tx(kv_full + st, 2 * ATOM);
                lb::tma_load_3d(sK + st * ATOM, &tmK, h * DH, j * ROWS, b, kv_full + st);
                lb::tma_load_3d(sV + st * ATOM, &tmV, h * DH, j * ROWS, b, kv_full + st);
            }
        }
        __syncwarp();
    } else if (warp == EW_WARPS + 1) {
        if (lane == 0) {
            const uint32_t idesc_s = make_idesc(NQ, false, false);
            const uint32_t idesc_kv = make_idesc(DH, false, true);  // dK, dV: A = P^T / dS^T K-major, B = dO / Q read MN-major
            const uint32_t idesc_dq = make_idesc(DH, true, true);   // dQ: A = dS^T read MN-major, B = K read MN-major
            const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sDO);
            const uint32_t pt_addr = smem_u32(sPT), dst_addr = smem_u32(sDST);
            auto issue_s = [&](int j) {  // S^T = K Q^T, dP^T = V dO^T of tile j
                const uint32_t k_addr = smem_u32(sK + (j % STAGES) * ATOM), v_addr = smem_u32(sV + (j % STAGES) * ATOM);
                wait(kv_full + j % STAGES, (uint32_t)(j / STAGES) & 1u);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(T_S, make_smem_desc(k_addr + k * 32, 16, 1024), make_smem_desc(q_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(T_P, make_smem_desc(v_addr + k * 32, 16, 1024), make_smem_desc(do_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
                umma_commit(s_done);
            };
            wait(q_full, 0);
            issue_s(0);
            for (int j = 0; j < tiles; ++j) {
                const uint32_t k_addr = smem_u32(sK + (j % STAGES) * ATOM);
                if (j + 1 < tiles) {  // S^T, dP^T of the next tile go in FRONT of this tile's gradient products: they only need
                    wait(s_free, (uint32_t)j & 1u);  // tile j's to have been read into registers, and the
                    issue_s(j + 1);                       // element-wise warps work on them while the products run
                }
                XT(8 * j + 5);
                wait(ew_p, (uint32_t)j & 1u);  // P^T, dS^T of tile j written; dK, dV of tile j - 1 drained
                tc_fence_after();
                XT(8 * j + 6);
                for (int s = 0; s < qsteps; ++s)  // dK = dS^T Q: 16 queries per step
                    umma_bf16(T_DK, make_smem_desc(dst_addr + (s >> 2) * ATOM + (s & 3) * 32, 16, 1024), make_smem_desc(q_addr + s * 2048, 8192, 1024),
                              idesc_kv, s > 0 ? 1u : 0u);
                for (int s = 0; s < qsteps; ++s)  // dV = P^T dO
                    umma_bf16(T_DV, make_smem_desc(pt_addr + (s >> 2) * ATOM + (s & 3) * 32, 16, 1024), make_smem_desc(do_addr + s * 2048, 8192, 1024),
                              idesc_kv, s > 0 ? 1u : 0u);
#pragma unroll
                for (int s = 0; s < 8; ++s)  // dQ += dS K: 16 keys per step, A = dS^T read MN-major (two 64-query atoms 16 KB apart)
                    umma_bf16(T_DQ, make_smem_desc(dst_addr + s * 2048, ATOM, 1024), make_smem_desc(k_addr + s * 2048, 8192, 1024), idesc_dq,
                              (j > 0 || s > 0) ? 1u : 0u);
                umma_commit(g_done);
                umma_commit(kv_free + j % STAGES);
                XT(8 * j + 7);
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3, cg = warp >> 2;
        const int row = q * 32 + lane;  // a query in the prologue and for dQ, a key of the tile everywhere else
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        auto signal = [&](uint64_t* bar) {
            tc_fence_before();
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) lb::mbar_arrive(bar);
        };
        for (int key = threadIdx.x; key < tiles * ROWS; key += EW_WARPS * 32)
            sOk[key] = (key < m && !(p.pad != nullptr && __ldg(p.pad + (int64_t)b * m + key) != 0)) ? 1 : 0;
        // delta = dO . O over the head's 64 columns (two threads per query)
        float part = 0.f;
        if (row < n) {
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
        sPart[cg * ROWS + row] = part;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (cg == 0) {
            sDelta[row] = sPart[row] + sPart[ROWS + row];
            sLse2[row] = row < n ? __ldg(p.lse + ((int64_t)b * p.heads + h) * n + row) * p.log2e : __int_as_float(0x7f800000);
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (threadIdx.x == 0) XT(8 * MAX_TILES + 1);

        // dK_h, dV_h of tile j (lanes = keys) leave through a staging tile: a thread's 64 bytes of a 1 KB-strided row are
        // 32 separate lines per store instruction (measured 1.25 us per tile in the load / store unit); staged, a warp writes
        // 4 whole 128-byte rows per instruction
        auto drain = [&](int j) {
            float v[32];
            uint32_t w[16];
            lb::ld_chunk(T_DK + lane_off + (uint32_t)(cg * 32), v);
            pack32(v, w);
            store_half_row_packed(sStK, row, cg, w);
            lb::ld_chunk(T_DV + lane_off + (uint32_t)(cg * 32), v);
            pack32(v, w);
            store_half_row_packed(sStV, row, cg, w);
            asm volatile("bar.sync 1, 256;" ::: "memory");
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int t = it * (EW_WARPS * 32) + (int)threadIdx.x;
                const int r = t >> 3, ci = t & 7;
                const int key = j * ROWS + r;
                if (key < m) {
                    const int so = r * 128 + ((ci ^ (r & 7)) << 4);
                    uint4 xk = *reinterpret_cast<const uint4*>(sStK + so), xv = *reinterpret_cast<const uint4*>(sStV + so);
                    bf16* ok = p.dk16 + (int64_t)b * p.k_bs + (int64_t)key * p.k_rs + h * DH + ci * 8;
                    bf16* ov = p.dv16 + (int64_t)b * p.v_bs + (int64_t)key * p.v_rs + h * DH + ci * 8;
                    if (p.dkv_add) {
                        xk = add8_bf16(xk, *reinterpret_cast<const uint4*>(ok));
                        xv = add8_bf16(xv, *reinterpret_cast<const uint4*>(ov));
                    }
                    *reinterpret_cast<uint4*>(ok) = xk;
                    *reinterpret_cast<uint4*>(ov) = xv;
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");  // (the staging tiles are rewritten by the next drain)
        };

        for (int j = 0; j < tiles; ++j) {
            uint32_t pw[CH][16], dw[CH][16];  // this thread's key against CH x 32 queries: probabilities and score gradients, bf16 pairs
            const bool key_ok = sOk[j * ROWS + row] != 0;
            wait(s_done, (uint32_t)j & 1u);
            tc_fence_after();
            if (threadIdx.x == 0) XT(8 * j);
#pragma unroll
            for (int cc = 0; cc < CH; ++cc) {
                const int chunk = cg + 2 * cc;  // queries chunk * 32 ...
                if (chunk * 32 < qsteps * 16) {  // (block-uniform) ... that the products read
                    float s[32], dp[32];
                    lb::ld_chunk(T_S + lane_off + (uint32_t)(chunk * 32), s);
                    lb::ld_chunk(T_P + lane_off + (uint32_t)(chunk * 32), dp);
                    const float4* l4 = reinterpret_cast<const float4*>(sLse2 + chunk * 32);
                    const float4* d4 = reinterpret_cast<const float4*>(sDelta + chunk * 32);
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        const float4 l = l4[i >> 2], d = d4[i >> 2];
                        const float ls[4] = {l.x, l.y, l.z, l.w}, ds[4] = {d.x, d.y, d.z, d.w};
                        float pv[4], gv[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            pv[u] = key_ok ? lb::ex2(fmaf(s[i + u], p.scale_log2e, -ls[u])) : 0.f;
                            gv[u] = pv[u] * (dp[i + u] - ds[u]) * p.scale;
                        }
                        const __nv_bfloat162 p0 = __floats2bfloat162_rn(pv[0], pv[1]), p1 = __floats2bfloat162_rn(pv[2], pv[3]);
                        const __nv_bfloat162 g0 = __floats2bfloat162_rn(gv[0], gv[1]), g1 = __floats2bfloat162_rn(gv[2], gv[3]);
                        pw[cc][i >> 1] = *reinterpret_cast<const uint32_t*>(&p0);
                        pw[cc][(i >> 1) + 1] = *reinterpret_cast<const uint32_t*>(&p1);
                        dw[cc][i >> 1] = *reinterpret_cast<const uint32_t*>(&g0);
                        dw[cc][(i >> 1) + 1] = *reinterpret_cast<const uint32_t*>(&g1);
                    }
                }
            }
            {   // S^T, dP^T are in registers: the tensor core may overwrite them with the next tile's
                tc_fence_before();
                __syncwarp();
                if (lane == 0) lb::mbar_arrive(s_free);
            }
            if (threadIdx.x == 0) XT(8 * j + 1);
            if (j > 0) {  // the previous tile's gradient products have run: its dK, dV leave, and sPT / sDST may be rewritten
                wait(g_done, (uint32_t)(j - 1) & 1u);
                tc_fence_after();
                if (threadIdx.x == 0) XT(8 * j + 2);
                drain(j - 1);
                if (threadIdx.x == 0) XT(8 * j + 3);
            }
#pragma unroll
            for (int cc = 0; cc < CH; ++cc) {
                const int chunk = cg + 2 * cc;
                if (chunk * 32 < qsteps * 16) {
                    store_half_row_packed(sPT + (chunk >> 1) * ATOM, row, chunk & 1, pw[cc]);
                    store_half_row_packed(sDST + (chunk >> 1) * ATOM, row, chunk & 1, dw[cc]);
                }
            }
            signal(ew_p);
            if (threadIdx.x == 0) XT(8 * j + 4);
        }
        wait(g_done, (uint32_t)(tiles - 1) & 1u);
        tc_fence_after();
        drain(tiles - 1);
        {   // dQ_h, accumulated over all tiles: lanes = queries
            float v[32];
            lb::ld_chunk(T_DQ + lane_off + (uint32_t)(cg * 32), v);
            if (row < n) lb::store32_bf16_global(p.dq16 + (int64_t)b * p.q_bs + (int64_t)row * p.q_rs + h * DH + cg * 32, v);
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

// all queries of a sample in one 128-row tile, more than one key tile, head size 64, no attention dropout; rows of q / o /
// dO / dq contiguous
bool xattn_bwd_tc_supported(const pio_attn_desc* d, const DropCfg* drop) {
    const bool off = getenv("PIO_NO_XATTN_TC") != nullptr;
    if (off || d == nullptr) return false;
    if (drop != nullptr && drop->threshold != 0) return false;
    const int64_t cols = (int64_t)d->heads * d->dh;
    // keys that fit ONE tile stay with the general kernels: there the row term is formed as delta = sum_k p dP from the
    // recomputed probabilities, which makes every row of dS sum to zero exactly — the decoder's query LayerNorm gradient
    // (a sum of dQ over all rows, mostly cancelling) is 4x more accurate that way; over several key tiles delta = dO . O
    // is the only one-pass form, here as there
    return d->dh == xat::DH && d->n >= 1 && d->n <= xat::ROWS && d->m > xat::ROWS && d->m <= xat::ROWS * xat::MAX_TILES && d->batch >= 1 &&
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
        PIO_CUDA(cudaFuncSetAttribute(xattn_bwd_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        PIO_CUDA(cudaFuncSetAttribute(xattn_bwd_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
        configured[dev].store(true);
    }
    const dim3 grid((unsigned)d->heads, (unsigned)d->batch);
    if (d->n <= 64) PIO_CUDA(launch_pdl(xattn_bwd_tc_kernel<64>, grid, dim3(THREADS), SMEM_BYTES, st, tq, tdo, tk, tv, p));
    else PIO_CUDA(launch_pdl(xattn_bwd_tc_kernel<128>, grid, dim3(THREADS), SMEM_BYTES, st, tq, tdo, tk, tv, p));
    count_launch();
    return PIO_OK;
}

}  // namespace pio
