// fp32 flash-style attention for PIO_MODE_FP32 and for head sizes the tensor-core kernel does not
// take.  Forward: online softmax over 32-key tiles, warp-shuffle max/sum reductions (lane = key
// for the score dot products, lane = channel for the P·V accumulation), nothing of size n x m
// ever leaves the SM.  Backward recomputes P from the saved log-sum-exp: one kernel owns query
// rows (dQ, and the row term delta = dO·O), one owns key rows (dK, dV) — no atomics.
#include "pio_ops_common.cuh"

namespace pio {

constexpr int AT_WARPS = 4;
constexpr int AT_RPW = 4;                 // rows (queries or keys) per warp
constexpr int AT_ROWS = AT_WARPS * AT_RPW;  // 16 rows owned by a CTA
constexpr int AT_TILE = 32;               // streamed rows per tile (one per lane)
constexpr int AT_THREADS = AT_WARPS * 32;

struct AttnP {
    const void *q, *k, *v;  // fp32, or bf16 when in16
    const float *o, *d_o, *lse_in;
    const uint8_t* pad;
    float *out, *lse, *dq, *dk, *dv, *delta;
    __nv_bfloat16 *out16, *dq16, *dk16, *dv16;  // optional bf16 copies, same element strides
    int64_t q_bs, q_rs, k_bs, k_rs, v_bs, v_rs, o_bs, o_rs;
    int batch, heads, n, m, dh;
    float scale;
    int in16;  // q, k, v are bf16 (same element strides) instead of fp32
    DropCfg drop;  // dropout on the attention probabilities (nn.MultiheadAttention(dropout=p)); threshold 0 = off
};

__device__ __forceinline__ float load_in(const void* base, int64_t idx, bool in16) {
    return in16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx])
                : __ldg(reinterpret_cast<const float*>(base) + idx);
}

// rows [r0, r0+nrows) of a (rows, dh) head slice -> smem transposed [DHP][AT_ROWS], zero padded
template <int DHP>
__device__ __forceinline__ void load_rows_T(float* sT, const float* __restrict__ g, int64_t rs, int r0, int rlim, int dh) {
    for (int idx = threadIdx.x; idx < AT_ROWS * DHP; idx += AT_THREADS) {
        const int r = idx / DHP, d = idx - r * DHP;
        float v = 0.f;
        if (r0 + r < rlim && d < dh) v = __ldg(g + (int64_t)(r0 + r) * rs + d);
        sT[d * AT_ROWS + r] = v;
    }
}
// same, source fp32 or bf16 at element offset `off` of `base`
template <int DHP>
__device__ __forceinline__ void load_rows_T(float* sT, const void* base, int64_t off, bool in16, int64_t rs, int r0, int rlim,
                                            int dh) {
    for (int idx = threadIdx.x; idx < AT_ROWS * DHP; idx += AT_THREADS) {
        const int r = idx / DHP, d = idx - r * DHP;
        float v = 0.f;
        if (r0 + r < rlim && d < dh) v = load_in(base, off + (int64_t)(r0 + r) * rs + d, in16);
        sT[d * AT_ROWS + r] = v;
    }
}
template <int DHP>
__device__ __forceinline__ void load_tile_rows(float* s, const void* base, int64_t off, bool in16, int64_t rs, int r0, int rlim,
                                               int dh) {
    for (int idx = threadIdx.x; idx < AT_TILE * DHP; idx += AT_THREADS) {
        const int r = idx / DHP, d = idx - r * DHP;
        float v = 0.f;
        if (r0 + r < rlim && d < dh) v = load_in(base, off + (int64_t)(r0 + r) * rs + d, in16);
        s[r * (DHP + 1) + d] = v;
    }
}
// rows [r0, r0+32) -> smem [32][DHP+1], zero padded
template <int DHP>
__device__ __forceinline__ void load_tile_rows(float* s, const float* __restrict__ g, int64_t rs, int r0, int rlim, int dh) {
    for (int idx = threadIdx.x; idx < AT_TILE * DHP; idx += AT_THREADS) {
        const int r = idx / DHP, d = idx - r * DHP;
        float v = 0.f;
        if (r0 + r < rlim && d < dh) v = __ldg(g + (int64_t)(r0 + r) * rs + d);
        s[r * (DHP + 1) + d] = v;
    }
}

template <int NCH>
__global__ void __launch_bounds__(AT_THREADS) attn_fwd_kernel(const AttnP p) {
    constexpr int DHP = 32 * NCH;
    extern __shared__ __align__(16) float smem[];
    float* QsT = smem;                         // [DHP][16]
    float* Ks = QsT + DHP * AT_ROWS;           // [32][DHP+1]
    float* Vs = Ks + AT_TILE * (DHP + 1);      // [32][DHP+1]
    float* Ps = Vs + AT_TILE * (DHP + 1);      // [warps][32][4]
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * AT_ROWS;
    const int64_t qoff = (int64_t)b * p.q_bs + (int64_t)h * p.dh;
    const int64_t koff = (int64_t)b * p.k_bs + (int64_t)h * p.dh;
    const int64_t voff = (int64_t)b * p.v_bs + (int64_t)h * p.dh;
    const bool in16 = p.in16 != 0;
    const uint8_t* padb = p.pad ? p.pad + (int64_t)b * p.m : nullptr;

    load_rows_T<DHP>(QsT, p.q, qoff, in16, p.q_rs, q0, p.n, p.dh);
    float m_run[AT_RPW], l_run[AT_RPW], acc[AT_RPW][NCH];
#pragma unroll
    for (int i = 0; i < AT_RPW; ++i) {
        m_run[i] = -INFINITY;
        l_run[i] = 0.f;
#pragma unroll
        for (int c = 0; c < NCH; ++c) acc[i][c] = 0.f;
    }
    float* Pw = Ps + wid * AT_TILE * 4;
    for (int j0 = 0; j0 < p.m; j0 += AT_TILE) {
        __syncthreads();  // previous tile fully consumed (also orders the Q load on the first pass)
        load_tile_rows<DHP>(Ks, p.k, koff, in16, p.k_rs, j0, p.m, p.dh);
        load_tile_rows<DHP>(Vs, p.v, voff, in16, p.v_rs, j0, p.m, p.dh);
        __syncthreads();
        float s[AT_RPW] = {0.f, 0.f, 0.f, 0.f};
        const float* krow = Ks + lane * (DHP + 1);
#pragma unroll 8
        for (int d = 0; d < DHP; ++d) {
            const float kv = krow[d];
            const float4 qv = *reinterpret_cast<const float4*>(QsT + d * AT_ROWS + wid * AT_RPW);
            s[0] = fmaf(qv.x, kv, s[0]);
            s[1] = fmaf(qv.y, kv, s[1]);
            s[2] = fmaf(qv.z, kv, s[2]);
            s[3] = fmaf(qv.w, kv, s[3]);
        }
        const int j = j0 + lane;
        const bool dead = (j >= p.m) || (padb != nullptr && padb[j] != 0);
        float pr[AT_RPW];
#pragma unroll
        for (int i = 0; i < AT_RPW; ++i) {
            const float sv = dead ? -INFINITY : s[i] * p.scale;
            const float mx = warp_max(sv);
            const float m_new = fmaxf(m_run[i], mx);
            float alpha = 1.f, pv = 0.f;
            if (m_new != -INFINITY) {
                alpha = expf(m_run[i] - m_new);  // exp(-inf) = 0 on the first live tile
                pv = dead ? 0.f : expf(sv - m_new);
            }
            l_run[i] = l_run[i] * alpha + warp_sum(pv);
            m_run[i] = m_new;
#pragma unroll
            for (int c = 0; c < NCH; ++c) acc[i][c] *= alpha;
            if (p.drop.threshold != 0) {  // the softmax normalisation sees every key; only P.V sees the mask
                const uint64_t idx = (((uint64_t)b * p.heads + h) * p.n + (q0 + wid * AT_RPW + i)) * (uint64_t)p.m + j;
                pv = drop_keep(p.drop, idx) ? pv * p.drop.inv_keep : 0.f;
            }
            pr[i] = pv;
        }
        *reinterpret_cast<float4*>(Pw + lane * 4) = make_float4(pr[0], pr[1], pr[2], pr[3]);
        __syncwarp();
#pragma unroll 4
        for (int kk = 0; kk < AT_TILE; ++kk) {
            const float4 pv = *reinterpret_cast<const float4*>(Pw + kk * 4);
            const float* vrow = Vs + kk * (DHP + 1) + lane;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const float vv = vrow[32 * c];
                acc[0][c] = fmaf(pv.x, vv, acc[0][c]);
                acc[1][c] = fmaf(pv.y, vv, acc[1][c]);
                acc[2][c] = fmaf(pv.z, vv, acc[2][c]);
                acc[3][c] = fmaf(pv.w, vv, acc[3][c]);
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int i = 0; i < AT_RPW; ++i) {
        const int qi = q0 + wid * AT_RPW + i;
        if (qi >= p.n) continue;
        const float inv = 1.0f / l_run[i];  // l = 0 (every key padded) -> inf -> NaN output, as torch
        float* og = p.out + (int64_t)b * p.o_bs + (int64_t)qi * p.o_rs + (int64_t)h * p.dh;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int d = 32 * c + lane;
            if (d < p.dh) {
                const float val = acc[i][c] * inv;
                og[d] = val;
                if (p.out16) p.out16[(int64_t)b * p.o_bs + (int64_t)qi * p.o_rs + (int64_t)h * p.dh + d] = __float2bfloat16(val);
            }
        }
        if (lane == 0) p.lse[((int64_t)b * p.heads + h) * p.n + qi] = m_run[i] + logf(l_run[i]);
    }
}

// dQ (and delta): CTA owns 16 query rows, streams key tiles.
template <int NCH>
__global__ void __launch_bounds__(AT_THREADS) attn_bwd_dq_kernel(const AttnP p) {
    constexpr int DHP = 32 * NCH;
    extern __shared__ __align__(16) float smem[];
    float* QsT = smem;                          // [DHP][16]
    float* dOsT = QsT + DHP * AT_ROWS;          // [DHP][16]
    float* Ks = dOsT + DHP * AT_ROWS;           // [32][DHP+1]
    float* Vs = Ks + AT_TILE * (DHP + 1);       // [32][DHP+1]
    float* dSs = Vs + AT_TILE * (DHP + 1);      // [warps][32][4]
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * AT_ROWS;
    const int64_t qoff = (int64_t)b * p.q_bs + (int64_t)h * p.dh;
    const int64_t koff = (int64_t)b * p.k_bs + (int64_t)h * p.dh;
    const int64_t voff = (int64_t)b * p.v_bs + (int64_t)h * p.dh;
    const bool in16 = p.in16 != 0;
    const float* og = p.o + (int64_t)b * p.o_bs + (int64_t)h * p.dh;
    const float* dog = p.d_o + (int64_t)b * p.o_bs + (int64_t)h * p.dh;
    const uint8_t* padb = p.pad ? p.pad + (int64_t)b * p.m : nullptr;

    load_rows_T<DHP>(QsT, p.q, qoff, in16, p.q_rs, q0, p.n, p.dh);
    load_rows_T<DHP>(dOsT, dog, p.o_rs, q0, p.n, p.dh);
    float delta[AT_RPW], lse[AT_RPW], acc[AT_RPW][NCH];
#pragma unroll
    for (int i = 0; i < AT_RPW; ++i) {
        const int qi = q0 + wid * AT_RPW + i;
        float dsum = 0.f;
        if (qi < p.n) {
            for (int d = lane; d < p.dh; d += 32)
                dsum += __ldg(dog + (int64_t)qi * p.o_rs + d) * __ldg(og + (int64_t)qi * p.o_rs + d);
        }
        delta[i] = warp_sum(dsum);
        lse[i] = (qi < p.n) ? p.lse_in[((int64_t)b * p.heads + h) * p.n + qi] : 0.f;
        if (lane == 0 && qi < p.n) p.delta[((int64_t)b * p.heads + h) * p.n + qi] = delta[i];
#pragma unroll
        for (int c = 0; c < NCH; ++c) acc[i][c] = 0.f;
    }
    float* dSw = dSs + wid * AT_TILE * 4;
    for (int j0 = 0; j0 < p.m; j0 += AT_TILE) {
        __syncthreads();
        load_tile_rows<DHP>(Ks, p.k, koff, in16, p.k_rs, j0, p.m, p.dh);
        load_tile_rows<DHP>(Vs, p.v, voff, in16, p.v_rs, j0, p.m, p.dh);
        __syncthreads();
        float s[AT_RPW] = {0.f, 0.f, 0.f, 0.f}, dp[AT_RPW] = {0.f, 0.f, 0.f, 0.f};
        const float* krow = Ks + lane * (DHP + 1);
        const float* vrow = Vs + lane * (DHP + 1);
#pragma unroll 4
        for (int d = 0; d < DHP; ++d) {
            const float kv = krow[d], vv = vrow[d];
            const float4 qv = *reinterpret_cast<const float4*>(QsT + d * AT_ROWS + wid * AT_RPW);
            const float4 gv = *reinterpret_cast<const float4*>(dOsT + d * AT_ROWS + wid * AT_RPW);
            s[0] = fmaf(qv.x, kv, s[0]); s[1] = fmaf(qv.y, kv, s[1]);
            s[2] = fmaf(qv.z, kv, s[2]); s[3] = fmaf(qv.w, kv, s[3]);
            dp[0] = fmaf(gv.x, vv, dp[0]); dp[1] = fmaf(gv.y, vv, dp[1]);
            dp[2] = fmaf(gv.z, vv, dp[2]); dp[3] = fmaf(gv.w, vv, dp[3]);
        }
        const int j = j0 + lane;
        const bool dead = (j >= p.m) || (padb != nullptr && padb[j] != 0);
        float ds[AT_RPW];
#pragma unroll
        for (int i = 0; i < AT_RPW; ++i) {
            const float pv = dead ? 0.f : expf(s[i] * p.scale - lse[i]);
            float dpi = dp[i];
            if (p.drop.threshold != 0) {
                const uint64_t idx = (((uint64_t)b * p.heads + h) * p.n + (q0 + wid * AT_RPW + i)) * (uint64_t)p.m + j;
                dpi = drop_keep(p.drop, idx) ? dpi * p.drop.inv_keep : 0.f;
            }
            ds[i] = pv * (dpi - delta[i]) * p.scale;
        }
        *reinterpret_cast<float4*>(dSw + lane * 4) = make_float4(ds[0], ds[1], ds[2], ds[3]);
        __syncwarp();
#pragma unroll 4
        for (int kk = 0; kk < AT_TILE; ++kk) {
            const float4 dv4 = *reinterpret_cast<const float4*>(dSw + kk * 4);
            const float* kr = Ks + kk * (DHP + 1) + lane;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const float kv = kr[32 * c];
                acc[0][c] = fmaf(dv4.x, kv, acc[0][c]);
                acc[1][c] = fmaf(dv4.y, kv, acc[1][c]);
                acc[2][c] = fmaf(dv4.z, kv, acc[2][c]);
                acc[3][c] = fmaf(dv4.w, kv, acc[3][c]);
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int i = 0; i < AT_RPW; ++i) {
        const int qi = q0 + wid * AT_RPW + i;
        if (qi >= p.n) continue;
        const int64_t dqo = (int64_t)b * p.q_bs + (int64_t)qi * p.q_rs + (int64_t)h * p.dh;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int d = 32 * c + lane;
            if (d < p.dh) {
                if (p.dq) p.dq[dqo + d] = acc[i][c];
                if (p.dq16) p.dq16[(int64_t)b * p.q_bs + (int64_t)qi * p.q_rs + (int64_t)h * p.dh + d] = __float2bfloat16(acc[i][c]);
            }
        }
    }
}

// dK, dV: CTA owns 16 key rows, streams query tiles (lane = query).
template <int NCH>
__global__ void __launch_bounds__(AT_THREADS) attn_bwd_dkv_kernel(const AttnP p) {
    constexpr int DHP = 32 * NCH;
    extern __shared__ __align__(16) float smem[];
    float* KsT = smem;                          // [DHP][16]
    float* VsT = KsT + DHP * AT_ROWS;           // [DHP][16]
    float* Qs = VsT + DHP * AT_ROWS;            // [32][DHP+1]
    float* dOs = Qs + AT_TILE * (DHP + 1);      // [32][DHP+1]
    float* Ps = dOs + AT_TILE * (DHP + 1);      // [warps][32][4]
    float* dSs = Ps + AT_WARPS * AT_TILE * 4;   // [warps][32][4]
    float* lse_s = dSs + AT_WARPS * AT_TILE * 4;  // [32]
    float* del_s = lse_s + AT_TILE;               // [32]
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * AT_ROWS;
    const int64_t qoff = (int64_t)b * p.q_bs + (int64_t)h * p.dh;
    const int64_t koff = (int64_t)b * p.k_bs + (int64_t)h * p.dh;
    const int64_t voff = (int64_t)b * p.v_bs + (int64_t)h * p.dh;
    const bool in16 = p.in16 != 0;
    const float* dog = p.d_o + (int64_t)b * p.o_bs + (int64_t)h * p.dh;
    const uint8_t* padb = p.pad ? p.pad + (int64_t)b * p.m : nullptr;

    load_rows_T<DHP>(KsT, p.k, koff, in16, p.k_rs, k0, p.m, p.dh);
    load_rows_T<DHP>(VsT, p.v, voff, in16, p.v_rs, k0, p.m, p.dh);
    bool kdead[AT_RPW];
    float dk[AT_RPW][NCH], dv[AT_RPW][NCH];
#pragma unroll
    for (int i = 0; i < AT_RPW; ++i) {
        const int kj = k0 + wid * AT_RPW + i;
        kdead[i] = (kj >= p.m) || (padb != nullptr && padb[kj] != 0);
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            dk[i][c] = 0.f;
            dv[i][c] = 0.f;
        }
    }
    float* Pw = Ps + wid * AT_TILE * 4;
    float* dSw = dSs + wid * AT_TILE * 4;
    const int64_t stat_base = ((int64_t)b * p.heads + h) * p.n;
    for (int i0 = 0; i0 < p.n; i0 += AT_TILE) {
        __syncthreads();
        load_tile_rows<DHP>(Qs, p.q, qoff, in16, p.q_rs, i0, p.n, p.dh);
        load_tile_rows<DHP>(dOs, dog, p.o_rs, i0, p.n, p.dh);
        if (threadIdx.x < AT_TILE) {
            const int qi = i0 + threadIdx.x;
            lse_s[threadIdx.x] = (qi < p.n) ? p.lse_in[stat_base + qi] : 0.f;
            del_s[threadIdx.x] = (qi < p.n) ? p.delta[stat_base + qi] : 0.f;
        }
        __syncthreads();
        float s[AT_RPW] = {0.f, 0.f, 0.f, 0.f}, dp[AT_RPW] = {0.f, 0.f, 0.f, 0.f};
        const float* qrow = Qs + lane * (DHP + 1);
        const float* grow = dOs + lane * (DHP + 1);
#pragma unroll 4
        for (int d = 0; d < DHP; ++d) {
            const float qv = qrow[d], gv = grow[d];
            const float4 kv = *reinterpret_cast<const float4*>(KsT + d * AT_ROWS + wid * AT_RPW);
            const float4 vv = *reinterpret_cast<const float4*>(VsT + d * AT_ROWS + wid * AT_RPW);
            s[0] = fmaf(kv.x, qv, s[0]); s[1] = fmaf(kv.y, qv, s[1]);
            s[2] = fmaf(kv.z, qv, s[2]); s[3] = fmaf(kv.w, qv, s[3]);
            dp[0] = fmaf(vv.x, gv, dp[0]); dp[1] = fmaf(vv.y, gv, dp[1]);
            dp[2] = fmaf(vv.z, gv, dp[2]); dp[3] = fmaf(vv.w, gv, dp[3]);
        }
        const bool qdead = (i0 + lane >= p.n);
        const float lse = lse_s[lane], del = del_s[lane];
        float pr[AT_RPW], ds[AT_RPW];
#pragma unroll
        for (int i = 0; i < AT_RPW; ++i) {
            const float pv = (qdead || kdead[i]) ? 0.f : expf(s[i] * p.scale - lse);
            float keep_scale = 1.f;
            if (p.drop.threshold != 0) {
                const uint64_t idx = (((uint64_t)b * p.heads + h) * p.n + (i0 + lane)) * (uint64_t)p.m + (k0 + wid * AT_RPW + i);
                keep_scale = drop_keep(p.drop, idx) ? p.drop.inv_keep : 0.f;
            }
            pr[i] = pv * keep_scale;
            ds[i] = pv * (dp[i] * keep_scale - del) * p.scale;
        }
        *reinterpret_cast<float4*>(Pw + lane * 4) = make_float4(pr[0], pr[1], pr[2], pr[3]);
        *reinterpret_cast<float4*>(dSw + lane * 4) = make_float4(ds[0], ds[1], ds[2], ds[3]);
        __syncwarp();
#pragma unroll 2
        for (int qq = 0; qq < AT_TILE; ++qq) {
            const float4 p4 = *reinterpret_cast<const float4*>(Pw + qq * 4);
            const float4 d4 = *reinterpret_cast<const float4*>(dSw + qq * 4);
            const float* gr = dOs + qq * (DHP + 1) + lane;
            const float* qr = Qs + qq * (DHP + 1) + lane;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                const float gv = gr[32 * c], qv = qr[32 * c];
                dv[0][c] = fmaf(p4.x, gv, dv[0][c]); dv[1][c] = fmaf(p4.y, gv, dv[1][c]);
                dv[2][c] = fmaf(p4.z, gv, dv[2][c]); dv[3][c] = fmaf(p4.w, gv, dv[3][c]);
                dk[0][c] = fmaf(d4.x, qv, dk[0][c]); dk[1][c] = fmaf(d4.y, qv, dk[1][c]);
                dk[2][c] = fmaf(d4.z, qv, dk[2][c]); dk[3][c] = fmaf(d4.w, qv, dk[3][c]);
            }
        }
        __syncwarp();
    }
#pragma unroll
    for (int i = 0; i < AT_RPW; ++i) {
        const int kj = k0 + wid * AT_RPW + i;
        if (kj >= p.m) continue;
        const int64_t dko = (int64_t)b * p.k_bs + (int64_t)kj * p.k_rs + (int64_t)h * p.dh;
        const int64_t dvo = (int64_t)b * p.v_bs + (int64_t)kj * p.v_rs + (int64_t)h * p.dh;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int d = 32 * c + lane;
            if (d < p.dh) {
                if (p.dk) p.dk[dko + d] = dk[i][c];
                if (p.dv) p.dv[dvo + d] = dv[i][c];
                if (p.dk16) p.dk16[(int64_t)b * p.k_bs + (int64_t)kj * p.k_rs + (int64_t)h * p.dh + d] = __float2bfloat16(dk[i][c]);
                if (p.dv16) p.dv16[(int64_t)b * p.v_bs + (int64_t)kj * p.v_rs + (int64_t)h * p.dh + d] = __float2bfloat16(dv[i][c]);
            }
        }
    }
}

static size_t fwd_smem(int dhp) { return (size_t)(dhp * AT_ROWS + 2 * AT_TILE * (dhp + 1) + AT_WARPS * AT_TILE * 4) * sizeof(float); }
static size_t dq_smem(int dhp) { return (size_t)(2 * dhp * AT_ROWS + 2 * AT_TILE * (dhp + 1) + AT_WARPS * AT_TILE * 4) * sizeof(float); }
static size_t dkv_smem(int dhp) {
    return (size_t)(2 * dhp * AT_ROWS + 2 * AT_TILE * (dhp + 1) + 2 * AT_WARPS * AT_TILE * 4 + 2 * AT_TILE) * sizeof(float);
}

template <typename K>
static int launch(K kernel, dim3 grid, size_t smem, const AttnP& p, cudaStream_t st) {
    if (smem > 48 * 1024) PIO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kernel<<<grid, AT_THREADS, smem, st>>>(p);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

static int check_desc(const pio_attn_desc* d, const char* who) {
    PIO_CHECK_ARG(d && d->q && d->k && d->v && d->o && d->lse, "%s: null pointer", who);
    PIO_CHECK_ARG(d->batch >= 0 && d->heads > 0 && d->n >= 0 && d->m >= 0 && d->dh > 0, "%s: bad dims", who);
    PIO_CHECK_ARG(d->dh <= 512, "%s: head dim %d > 512 unsupported", who, d->dh);
    PIO_CHECK_ARG(d->heads <= 65535 && d->batch <= 65535, "%s: batch/heads exceed grid limits", who);
    return PIO_OK;
}

static void fill(AttnP& p, const pio_attn_desc* d) {
    p.q = d->q; p.k = d->k; p.v = d->v; p.pad = d->pad;
    p.q_bs = d->q_bs; p.q_rs = d->q_rs; p.k_bs = d->k_bs; p.k_rs = d->k_rs;
    p.v_bs = d->v_bs; p.v_rs = d->v_rs; p.o_bs = d->o_bs; p.o_rs = d->o_rs;
    p.batch = d->batch; p.heads = d->heads; p.n = d->n; p.m = d->m; p.dh = d->dh; p.scale = d->scale;
}

#define PIO_DISPATCH_NCH(dh, EXPR)        \
    do {                                  \
        if ((dh) <= 32) { constexpr int NCH = 1; EXPR; }        \
        else if ((dh) <= 64) { constexpr int NCH = 2; EXPR; }   \
        else if ((dh) <= 128) { constexpr int NCH = 4; EXPR; }  \
        else if ((dh) <= 256) { constexpr int NCH = 8; EXPR; }  \
        else { constexpr int NCH = 16; EXPR; }                  \
    } while (0)

int attn_fwd_simt(const pio_attn_desc* d, cudaStream_t st, const AttnBf16Out* extra) {
    PIO_TRY(check_desc(d, "pio_attn_fwd"));
    if (d->batch == 0 || d->n == 0) return PIO_OK;
    AttnP p{};
    fill(p, d);
    p.out = d->o;
    p.lse = d->lse;
    if (extra) {
        p.out16 = extra->o16;
        p.in16 = extra->in16;
        p.drop = extra->drop;
    }
    dim3 grid((unsigned)cdiv(d->n, AT_ROWS), (unsigned)d->heads, (unsigned)d->batch);
    int rc = PIO_OK;
    PIO_DISPATCH_NCH(d->dh, rc = launch(attn_fwd_kernel<NCH>, grid, fwd_smem(32 * NCH), p, st));
    return rc;
}

int attn_bwd_simt(const pio_attn_bwd_desc* bd, cudaStream_t st, const AttnBf16Out* extra) {
    const pio_attn_desc* d = &bd->fwd;
    PIO_TRY(check_desc(d, "pio_attn_bwd"));
    PIO_CHECK_ARG(bd->d_o && bd->delta, "pio_attn_bwd: null pointer");
    PIO_CHECK_ARG((bd->dq || (extra && extra->dq16)) && (bd->dk || (extra && extra->dk16)) && (bd->dv || (extra && extra->dv16)),
                  "pio_attn_bwd: every gradient needs an fp32 or a bf16 destination");
    if (d->batch == 0) return PIO_OK;
    AttnP p{};
    fill(p, d);
    p.o = d->o;
    p.lse_in = d->lse;
    p.d_o = bd->d_o;
    p.dq = bd->dq; p.dk = bd->dk; p.dv = bd->dv; p.delta = bd->delta;
    if (extra) {
        p.drop = extra->drop;
        p.in16 = extra->in16;
        p.dq16 = extra->dq16;
        p.dk16 = extra->dk16;
        p.dv16 = extra->dv16;
    }
    int rc = PIO_OK;
    if (d->n > 0) {
        dim3 gq((unsigned)cdiv(d->n, AT_ROWS), (unsigned)d->heads, (unsigned)d->batch);
        PIO_DISPATCH_NCH(d->dh, rc = launch(attn_bwd_dq_kernel<NCH>, gq, dq_smem(32 * NCH), p, st));
        PIO_TRY(rc);
    }
    if (d->m > 0) {
        dim3 gk((unsigned)cdiv(d->m, AT_ROWS), (unsigned)d->heads, (unsigned)d->batch);
        PIO_DISPATCH_NCH(d->dh, rc = launch(attn_bwd_dkv_kernel<NCH>, gk, dkv_smem(32 * NCH), p, st));
    }
    return rc;
}

}  // namespace pio

extern "C" {

int pio_attn_fwd(const pio_attn_desc* d, int mode, pio_stream_t stream) {
    if (mode == PIO_MODE_BF16) {  // q, k, v are bf16 behind the float* type
        PIO_CHECK_ARG(d != nullptr, "pio_attn_fwd: null descriptor");
        if (pio::attn_fwd_tc_supported(d)) return pio::attn_fwd_tc(d, nullptr, (cudaStream_t)stream);
        pio::AttnBf16Out ex;
        ex.in16 = 1;
        return pio::attn_fwd_simt(d, (cudaStream_t)stream, &ex);
    }
    return pio::attn_fwd_simt(d, (cudaStream_t)stream);
}

int pio_attn_bwd(const pio_attn_bwd_desc* d, int mode, pio_stream_t stream) {
    if (mode == PIO_MODE_BF16) {
        pio::AttnBf16Out ex;
        ex.in16 = 1;
        return pio::attn_bwd_simt(d, (cudaStream_t)stream, &ex);
    }
    return pio::attn_bwd_simt(d, (cudaStream_t)stream);
}

}  // extern "C"
