// fp32 FFMA GEMM for PIO_MODE_FP32 (the rel-1e-4 parity mode): C[m,n] = sum_k A(m,k) B(n,k)
// with a fused epilogue (bias / GELU / GELU' / residual / accumulate), operand layouts chosen
// per call so forward (X W^T), input-gradient (dY W) and weight-gradient (dY^T X) products all
// read their operands in place, and deterministic split-K for the weight-gradient shapes
// (tiny m x n, huge k) so the grid still covers the 148 SMs.
#include "pio_common.cuh"

namespace pio {

constexpr int GEMM_BK = 16;
constexpr int GEMM_THREADS = 256;

struct GemmParams {
    const float* a;
    const float* b;
    float* c;
    int64_t m, n, k;
    int64_t lda, ldb, ldc;
    const float* bias;
    const float* residual;
    int64_t ldr;
    float* aux;
    int64_t ldaux;
    int act;
    int accumulate;
    float* rowsum_a;
    // split-K
    int splits;
    int64_t k_per_split;
    float* partial;          // [splits][m*n + (rowsum ? m : 0)]
    int64_t partial_stride;
    DropCfg drop;            // dropout on the value before the residual add
};

template <bool KMAJOR, int BMN>
__device__ __forceinline__ void load_tile(const float* __restrict__ p, int64_t ld, int64_t mn_base, int64_t mn_lim,
                                          int64_t k_base, int64_t k_lim, float (&reg)[BMN * GEMM_BK / GEMM_THREADS]) {
    constexpr int PER = BMN * GEMM_BK / GEMM_THREADS;
    const int t = threadIdx.x;
    if (KMAJOR) {
        const int kk = t & (GEMM_BK - 1);
        const int r0 = t >> 4;  // 16 rows per pass
        const int64_t kg = k_base + kk;
#pragma unroll
        for (int p2 = 0; p2 < PER; ++p2) {
            const int64_t r = mn_base + r0 + 16 * p2;
            reg[p2] = (r < mn_lim && kg < k_lim) ? __ldg(p + r * ld + kg) : 0.f;
        }
    } else {
        const int rr = t % BMN;
        const int k0 = t / BMN;
        constexpr int KSTEP = GEMM_THREADS / BMN;
        const int64_t r = mn_base + rr;
#pragma unroll
        for (int p2 = 0; p2 < PER; ++p2) {
            const int64_t kg = k_base + k0 + KSTEP * p2;
            reg[p2] = (r < mn_lim && kg < k_lim) ? __ldg(p + kg * ld + r) : 0.f;
        }
    }
}

template <bool KMAJOR, int BMN>
__device__ __forceinline__ void store_tile(float (*s)[BMN + 4], const float (&reg)[BMN * GEMM_BK / GEMM_THREADS]) {
    constexpr int PER = BMN * GEMM_BK / GEMM_THREADS;
    const int t = threadIdx.x;
    if (KMAJOR) {
        const int kk = t & (GEMM_BK - 1);
        const int r0 = t >> 4;
#pragma unroll
        for (int p2 = 0; p2 < PER; ++p2) s[kk][r0 + 16 * p2] = reg[p2];
    } else {
        const int rr = t % BMN;
        const int k0 = t / BMN;
        constexpr int KSTEP = GEMM_THREADS / BMN;
#pragma unroll
        for (int p2 = 0; p2 < PER; ++p2) s[k0 + KSTEP * p2][rr] = reg[p2];
    }
}

template <bool A_KMAJOR, bool B_KMAJOR, int BM, int BN>
__global__ void __launch_bounds__(GEMM_THREADS) gemm_f32_kernel(const GemmParams g) {
    constexpr int TM = BM / 16, TN = BN / 16;  // 8 or 4
    constexpr int GM = TM / 4, GN = TN / 4;    // groups of 4 rows / cols
    __shared__ __align__(16) float As[2][GEMM_BK][BM + 4];
    __shared__ __align__(16) float Bs[2][GEMM_BK][BN + 4];

    const int t = threadIdx.x;
    const int tx = t & 15, ty = t >> 4;
    const int64_t m_base = (int64_t)blockIdx.y * BM;
    const int64_t n_base = (int64_t)blockIdx.x * BN;
    const int z = blockIdx.z;
    const int64_t k_begin = (int64_t)z * g.k_per_split;
    const int64_t k_end = (g.splits > 1 && k_begin + g.k_per_split < g.k) ? (k_begin + g.k_per_split) : g.k;

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    float rs = 0.f;  // rowsum of A for row m_base + t (t < BM), only in n-tile 0
    const bool do_rowsum = (g.rowsum_a != nullptr) && (blockIdx.x == 0);

    float ra[BM * GEMM_BK / GEMM_THREADS], rb[BN * GEMM_BK / GEMM_THREADS];
    const int64_t nk = (k_end > k_begin) ? (k_end - k_begin + GEMM_BK - 1) / GEMM_BK : 0;
    if (nk > 0) {
        load_tile<A_KMAJOR, BM>(g.a, g.lda, m_base, g.m, k_begin, k_end, ra);
        load_tile<B_KMAJOR, BN>(g.b, g.ldb, n_base, g.n, k_begin, k_end, rb);
        store_tile<A_KMAJOR, BM>(As[0], ra);
        store_tile<B_KMAJOR, BN>(Bs[0], rb);
    }
    __syncthreads();
    for (int64_t it = 0; it < nk; ++it) {
        const int cur = (int)(it & 1);
        if (it + 1 < nk) {
            const int64_t kb = k_begin + (it + 1) * GEMM_BK;
            load_tile<A_KMAJOR, BM>(g.a, g.lda, m_base, g.m, kb, k_end, ra);
            load_tile<B_KMAJOR, BN>(g.b, g.ldb, n_base, g.n, kb, k_end, rb);
        }
#pragma unroll
        for (int kk = 0; kk < GEMM_BK; ++kk) {
            float av[TM], bv[TN];
#pragma unroll
            for (int gi = 0; gi < GM; ++gi) {
                const float4 v = *reinterpret_cast<const float4*>(&As[cur][kk][gi * (BM / 2) + ty * 4]);
                av[gi * 4 + 0] = v.x; av[gi * 4 + 1] = v.y; av[gi * 4 + 2] = v.z; av[gi * 4 + 3] = v.w;
            }
#pragma unroll
            for (int gj = 0; gj < GN; ++gj) {
                const float4 v = *reinterpret_cast<const float4*>(&Bs[cur][kk][gj * (BN / 2) + tx * 4]);
                bv[gj * 4 + 0] = v.x; bv[gj * 4 + 1] = v.y; bv[gj * 4 + 2] = v.z; bv[gj * 4 + 3] = v.w;
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (do_rowsum && t < BM) {
#pragma unroll
            for (int kk = 0; kk < GEMM_BK; ++kk) rs += As[cur][kk][t];
        }
        if (it + 1 < nk) {
            store_tile<A_KMAJOR, BM>(As[cur ^ 1], ra);
            store_tile<B_KMAJOR, BN>(Bs[cur ^ 1], rb);
        }
        __syncthreads();
    }

    // ---- epilogue ----
    if (g.splits > 1) {
        float* out = g.partial + (int64_t)z * g.partial_stride;
#pragma unroll
        for (int i = 0; i < TM; ++i) {
            const int64_t m = m_base + (i / 4) * (BM / 2) + ty * 4 + (i & 3);
            if (m >= g.m) continue;
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                const int64_t n = n_base + (j / 4) * (BN / 2) + tx * 4 + (j & 3);
                if (n < g.n) out[m * g.n + n] = acc[i][j];
            }
        }
        if (do_rowsum && t < BM && m_base + t < g.m) out[g.m * g.n + m_base + t] = rs;
        return;
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int64_t m = m_base + (i / 4) * (BM / 2) + ty * 4 + (i & 3);
        if (m >= g.m) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int64_t n = n_base + (j / 4) * (BN / 2) + tx * 4 + (j & 3);
            if (n >= g.n) continue;
            float v = acc[i][j];
            if (g.bias) v += __ldg(g.bias + n);
            if (g.act == 1) {
                if (g.aux) g.aux[m * g.ldaux + n] = v;
                v = gelu_erf(v);
            } else if (g.act == 2) {
                v *= gelu_erf_grad(g.aux[m * g.ldaux + n]);
            }
            if (g.drop.threshold != 0) v = drop_keep(g.drop, (uint64_t)m * (uint64_t)g.n + (uint64_t)n) ? v * g.drop.inv_keep : 0.f;
            if (g.residual) v += g.residual[m * g.ldr + n];
            float* dst = g.c + m * g.ldc + n;
            if (g.accumulate) v += *dst;
            *dst = v;
        }
    }
    if (do_rowsum && t < BM && m_base + t < g.m) {
        float* dst = g.rowsum_a + m_base + t;
        *dst = g.accumulate ? (*dst + rs) : rs;
    }
}

static void choose_split(int64_t m, int64_t n, int64_t k, bool splittable, int* splits, int64_t* k_per, bool* small_tile) {
    const int sms = num_sms();
    const int64_t tiles128 = cdiv(m, 128) * cdiv(n, 128);
    *small_tile = tiles128 < sms;
    const int bm = *small_tile ? 64 : 128;
    const int64_t tiles = cdiv(m, bm) * cdiv(n, bm);
    int s = 1;
    if (splittable && k >= 1024 && tiles < 2 * sms) {
        s = (int)std::min<int64_t>(cdiv(2 * sms, tiles), k / 256);
        s = std::max(s, 1);
    }
    int64_t kp = cdiv(cdiv(k, s), GEMM_BK) * GEMM_BK;
    s = (int)cdiv(k, kp);
    *splits = s;
    *k_per = kp;
}

}  // namespace pio

using namespace pio;

extern "C" {

size_t pio_gemm_f32_ws_bytes(int64_t m, int64_t n, int64_t k) {
    int splits;
    int64_t kp;
    bool small;
    choose_split(m, n, k, true, &splits, &kp, &small);
    if (splits <= 1) return 256;
    return align_up((size_t)splits * (size_t)(m * n + m) * sizeof(float), 256) + 256;
}

int pio_gemm_f32(const pio_gemm_desc* d, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_CHECK_ARG(d && d->a && d->b && d->c, "pio_gemm_f32: null pointer");
    PIO_CHECK_ARG(d->m >= 0 && d->n >= 0 && d->k >= 0, "pio_gemm_f32: negative dims");
    PIO_CHECK_ARG(d->act >= 0 && d->act <= 2, "pio_gemm_f32: bad act %d", d->act);
    PIO_CHECK_ARG(d->act != 2 || d->aux, "pio_gemm_f32: act=2 needs aux");
    if (d->m == 0 || d->n == 0) return PIO_OK;
    cudaStream_t st = (cudaStream_t)stream;

    GemmParams g;
    g.a = d->a; g.b = d->b; g.c = d->c;
    g.m = d->m; g.n = d->n; g.k = d->k;
    g.lda = d->lda; g.ldb = d->ldb; g.ldc = d->ldc;
    g.bias = d->bias; g.residual = d->residual; g.ldr = d->ldr;
    g.aux = d->aux; g.ldaux = d->ldaux; g.act = d->act; g.accumulate = d->accumulate;
    g.rowsum_a = d->rowsum_a;
    g.drop = make_drop(d->dropout_p, d->dropout_seed);

    const bool splittable = !d->bias && !d->residual && d->act == 0 && d->ldc == d->n && !(d->dropout_p > 0.f);
    int splits;
    int64_t kp;
    bool small;
    choose_split(d->m, d->n, d->k, splittable, &splits, &kp, &small);
    g.partial = nullptr;
    g.partial_stride = d->m * d->n + (d->rowsum_a ? d->m : 0);
    if (splits > 1) {
        Workspace W(ws, ws_bytes);
        g.partial = W.takef((size_t)splits * (size_t)g.partial_stride);
        if (!g.partial) {  // not enough scratch: fall back to a single pass over k
            splits = 1;
            kp = cdiv(d->k, GEMM_BK) * GEMM_BK;
        }
    }
    g.splits = splits;
    g.k_per_split = kp;

    const int bm = small ? 64 : 128;
    dim3 grid((unsigned)cdiv(d->n, bm), (unsigned)cdiv(d->m, bm), (unsigned)splits);
    PIO_CHECK_ARG(grid.y <= 65535 && grid.z <= 65535, "pio_gemm_f32: grid too large");
#define PIO_GEMM_LAUNCH(AK, BKM)                                                       \
    do {                                                                               \
        if (small) gemm_f32_kernel<AK, BKM, 64, 64><<<grid, GEMM_THREADS, 0, st>>>(g);  \
        else gemm_f32_kernel<AK, BKM, 128, 128><<<grid, GEMM_THREADS, 0, st>>>(g);      \
    } while (0)
    if (d->a_kmajor && d->b_kmajor) PIO_GEMM_LAUNCH(true, true);
    else if (d->a_kmajor && !d->b_kmajor) PIO_GEMM_LAUNCH(true, false);
    else if (!d->a_kmajor && d->b_kmajor) PIO_GEMM_LAUNCH(false, true);
    else PIO_GEMM_LAUNCH(false, false);
#undef PIO_GEMM_LAUNCH
    PIO_LAUNCH_CHECK();
    if (splits > 1) {
        PIO_TRY(reduce_partials(g.partial, splits, g.partial_stride, d->c, d->m * d->n, d->rowsum_a,
                                d->rowsum_a ? d->m : 0, d->accumulate, st));
    }
    return PIO_OK;
}

}  // extern "C"
