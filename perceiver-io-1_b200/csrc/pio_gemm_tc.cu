// bf16 tensor-core GEMM for PIO_MODE_BF16: C[m,n] = sum_k A(m,k) B(n,k), fp32 accumulation.
//
//   * operands arrive by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into a 3-stage
//     shared-memory ring guarded by full/empty mbarriers;
//   * one elected thread issues tcgen05.mma (kind::f16, M=128, N=BN, K=16) with the fp32
//     accumulator tile in TMEM; tcgen05.commit releases ring slots and signals the epilogue;
//   * eight epilogue warps (two per TMEM lane quadrant, half the tile's columns each) read the accumulator with
//     tcgen05.ld (lane = row), stage it through shared memory and finish it with coalesced row segments:
//     bias / GELU / GELU' / residual / accumulate / atomics, fp32 and/or bf16; operands the epilogue reads from
//     global memory are requested before the accumulator is waited for (64-wide tiles).
//
// Either operand may be K-major (row = m|n, contiguous k: activations X, weights W for X W^T) or
// MN-major (row = k, contiguous m|n: W for dY W, and dY / X for the weight gradient dY^T X), so
// every product of the forward and backward pass reads its operands where they already lie.
// Warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..9 = epilogue.
#include <cuda.h>

#include <atomic>
#include <stdlib.h>

#include <mutex>

#include "pio_common.cuh"
#include "pio_tc_common.cuh"

namespace pio {

constexpr int TC_BK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int TC_STAGES = 3;  // 3 x 32 KB + barriers: two CTAs per SM, so one tile's epilogue overlaps another's mainloop
constexpr int TC_THREADS = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two per TMEM lane quadrant, half the columns each)

struct TcParams {
    int M, N, K;
    int num_kb, kb_per_split, splits;
    float* c32;
    int64_t ldc32;
    __nv_bfloat16* c16;
    int64_t ldc16;
    const float* bias;
    const float* residual;
    int64_t ldr;
    __nv_bfloat16* aux16;
    int64_t ldaux;
    int act;
    int accumulate;
    float* partial;
    int64_t partial_stride;
    int epi_kind;
    DropCfg drop;  // dropout on the value before the residual add (generic epilogue only)
    const int* dyn_rows;  // device-side row count (see pio_gemm_bf16_desc), or null
    int dyn_off, dyn_axis;
};

// Epilogue flavours, resolved on the host so the row loop below carries no run-time flag tests
// (with four epilogue warps per CTA there is nothing to hide a branchy, latency-bound loop behind).
enum EpiKind : int {
    EPI_F32 = 0,        // c32 = acc + bias
    EPI_F32_RES = 1,    // c32 = acc + bias + residual
    EPI_F32_ACC = 2,    // c32 += acc + bias
    EPI_B16 = 3,        // c16 = acc + bias
    EPI_B16_GELU = 4,   // pre = acc + bias ; c16 = gelu(pre) ; aux16 = gelu'(pre)
    EPI_B16_DGELU = 5,  // c16 = acc * aux16   (aux16 = gelu' saved by the forward)
    EPI_PARTIAL = 6,    // split-K partial sums
    EPI_GENERIC = 7,    // any other combination of outputs / flags
    EPI_F32_ATOMIC = 8  // c32 += acc with fp32 atomics (split-K without a reduction pass; order not deterministic)
};

__device__ __forceinline__ void store4_f32(float* o, const float (&v)[4], bool vec, int valid) {
    if (vec && valid == 4) {
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < valid) o[j] = v[j];
    }
}
__device__ __forceinline__ void store4_b16(__nv_bfloat16* o, const float (&v)[4], bool vec, int valid) {
    if (vec && valid == 4) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]);
        __nv_bfloat162 hi = __floats2bfloat162_rn(v[2], v[3]);
        uint2 u;
        u.x = *reinterpret_cast<uint32_t*>(&lo);
        u.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(o) = u;
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < valid) o[j] = __float2bfloat16(v[j]);
    }
}
__device__ __forceinline__ void load4_b16(const __nv_bfloat16* a, float (&v)[4], bool vec, int valid) {
    if (vec && valid == 4) {
        const uint2 u = *reinterpret_cast<const uint2*>(a);
        const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
        const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
        v[0] = __low2float(lo); v[1] = __high2float(lo); v[2] = __low2float(hi); v[3] = __high2float(hi);
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (j < valid) ? __bfloat162float(a[j]) : 0.f;
    }
}
__device__ __forceinline__ void load4_f32(const float* a, float (&v)[4], bool vec, int valid) {
    if (vec && valid == 4) {
        const float4 t = *reinterpret_cast<const float4*>(a);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = (j < valid) ? a[j] : 0.f;
    }
}

// The global operand that some epilogue flavours read per output row (residual / running sum / saved GELU'): the loads of
// one batch of BATCH row-iterations, issued together.
constexpr int EPI_BATCH = 8;
template <int BN, int KIND>
__device__ __forceinline__ void epilogue_preload(const TcParams& p, int lane, int n0, int64_t gm0, int rows_here, int it0,
                                                 float (&pre)[EPI_BATCH][4]) {
    constexpr int LANES_PER_ROW = BN / 4;
    constexpr int ROWS_PER_IT = 32 / LANES_PER_ROW;
    const int col = (lane % LANES_PER_ROW) * 4;
    const int rsub = lane / LANES_PER_ROW;
    const int n = n0 + col;
    const int valid = min(4, p.N - n);
    const bool v32 = ((p.ldc32 & 3) == 0) && ((n & 3) == 0) && ((((uintptr_t)p.c32) & 15) == 0);
    const bool vaux = ((p.ldaux & 3) == 0) && ((((uintptr_t)p.aux16) & 7) == 0);
    const bool vres = ((p.ldr & 3) == 0) && ((((uintptr_t)p.residual) & 15) == 0);
#pragma unroll
    for (int u = 0; u < EPI_BATCH; ++u) {
        const int rr = (it0 + u) * ROWS_PER_IT + rsub;
        const bool ok = rr < rows_here && valid > 0;
        const int64_t gm = gm0 + (ok ? rr : 0);
        const int vv = ok ? valid : 0;
        if (KIND == EPI_F32_RES) load4_f32(p.residual + gm * p.ldr + n, pre[u], vres, vv);
        else if (KIND == EPI_F32_ACC) load4_f32(p.c32 + gm * p.ldc32 + n, pre[u], v32, vv);
        else load4_b16(p.aux16 + gm * p.ldaux + n, pre[u], vaux, vv);
    }
}

// One warp's 32 accumulator rows x BN columns, staged in shared memory (row pitch BN+4 floats).
// Each lane owns 4 consecutive columns of a row, so every global access is a coalesced row
// segment (16 B per lane when the leading dimensions allow it).
template <int BN, int KIND>
__device__ __forceinline__ void epilogue_rows(const TcParams& p, const float* stile, int lane, int n0, int64_t gm0,
                                              int rows_here, int zsplit, const float (&early)[EPI_BATCH][4], bool use_early) {
    constexpr int PITCH = BN + 4;
    constexpr int LANES_PER_ROW = BN / 4;            // 32 (BN=128) or 16 (BN=64)
    constexpr int ROWS_PER_IT = 32 / LANES_PER_ROW;  // 1 or 2
    const int col = (lane % LANES_PER_ROW) * 4;
    const int rsub = lane / LANES_PER_ROW;
    const int n = n0 + col;
    const int valid = min(4, p.N - n);  // <= 0: this lane has no columns
    float bias[4] = {0.f, 0.f, 0.f, 0.f};
    if (p.bias && KIND != EPI_PARTIAL && KIND != EPI_B16_DGELU && KIND != EPI_F32_ATOMIC) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < valid) bias[j] = __ldg(p.bias + n + j);
    }
    const bool v32 = ((p.ldc32 & 3) == 0) && ((n & 3) == 0) && ((((uintptr_t)p.c32) & 15) == 0);
    const bool v16 = ((p.ldc16 & 3) == 0) && ((((uintptr_t)p.c16) & 7) == 0);
    const bool vaux = ((p.ldaux & 3) == 0) && ((((uintptr_t)p.aux16) & 7) == 0);
    const bool vres = ((p.ldr & 3) == 0) && ((((uintptr_t)p.residual) & 15) == 0);
    const bool vpart = ((p.N & 3) == 0);
    // Flavours that read a global operand per output row (residual, running sum, saved GELU') first issue the
    // loads of a batch of rows and only then consume them: with one dependent load per iteration the four epilogue
    // warps of a CTA had a single 512-byte request in flight each and the kernel ran at the speed of 32 serial
    // memory round trips (ncu: long-scoreboard 22 per issue, tensor pipe 4.6 % on the 32768 x 256 x 256 products).
    constexpr int ITERS = 32 / ROWS_PER_IT;
    constexpr int BATCH = EPI_BATCH;
    constexpr bool PRELOAD = (KIND == EPI_F32_RES || KIND == EPI_F32_ACC || KIND == EPI_B16_DGELU);
#pragma unroll 1
    for (int it0 = 0; it0 < ITERS; it0 += BATCH) {
        float pre[BATCH][4];
        if (PRELOAD) {
            if (it0 == 0 && use_early) {  // the first batch was loaded before the accumulator was waited for
#pragma unroll
                for (int u = 0; u < BATCH; ++u)
#pragma unroll
                    for (int j = 0; j < 4; ++j) pre[u][j] = early[u][j];
            } else {
                epilogue_preload<BN, KIND>(p, lane, n0, gm0, rows_here, it0, pre);
            }
        }
#pragma unroll
        for (int u = 0; u < BATCH; ++u) {
            const int it = it0 + u;
        const int rr = it * ROWS_PER_IT + rsub;
        if (rr >= rows_here || valid <= 0) continue;
        const int64_t gm = gm0 + rr;
        const float4 t = *reinterpret_cast<const float4*>(stile + rr * PITCH + col);
        float v[4] = {t.x + bias[0], t.y + bias[1], t.z + bias[2], t.w + bias[3]};
        if (KIND == EPI_PARTIAL) {
            store4_f32(p.partial + (int64_t)zsplit * p.partial_stride + gm * p.N + n, v, vpart, valid);
        } else if (KIND == EPI_F32) {
            store4_f32(p.c32 + gm * p.ldc32 + n, v, v32, valid);
        } else if (KIND == EPI_F32_RES) {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] += pre[u][j];
            store4_f32(p.c32 + gm * p.ldc32 + n, v, v32, valid);
        } else if (KIND == EPI_F32_ATOMIC) {
            float* o = p.c32 + gm * p.ldc32 + n;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (j < valid) atomicAdd(o + j, v[j]);
        } else if (KIND == EPI_F32_ACC) {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] += pre[u][j];
            store4_f32(p.c32 + gm * p.ldc32 + n, v, v32, valid);
        } else if (KIND == EPI_B16) {
            store4_b16(p.c16 + gm * p.ldc16 + n, v, v16, valid);
        } else if (KIND == EPI_B16_GELU) {
            float dg[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) gelu_erf_both(v[j], v[j], dg[j]);
            if (p.aux16) store4_b16(p.aux16 + gm * p.ldaux + n, dg, vaux, valid);
            store4_b16(p.c16 + gm * p.ldc16 + n, v, v16, valid);
        } else if (KIND == EPI_B16_DGELU) {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] *= pre[u][j];
            store4_b16(p.c16 + gm * p.ldc16 + n, v, v16, valid);
        } else {  // EPI_GENERIC
            if (p.act == 1) {
                float dg[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) gelu_erf_both(v[j], v[j], dg[j]);
                if (p.aux16) store4_b16(p.aux16 + gm * p.ldaux + n, dg, vaux, valid);
            } else if (p.act == 2) {
                float a[4];
                load4_b16(p.aux16 + gm * p.ldaux + n, a, vaux, valid);
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] *= a[j];
            }
            if (p.drop.threshold != 0) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    v[j] = drop_keep(p.drop, (uint64_t)gm * (uint64_t)p.N + (uint64_t)(n + j)) ? v[j] * p.drop.inv_keep : 0.f;
            }
            if (p.residual) {
                float r[4];
                load4_f32(p.residual + gm * p.ldr + n, r, vres, valid);
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] += r[j];
            }
            if (p.c32) {
                if (p.accumulate) {
                    float r[4];
                    load4_f32(p.c32 + gm * p.ldc32 + n, r, v32, valid);
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] += r[j];
                }
                store4_f32(p.c32 + gm * p.ldc32 + n, v, v32, valid);
            }
            if (p.c16) store4_b16(p.c16 + gm * p.ldc16 + n, v, v16, valid);
        }
        }
    }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(TC_THREADS, 2)
    gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcParams p) {
    constexpr int A_STAGE = TC_BM * TC_BK * 2;  // 16 KB
    constexpr int B_STAGE = BN * TC_BK * 2;
    constexpr uint32_t STAGE_BYTES = A_STAGE + B_STAGE;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;
    uint8_t* sB = smem + TC_STAGES * A_STAGE;
    uint64_t* full_bar = (uint64_t*)(sB + TC_STAGES * B_STAGE);
    uint64_t* empty_bar = full_bar + TC_STAGES;
    uint64_t* tmem_full = empty_bar + TC_STAGES;
    uint32_t* tmem_slot = (uint32_t*)(tmem_full + 1);

    pdl_trigger();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * TC_BM, n0 = blockIdx.x * BN;
    int m_eff = p.M, num_kb = p.num_kb;
    if (p.dyn_rows != nullptr) {  // block-uniform; the count comes from an earlier kernel, so no prologue overlap here
        pdl_wait();
        const int v = max(0, __ldg(p.dyn_rows) - p.dyn_off);
        if (p.dyn_axis == 0) m_eff = min(m_eff, v);
        else num_kb = min(num_kb, (v + TC_BK - 1) / TC_BK);
        if (m0 >= m_eff || blockIdx.z * p.kb_per_split >= num_kb) return;  // nothing to do: leave before any set-up
    }

    const int kb_begin = blockIdx.z * p.kb_per_split;
    const int kb_end = min(num_kb, kb_begin + p.kb_per_split);
    const int nkb = kb_end - kb_begin;

    // one k-block of both operands into ring stage i % TC_STAGES
    auto load_kb = [&](int i) {
        const int s = i % TC_STAGES;
        const uint32_t ph = (uint32_t)(i / TC_STAGES) & 1u;
        mbar_wait(&empty_bar[s], ph ^ 1u);
        mbar_expect_tx(&full_bar[s], STAGE_BYTES);
        const int k0 = (kb_begin + i) * TC_BK;
        uint8_t* a_dst = sA + s * A_STAGE;
        uint8_t* b_dst = sB + s * B_STAGE;
        if (!A_MN) {
            tma_load_2d(a_dst, &tmA, k0, m0, &full_bar[s]);
        } else {
#pragma unroll
            for (int j = 0; j < TC_BM / 64; ++j) tma_load_2d(a_dst + j * 8192, &tmA, m0 + j * 64, k0, &full_bar[s]);
        }
        if (!B_MN) {
            tma_load_2d(b_dst, &tmB, k0, n0, &full_bar[s]);
        } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_2d(b_dst + j * 8192, &tmB, n0 + j * 64, k0, &full_bar[s]);
        }
    };
    const int early_kb = min(nkb, TC_STAGES);

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
        // fill the ring before the TMEM allocation and the block barrier: the first loads' latency
        // then overlaps the rest of the prologue (the operands come from earlier kernels, hence the wait)
        pdl_wait();
        for (int i = 0; i < early_kb; ++i) load_kb(i);
    }
    if (warp == 1) tmem_alloc(tmem_slot, BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // prologue done; from here on the kernel reads / writes what earlier kernels produced

    if (warp == 0) {
        if (lane == 0) {
            for (int i = early_kb; i < nkb; ++i) load_kb(i);
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(BN, A_MN, B_MN);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % TC_STAGES;
                const uint32_t ph = (uint32_t)(i / TC_STAGES) & 1u;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(sA + s * A_STAGE);
                const uint32_t b_addr = smem_u32(sB + s * B_STAGE);
#pragma unroll
                for (int k = 0; k < TC_BK / 16; ++k) {
                    // K-major: step 16 elements (32 B) inside the 128-byte swizzle row;
                    // MN-major: step 16 k-rows of 128 B (two 8-row swizzle atoms)
                    const uint64_t da = A_MN ? make_smem_desc(a_addr + k * 2048, 8192, 1024) : make_smem_desc(a_addr + k * 32, 16, 1024);
                    const uint64_t db = B_MN ? make_smem_desc(b_addr + k * 2048, 8192, 1024) : make_smem_desc(b_addr + k * 32, 16, 1024);
                    umma_bf16(tmem_base, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);  // frees the ring slot once these MMAs have read it
            }
            umma_commit(tmem_full);  // accumulator complete
        }
        __syncwarp();
    } else {
        // ---------------- epilogue: TMEM lane quadrant = warp id mod 4
        // tcgen05.ld hands each thread one accumulator ROW; the rows are staged in shared memory
        // (aliasing the now idle operand ring) and read back with lane = 4 consecutive columns.
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;  // which half of the tile's columns this warp finishes
        constexpr int HN = BN / 2;
        constexpr int PITCH = HN + 4;
        // Flavours that read a global operand per output row: the epilogue warps have nothing to do while the MMAs run,
        // so the first batch of those loads (all of them for 64-wide tiles) is issued BEFORE waiting for the accumulator
        // — the operand (a residual, the saved GELU' of the forward pass) usually comes from HBM, and that round trip
        // used to sit between the last MMA and the first store (dX with GELU': 10.9 us against 5.5 us for a plain dX).
        float early[EPI_BATCH][4];
        bool have_early = false;
        {
            const int64_t gm0e = (int64_t)m0 + q * 32;
            const int rows_e = (int)min((int64_t)32, (int64_t)m_eff - gm0e);
            const int nhe = n0 + half * HN;
            if constexpr (BN == 64) {  // (128-wide tiles: the 32 extra live registers across the accumulator copy would spill)
                if (p.epi_kind == EPI_F32_RES) { epilogue_preload<HN, EPI_F32_RES>(p, lane, nhe, gm0e, rows_e, 0, early); have_early = true; }
                else if (p.epi_kind == EPI_B16_DGELU) { epilogue_preload<HN, EPI_B16_DGELU>(p, lane, nhe, gm0e, rows_e, 0, early); have_early = true; }
            }
        }
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        float* stile = reinterpret_cast<float*>(smem) + (half * 4 + q) * (32 * PITCH);
#pragma unroll
        for (int c0 = 0; c0 < HN; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * HN + c0), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(stile + lane * PITCH + c0 + j) =
                    make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        }
        __syncwarp();
        const int64_t gm0 = (int64_t)m0 + q * 32;
        const int rows_here = (int)min((int64_t)32, (int64_t)m_eff - gm0);  // may be <= 0
        if (p.dyn_rows != nullptr && p.dyn_axis == 0 && p.c16 != nullptr && rows_here < 32) {
            // rows between the device-side count and the end of this 128-row tile: zero in the bf16 output, so a later
            // product over the row axis can read them as operand padding (its k-blocks end at a multiple of 64 rows)
            const int first = max(rows_here, 0);
            const int last = (int)min((int64_t)32, (int64_t)p.M - gm0);
            const int nh0 = n0 + half * (BN / 2);
            for (int rr = first; rr < last; ++rr)
                for (int cc = lane; cc < BN / 2; cc += 32)
                    if (nh0 + cc < p.N) p.c16[(gm0 + rr) * p.ldc16 + nh0 + cc] = __float2bfloat16(0.f);
        }
        const int nh = n0 + half * HN;
        switch (p.epi_kind) {
            case EPI_F32: epilogue_rows<HN, EPI_F32>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            case EPI_F32_RES: epilogue_rows<HN, EPI_F32_RES>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            case EPI_F32_ACC: epilogue_rows<HN, EPI_F32_ACC>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            case EPI_B16: epilogue_rows<HN, EPI_B16>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            case EPI_B16_GELU: epilogue_rows<HN, EPI_B16_GELU>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            case EPI_B16_DGELU: epilogue_rows<HN, EPI_B16_DGELU>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            case EPI_PARTIAL: epilogue_rows<HN, EPI_PARTIAL>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            case EPI_F32_ATOMIC: epilogue_rows<HN, EPI_F32_ATOMIC>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
            default: epilogue_rows<HN, EPI_GENERIC>(p, stile, lane, nh, gm0, rows_here, blockIdx.z, early, have_early); break;
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, BN);
    }
}

// ------------------------------------------------------------------ host side
// 2-D bf16 tensor map over a row-major (rows, cols) matrix with leading dimension ld (elements);
// box = box_rows x 64 columns, 128-byte swizzle, zero fill out of bounds.
int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled entry point not available from the driver");
        return PIO_ERR_CUDA;
    }
    PIO_CHECK_ARG((((uintptr_t)base) & 15) == 0, "bf16 GEMM operand base %p is not 16-byte aligned", base);
    PIO_CHECK_ARG(ld % 8 == 0, "bf16 GEMM operand leading dimension %lld is not a multiple of 8 elements", (long long)ld);
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows %lld cols %lld ld %lld)", (int)r, (long long)rows,
                  (long long)cols, (long long)ld);
        return PIO_ERR_CUDA;
    }
    return PIO_OK;
}

static void tc_choose(int64_t m, int64_t n, int64_t k, bool splittable, int* bn, int* splits, int* kb_per) {
    const int sms = num_sms();
    // 64-wide tiles when 128-wide ones would leave most SMs idle (the latent GEMMs: 4096 x 256 outputs)
    static const bool narrow = getenv("PIO_TC_NO_NARROW") == nullptr;
    *bn = (n <= 64 || (narrow && cdiv(m, TC_BM) * cdiv(n, 128) < sms / 2)) ? 64 : 128;
    const int64_t tiles = cdiv(m, TC_BM) * cdiv(n, *bn);
    const int num_kb = (int)cdiv(k, TC_BK);
    int s = 1;
    // k-blocks per split: 8 (512 rows of the contraction).  Measured on the cfg2 step, where the weight-gradient
    // products (256 x 256 outputs over 4096 rows, on the side stream) are the longer of the two backward chains:
    // 16 -> 3.73 ms, 8 -> 3.61 ms, 4 -> 3.73 ms, 2 -> 3.74 ms (more, smaller CTAs start to take SMs from the main chain).
    static const int min_kb = getenv("PIO_TC_SPLIT_MIN_KB") ? atoi(getenv("PIO_TC_SPLIT_MIN_KB")) : 8;
    if (splittable && num_kb >= 16 && tiles < sms) {
        s = (int)std::min<int64_t>(cdiv(sms, tiles), num_kb / min_kb);
        s = std::max(s, 1);
    }
    int per = (int)cdiv(num_kb, s);
    s = (int)cdiv(num_kb, per);
    *splits = s;
    *kb_per = per;
}

template <int BN, bool A_MN, bool B_MN>
static int launch_tc(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, dim3 grid, cudaStream_t st) {
    const size_t smem = (size_t)TC_STAGES * (TC_BM * TC_BK * 2 + BN * TC_BK * 2) + 1024 + 256;
    static std::atomic<bool> configured[64];  // per instantiation; attribute is sticky per function
    int dev_ = 0;
    PIO_CUDA(cudaGetDevice(&dev_));  // the attribute is per function AND per device
    dev_ = (dev_ >= 0 && dev_ < 64) ? dev_ : 0;
    if (!configured[dev_].load()) {
        PIO_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev_].store(true);
    }
    PIO_CUDA(launch_pdl(gemm_tc_kernel<BN, A_MN, B_MN>, grid, dim3(TC_THREADS), smem, st, ta, tb, p));
    count_launch();
    return PIO_OK;
}

int gemm_bf16(const pio_gemm_bf16_desc* d, void* ws, size_t ws_bytes, cudaStream_t st) {
    PIO_CHECK_ARG(d && d->a && d->b && (d->c32 || d->c16), "pio_gemm_bf16: null pointer");
    PIO_CHECK_ARG(d->m >= 0 && d->n >= 0 && d->k >= 1, "pio_gemm_bf16: bad dims");
    PIO_CHECK_ARG(d->act >= 0 && d->act <= 2 && (d->act != 2 || d->aux16), "pio_gemm_bf16: bad act/aux");
    PIO_CHECK_ARG(d->m < (1ll << 31) && d->n < (1ll << 31) && d->k < (1ll << 31), "pio_gemm_bf16: dims exceed 2^31");
    if (d->m == 0 || d->n == 0) return PIO_OK;
    const bool a_mn = !d->a_kmajor, b_mn = !d->b_kmajor;
    const bool dropping = d->dropout_p > 0.f;
    // (a device-side row count rules out the partial-sum reduction pass, which would touch skipped rows)
    const bool splittable = !d->bias && !d->residual && d->act == 0 && d->c32 && !d->c16 && d->ldc32 == d->n && !dropping &&
                            !(d->dyn_rows && d->accumulate != 2);
    int bn, splits, kb_per;
    tc_choose(d->m, d->n, d->k, splittable, &bn, &splits, &kb_per);

    TcParams p{};
    p.M = (int)d->m; p.N = (int)d->n; p.K = (int)d->k;
    p.num_kb = (int)cdiv(d->k, TC_BK);
    p.c32 = d->c32; p.ldc32 = d->ldc32;
    p.c16 = (__nv_bfloat16*)d->c16; p.ldc16 = d->ldc16;
    p.bias = d->bias; p.residual = d->residual; p.ldr = d->ldr;
    p.aux16 = (__nv_bfloat16*)d->aux16; p.ldaux = d->ldaux;
    p.act = d->act; p.accumulate = d->accumulate;
    p.drop = make_drop(d->dropout_p, d->dropout_seed);
    p.partial_stride = d->m * d->n;
    p.dyn_rows = d->dyn_rows; p.dyn_off = (int)d->dyn_offset; p.dyn_axis = d->dyn_axis;
    PIO_CHECK_ARG(!d->dyn_rows || (d->dyn_axis == 0 || (d->dyn_axis == 1 && d->accumulate == 2)),
                  "pio_gemm_bf16: dyn_axis must be 0, or 1 with accumulate == 2");
    PIO_CHECK_ARG(!d->dyn_rows || (d->dyn_offset >= 0 && d->dyn_offset < (1ll << 31)), "pio_gemm_bf16: bad dyn_offset");
    const bool atomic = (d->accumulate == 2);
    if (atomic) PIO_CHECK_ARG(splittable, "pio_gemm_bf16: accumulate=2 (atomic) needs a plain fp32 output without epilogue");
    if (splits > 1 && !atomic) {
        Workspace W(ws, ws_bytes);
        p.partial = W.takef((size_t)splits * (size_t)p.partial_stride);
        if (!p.partial) {
            splits = 1;
            kb_per = p.num_kb;
        }
    }
    p.splits = splits;
    p.kb_per_split = kb_per;
    {
        const bool o32 = d->c32 != nullptr, o16 = d->c16 != nullptr;
        int kind = EPI_GENERIC;
        if (atomic) kind = EPI_F32_ATOMIC;
        else if (splits > 1) kind = EPI_PARTIAL;
        else if (dropping) kind = EPI_GENERIC;
        else if (o32 && !o16 && d->act == 0 && !d->residual && !d->accumulate) kind = EPI_F32;
        else if (o32 && !o16 && d->act == 0 && d->residual && !d->accumulate) kind = EPI_F32_RES;
        else if (o32 && !o16 && d->act == 0 && !d->residual && d->accumulate) kind = EPI_F32_ACC;
        else if (!o32 && o16 && d->act == 0 && !d->residual) kind = EPI_B16;
        else if (!o32 && o16 && d->act == 1 && !d->residual) kind = EPI_B16_GELU;
        else if (!o32 && o16 && d->act == 2 && !d->residual && !d->bias) kind = EPI_B16_DGELU;
        p.epi_kind = kind;
    }

    // tensor maps: K-major operand = (rows = m|n, cols = k); MN-major operand = (rows = k, cols = m|n)
    CUtensorMap ta, tb;
    if (!a_mn) PIO_TRY(make_tmap_bf16(&ta, d->a, d->m, d->k, d->lda, TC_BM));
    else PIO_TRY(make_tmap_bf16(&ta, d->a, d->k, d->m, d->lda, TC_BK));
    if (!b_mn) PIO_TRY(make_tmap_bf16(&tb, d->b, d->n, d->k, d->ldb, bn));
    else PIO_TRY(make_tmap_bf16(&tb, d->b, d->k, d->n, d->ldb, TC_BK));

    dim3 grid((unsigned)cdiv(d->n, bn), (unsigned)cdiv(d->m, TC_BM), (unsigned)splits);
    PIO_CHECK_ARG(grid.y <= 65535 && grid.z <= 65535, "pio_gemm_bf16: grid too large");
    int rc;
#define PIO_TC(BNV)                                                        \
    do {                                                                   \
        if (!a_mn && !b_mn) rc = launch_tc<BNV, false, false>(ta, tb, p, grid, st);      \
        else if (!a_mn && b_mn) rc = launch_tc<BNV, false, true>(ta, tb, p, grid, st);   \
        else if (a_mn && !b_mn) rc = launch_tc<BNV, true, false>(ta, tb, p, grid, st);   \
        else rc = launch_tc<BNV, true, true>(ta, tb, p, grid, st);                        \
    } while (0)
    if (bn == 64) PIO_TC(64);
    else PIO_TC(128);
#undef PIO_TC
    PIO_TRY(rc);
    if (splits > 1 && !atomic)
        PIO_TRY(reduce_partials(p.partial, splits, p.partial_stride, d->c32, d->m * d->n, nullptr, 0, d->accumulate, st));
    return PIO_OK;
}

// fp32 -> bf16 with a (possibly padded) destination pitch; pad columns are zero-filled
__global__ void cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t rows, int cols,
                                 int64_t ld_src, int64_t ld_dst) {
    pdl_trigger();
    pdl_wait();
    const int64_t total = rows * ld_dst;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / ld_dst;
        const int c = (int)(i - r * ld_dst);
        dst[i] = __float2bfloat16(c < cols ? src[r * ld_src + c] : 0.f);
    }
}

int cast_bf16(const float* src, void* dst, int64_t rows, int cols, int64_t ld_src, int64_t ld_dst, cudaStream_t st) {
    PIO_CHECK_ARG(src && dst && rows >= 0 && cols > 0 && ld_dst >= cols && ld_src >= cols, "pio_cast_bf16: bad arguments");
    if (rows == 0) return PIO_OK;
    const int64_t total = rows * ld_dst;
    const int blocks = (int)std::min<int64_t>(cdiv(total, 256 * 4), (int64_t)num_sms() * 16);
    PIO_CUDA(launch_pdl(cast_bf16_kernel, dim3(std::max(blocks, 1)), dim3(256), 0, st, src, (__nv_bfloat16*)dst, rows, cols, ld_src,
                        ld_dst));
    count_launch();
    return PIO_OK;
}

// one launch for a table of matrices: blockIdx.y = table entry, blockIdx.x strides over its elements
__global__ void cast_bf16_multi_kernel(const pio_cast_item* __restrict__ items) {
    pdl_trigger();
    pdl_wait();
    const pio_cast_item it = items[blockIdx.y];
    const float* __restrict__ src = it.src;
    __nv_bfloat16* __restrict__ dst = (__nv_bfloat16*)it.dst;
    const int64_t total = it.rows * it.cols;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (it.cols == it.ld_dst && (it.cols & 3) == 0) {  // dense: 4 elements per thread per trip
        const int64_t n4 = total >> 2;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += stride) {
            const float4 v = reinterpret_cast<const float4*>(src)[i];
            __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            uint2 o;
            o.x = *reinterpret_cast<uint32_t*>(&lo);
            o.y = *reinterpret_cast<uint32_t*>(&hi);
            reinterpret_cast<uint2*>(dst)[i] = o;
        }
        return;
    }
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += stride) {
        const int64_t r = i / it.cols;
        const int c = (int)(i - r * it.cols);
        dst[r * it.ld_dst + c] = __float2bfloat16(src[i]);
    }
}

int cast_bf16_multi(const pio_cast_item* items, int count, cudaStream_t st) {
    PIO_CHECK_ARG(count >= 0 && (items || count == 0), "pio_cast_bf16_multi: bad arguments");
    if (count == 0) return PIO_OK;
    PIO_CHECK_ARG(count <= 65535, "pio_cast_bf16_multi: at most 65535 entries per call");
    PIO_CUDA(launch_pdl(cast_bf16_multi_kernel, dim3(64, count), dim3(256), 0, st, items));
    count_launch();
    return PIO_OK;
}

}  // namespace pio

using namespace pio;

extern "C" {

int pio_cast_bf16_multi(const pio_cast_item* items, int count, pio_stream_t stream) {
    return cast_bf16_multi(items, count, (cudaStream_t)stream);
}

size_t pio_gemm_bf16_ws_bytes(int64_t m, int64_t n, int64_t k) {
    int bn, splits, kb_per;
    tc_choose(m, n, k, true, &bn, &splits, &kb_per);
    if (splits <= 1) return 256;
    return align_up((size_t)splits * (size_t)(m * n) * sizeof(float), 256) + 256;
}

int pio_gemm_bf16(const pio_gemm_bf16_desc* d, void* ws, size_t ws_bytes, pio_stream_t stream) {
    return gemm_bf16(d, ws, ws_bytes, (cudaStream_t)stream);
}

int pio_cast_bf16(const float* src, void* dst, int64_t rows, int cols, int64_t ld_src, int64_t ld_dst, pio_stream_t stream) {
    return cast_bf16(src, dst, rows, cols, ld_src, ld_dst, (cudaStream_t)stream);
}

}  // extern "C"
