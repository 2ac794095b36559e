// PIO_MODE_BF16 pipelines of the fused operators: every contraction runs on the tcgen05 GEMM
// (pio_gemm_tc.cu) with bf16 operands and fp32 accumulation; LayerNorm statistics, the residual
// stream, softmax and all parameter gradients stay fp32.  Matrix weights arrive as bf16 shadows
// (row pitch padded to 8 elements) in the same pio_*_weights structs; LayerNorm parameters and
// biases stay fp32.  Saved activations that only feed GEMMs are kept in bf16.
#include <stdlib.h>

#include <map>
#include <mutex>
#include <utility>
#include <vector>

#include "pio_ops_common.cuh"

namespace pio {
namespace bf {

using bf16 = __nv_bfloat16;

struct ByteCarver {
    char* base;
    size_t off;
    void* take(size_t bytes) {
        void* r = base ? base + off : nullptr;
        off += align_up(bytes, 256);
        return r;
    }
    float* f32(size_t n) { return (float*)take(n * 4); }
    bf16* b16(size_t n) { return (bf16*)take(n * 2); }
};

#define PIO_TAKE_BYTES(type, var, ctx, bytes, who)                                        \
    type* var = (type*)(ctx).W.take((size_t)(bytes));                                     \
    if (!var) {                                                                           \
        set_error("%s: workspace too small (%zu bytes given)", who, (ctx).W.size);       \
        return PIO_ERR_WORKSPACE;                                                         \
    }

static int tc_gemm(Ctx& c, const pio_gemm_bf16_desc& d) { return gemm_bf16(&d, c.rest(), c.rest_bytes(), c.st); }

// ---------------------------------------------------------------- side stream for weight gradients
// In the backward pass the weight / bias gradients (dW = dY^T X, db = colsum dY) hang off the critical
// dX chain: nothing downstream in the same operator reads them.  With atomic accumulation (accumulate
// == 2) they need no scratch either, so they are forked onto a per-device auxiliary stream and joined at
// the end of the operator; the small, latency-bound kernels of the two chains then overlap on the GPU
// (under stream capture the fork/join becomes two parallel branches of the graph).
struct SideRes {
    cudaStream_t stream = nullptr;   // weight-gradient GEMMs
    cudaStream_t stream2 = nullptr;  // bias-gradient column sums (beside the GEMMs: one chain was the longer of the operator's two)
    cudaEvent_t fork = nullptr, join = nullptr, join2 = nullptr;
};
// One set of side streams / events per (device, caller stream): two host threads (or two streams) running backward
// operators on the same GPU must not share fork / join events — a second caller's record would overwrite the first's and
// let side-stream kernels start before their producers.
static SideRes* side_res(cudaStream_t caller) {
    static std::map<std::pair<int, cudaStream_t>, SideRes*> table;
    static std::mutex mu;
    static const bool disabled = getenv("PIO_NO_SIDE_STREAM") != nullptr;
    if (disabled) return nullptr;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    SideRes*& slot = table[std::make_pair(dev, caller)];
    if (slot == nullptr) {
        if (table.size() > 256) {  // (a caller that keeps making streams: stay on its own stream rather than leak)
            table.erase(std::make_pair(dev, caller));
            return nullptr;
        }
        SideRes* r = new SideRes();
        if (cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&r->stream2, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&r->fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&r->join2, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&r->join, cudaEventDisableTiming) != cudaSuccess) {
            delete r;
            table.erase(std::make_pair(dev, caller));
            return nullptr;
        }
        slot = r;
    }
    return slot;
}

struct Side {
    Ctx& c;
    SideRes* r;
    bool used = false;
    Side(Ctx& ctx, int acc) : c(ctx), r(acc == 2 ? side_res(ctx.st) : nullptr) {}
    bool active() const { return r != nullptr; }
    cudaStream_t stream() const { return r ? r->stream : c.st; }
    cudaStream_t stream2() const { return r ? r->stream2 : c.st; }
    // everything enqueued on the main stream so far becomes visible to the side streams
    int fork() {
        if (!r) return PIO_OK;
        PIO_CUDA(cudaEventRecord(r->fork, c.st));
        PIO_CUDA(cudaStreamWaitEvent(r->stream, r->fork, 0));
        PIO_CUDA(cudaStreamWaitEvent(r->stream2, r->fork, 0));
        used = true;
        return PIO_OK;
    }
    // the main stream waits for everything enqueued on the side streams
    int join() {
        if (!r || !used) return PIO_OK;
        PIO_CUDA(cudaEventRecord(r->join, r->stream));
        PIO_CUDA(cudaStreamWaitEvent(c.st, r->join, 0));
        PIO_CUDA(cudaEventRecord(r->join2, r->stream2));
        PIO_CUDA(cudaStreamWaitEvent(c.st, r->join2, 0));
        used = false;
        return PIO_OK;
    }
};

// Deferred join (pio_set_side_join_deferred / pio_side_join): an operator that only has parameter-gradient kernels
// left on the side streams may return without making the caller's stream wait for them, so that they overlap with the
// NEXT operator's kernels; the caller then joins (pio_side_join) before anything on its stream reads those gradients
// or reuses the operator's workspace and saved buffers.  Per host thread; off unless the caller opts in.
static thread_local int t_defer_side_join = 0;
void set_side_join_deferred(int on) { t_defer_side_join = on ? 1 : 0; }
int side_join(cudaStream_t st) {
    SideRes* r = side_res(st);
    if (!r) return PIO_OK;
    PIO_CUDA(cudaEventRecord(r->join, r->stream));
    PIO_CUDA(cudaStreamWaitEvent(st, r->join, 0));
    PIO_CUDA(cudaEventRecord(r->join2, r->stream2));
    PIO_CUDA(cudaStreamWaitEvent(st, r->join2, 0));
    return PIO_OK;
}

// y = act(x16 W16^T + bias) (+ residual)
static int tc_xwT(Ctx& c, const bf16* x, int64_t ldx, const bf16* w, int64_t ldw, int64_t rows, int64_t n, int64_t k,
                  float* c32, int64_t ldc32, bf16* c16, int64_t ldc16, const float* bias, const float* res, int64_t ldr,
                  int act, bf16* aux, int64_t ldaux, float drop_p = 0.f, uint64_t drop_seed = 0) {
    pio_gemm_bf16_desc d{};
    d.dropout_p = drop_p; d.dropout_seed = drop_seed;
    d.a = x; d.b = w; d.m = rows; d.n = n; d.k = k; d.lda = ldx; d.ldb = ldw; d.a_kmajor = 1; d.b_kmajor = 1;
    d.c32 = c32; d.ldc32 = ldc32; d.c16 = c16; d.ldc16 = ldc16;
    d.bias = bias; d.residual = res; d.ldr = ldr; d.act = act; d.aux16 = aux; d.ldaux = ldaux;
    d.dyn_rows = c.dyn; d.dyn_axis = 0;
    return tc_gemm(c, d);
}
// dx (+)= dy16[rows, nout] W16[nout, kin]   (B read MN-major, in place)
static int tc_dyW(Ctx& c, const bf16* dy, int64_t lddy, const bf16* w, int64_t ldw, int64_t rows, int64_t nout, int64_t kin,
                  float* c32, int64_t ldc32, bf16* c16, int64_t ldc16, int act, bf16* aux, int64_t ldaux, int accumulate) {
    pio_gemm_bf16_desc d{};
    d.a = dy; d.b = w; d.m = rows; d.n = kin; d.k = nout; d.lda = lddy; d.ldb = ldw; d.a_kmajor = 1; d.b_kmajor = 0;
    d.c32 = c32; d.ldc32 = ldc32; d.c16 = c16; d.ldc16 = ldc16; d.act = act; d.aux16 = aux; d.ldaux = ldaux;
    d.accumulate = accumulate;
    d.dyn_rows = c.dyn; d.dyn_axis = 0;
    return tc_gemm(c, d);
}
// dw[nout, kin] (+)= dy16[rows, nout]^T x16[rows, kin]   (both read MN-major, in place; split-K)
static int tc_dW(Ctx& c, const bf16* dy, int64_t lddy, const bf16* x, int64_t ldx, float* dw, int64_t rows, int64_t nout,
                 int64_t kin, int accumulate, cudaStream_t st = nullptr) {
    pio_gemm_bf16_desc d{};
    d.a = dy; d.b = x; d.m = nout; d.n = kin; d.k = rows; d.lda = lddy; d.ldb = ldx; d.a_kmajor = 0; d.b_kmajor = 0;
    d.c32 = dw; d.ldc32 = kin; d.accumulate = accumulate;
    if (c.dyn != nullptr) {  // contraction over the live rows only (the operands are zero up to the next multiple of 64)
        PIO_CHECK_ARG(accumulate == 2, "a device-side row count needs atomic weight-gradient accumulation (accumulate = 2)");
        d.dyn_rows = c.dyn; d.dyn_axis = 1;
    }
    if (st == nullptr || st == c.st) return tc_gemm(c, d);
    return gemm_bf16(&d, nullptr, 0, st);  // side stream: atomic accumulation, no scratch
}
static int side_colsum(Ctx& c, const Side& sd, const bf16* src, int64_t ld, float* out, int acc, int64_t rows, int cols) {
    if (sd.active()) return colsum_bf16(src, ld, out, acc, rows, cols, nullptr, 0, sd.stream2(), c.dyn);
    return colsum_bf16(src, ld, out, acc, rows, cols, c.rest(), c.rest_bytes(), c.st, c.dyn);
}

static size_t scratch_bytes(int64_t rows, int64_t nout, int64_t kin) {
    return std::max(std::max(pio_gemm_bf16_ws_bytes(nout, kin, rows), colsum_ws_bytes(rows, (int)std::max(nout, kin))),
                    pio_layernorm_bwd_ws_bytes(rows, (int)kin)) + 1024;
}

// gradient entering the branch of a Residual with dropout (same mask as the forward epilogue)
static int branch_grad(Ctx& c, const float* dy, int64_t count, int residual, const float** out, const char* who) {
    *out = dy;
    if (!(residual && c.drop_p > 0.f)) return PIO_OK;
    float* tmp = c.W.takef((size_t)count);
    if (!tmp) {
        set_error("%s: workspace too small (%zu bytes given)", who, c.W.size);
        return PIO_ERR_WORKSPACE;
    }
    PIO_TRY(dropout_mask(dy, tmp, count, c.res_drop(residual), c.st));
    *out = tmp;
    return PIO_OK;
}

static int need_mult8(int v, const char* what, const char* who) {
    PIO_CHECK_ARG(v % 8 == 0, "%s: %s = %d must be a multiple of 8 in bf16 mode", who, what, v);
    return PIO_OK;
}

// ---------------------------------------------------------------- mlp
struct MlpSaved {
    float* stats;
    bf16 *h0, *h1, *g;  // LayerNorm output, GELU'(pre-activation) (written by the GELU epilogue), GELU output
    size_t total;
    MlpSaved(void* p, int64_t rows, int ch) {
        ByteCarver k{(char*)p, 0};
        stats = k.f32((size_t)rows * 2);
        h0 = k.b16((size_t)rows * ch);
        h1 = k.b16((size_t)rows * ch);
        g = k.b16((size_t)rows * ch);
        total = k.off;
    }
};
size_t mlp_saved_bytes(int64_t rows, int ch) { return MlpSaved(nullptr, rows, ch).total + 256; }
size_t mlp_ws_bytes(int64_t rows, int ch) {
    return 2 * align_up((size_t)rows * ch * 2, 256) + 2 * align_up((size_t)rows * ch * 4, 256) + scratch_bytes(rows, ch, ch);
}

int mlp_fwd(Ctx& c, const float* x, float* y, const pio_mlp_weights* w, void* saved, int64_t rows, int ch, int residual) {
    PIO_TRY(need_mult8(ch, "channels", "pio_mlp_fwd"));
    MlpSaved s(saved, rows, ch);
    PIO_TRY(layernorm_fwd_bf16(x, w->ln_w, w->ln_b, s.h0, ch, s.stats, rows, ch, LN_EPS, c.st, c.dyn));
    PIO_TRY(tc_xwT(c, s.h0, ch, (const bf16*)w->w1, ch, rows, ch, ch, nullptr, 0, s.g, ch, w->b1, nullptr, 0, 1, s.h1, ch));
    PIO_TRY(tc_xwT(c, s.g, ch, (const bf16*)w->w2, ch, rows, ch, ch, y, ch, nullptr, 0, w->b2, residual ? x : nullptr, ch, 0,
                   nullptr, 0, residual ? c.drop_p : 0.f, c.drop_seed + 1));
    return PIO_OK;
}

int mlp_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_mlp_weights* w, const void* saved,
            const pio_mlp_grads* g, int acc, int64_t rows, int ch, int residual, const bf16* dy16_in, bf16* dx16_out) {
    PIO_TRY(need_mult8(ch, "channels", "pio_mlp_bwd"));
    MlpSaved s((void*)saved, rows, ch);
    const size_t mark = c.W.off;
    PIO_TAKE_BYTES(bf16, dy16_own, c, (size_t)rows * ch * 2, "pio_mlp_bwd");
    PIO_TAKE_BYTES(bf16, dh1, c, (size_t)rows * ch * 2, "pio_mlp_bwd");
    PIO_TAKE_BYTES(float, dh0, c, (size_t)rows * ch * 4, "pio_mlp_bwd");
    Side side(c, acc);
    const bf16* dy16 = dy16_own;
    if (dy16_in != nullptr && !(residual && c.drop_p > 0.f) && side.active()) {
        // the producer of dy (the previous operator's LayerNorm backward) already wrote it as bf16: no cast pass on the
        // critical chain, and the bias-gradient column sums move to the side stream with the weight gradient
        dy16 = dy16_in;
        PIO_TRY(side.fork());
        PIO_TRY(side_colsum(c, side, dy16, ch, g->b2, acc, rows, ch));
    } else {
        const float* dyb;
        PIO_TRY(branch_grad(c, dy, rows * ch, residual, &dyb, "pio_mlp_bwd"));
        PIO_TRY(cast_colsum_f32(dyb, ch, dy16_own, ch, g->b2, acc, rows, ch, c.rest(), c.rest_bytes(), c.st, c.dyn));
        PIO_TRY(side.fork());
    }
    PIO_TRY(tc_dW(c, dy16, ch, s.g, ch, g->w2, rows, ch, ch, acc, side.stream()));
    PIO_TRY(tc_dyW(c, dy16, ch, (const bf16*)w->w2, ch, rows, ch, ch, nullptr, 0, dh1, ch, 2, s.h1, ch, 0));
    PIO_TRY(side.fork());
    PIO_TRY(side_colsum(c, side, dh1, ch, g->b1, acc, rows, ch));
    PIO_TRY(tc_dW(c, dh1, ch, s.h0, ch, g->w1, rows, ch, ch, acc, side.stream()));
    PIO_TRY(tc_dyW(c, dh1, ch, (const bf16*)w->w1, ch, rows, ch, ch, dh0, ch, nullptr, 0, 0, nullptr, 0, 0));
    PIO_TRY(layernorm_bwd(dh0, x, s.stats, w->ln_w, residual ? dy : nullptr, dx, g->ln_w, g->ln_b, acc, rows, ch,
                          c.rest(), c.rest_bytes(), c.st, c.dyn, dx16_out));
    PIO_TRY(side.join());
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- kv projection
struct KvSaved {
    float* stats;
    bf16* h0;
    size_t total;
    KvSaved(void* p, int64_t rows, int ckv) {
        ByteCarver k{(char*)p, 0};
        stats = k.f32((size_t)rows * 2);
        h0 = k.b16((size_t)rows * pad8(ckv));
        total = k.off;
    }
};
size_t kv_saved_bytes(int64_t rows, int ckv) { return KvSaved(nullptr, rows, ckv).total + 256; }
size_t kv_ws_bytes(int64_t rows, int ckv, int e) {
    return align_up((size_t)rows * 2 * e * 2, 256) + align_up((size_t)rows * ckv * 4, 256) + scratch_bytes(rows, 2 * e, ckv);
}

static bool packed16(const void* wk, const void* wv, int e, int64_t ldw) {
    return (const char*)wv == (const char*)wk + (size_t)e * ldw * 2;
}

int kv_project_fwd(Ctx& c, const float* x_kv, float* kv, const pio_attn_weights* w, void* saved, int64_t rows, int ckv, int e) {
    PIO_TRY(need_mult8(e, "embed dim", "pio_kv_project_fwd"));
    KvSaved s(saved, rows, ckv);
    const int64_t ckp = pad8(ckv);
    bf16* kv16 = reinterpret_cast<bf16*>(kv);  // bf16 mode: the kv tensor is bf16 (B, M, 2E)
    if (packed16(w->wk, w->wv, e, ckp) && w->bv == w->bk + e && c.dyn == nullptr && ln_linear_supported(rows, 2 * e, ckv, ckp, 2 * e, ckp)) {
        // kv_norm + K/V in-projection in one persistent kernel (pio_logits_tc.cu): x_kv is read once, the normalised operand
        // goes straight into the product's shared-memory tile (its bf16 copy and the statistics are still saved)
        return ln_linear(x_kv, w->kv_ln_w, w->kv_ln_b, LN_EPS, w->wk, ckp, w->bk, kv16, 2 * e, s.h0, ckp, s.stats, rows, 2 * e, c.st);
    }
    PIO_TRY(layernorm_fwd_bf16(x_kv, w->kv_ln_w, w->kv_ln_b, s.h0, ckp, s.stats, rows, ckv, LN_EPS, c.st));
    if (packed16(w->wk, w->wv, e, ckp) && w->bv == w->bk + e) {
        PIO_TRY(tc_xwT(c, s.h0, ckp, (const bf16*)w->wk, ckp, rows, 2 * e, ckv, nullptr, 0, kv16, 2 * e, w->bk, nullptr, 0, 0,
                       nullptr, 0));
    } else {
        PIO_TRY(tc_xwT(c, s.h0, ckp, (const bf16*)w->wk, ckp, rows, e, ckv, nullptr, 0, kv16, 2 * e, w->bk, nullptr, 0, 0, nullptr, 0));
        PIO_TRY(tc_xwT(c, s.h0, ckp, (const bf16*)w->wv, ckp, rows, e, ckv, nullptr, 0, kv16 + e, 2 * e, w->bv, nullptr, 0, 0,
                       nullptr, 0));
    }
    return PIO_OK;
}

int kv_project_bwd(Ctx& c, const float* dkv, const float* x_kv, float* dx_kv, const pio_attn_weights* w, const void* saved,
                   const pio_attn_grads* g, int acc, int64_t rows, int ckv, int e) {
    PIO_TRY(need_mult8(e, "embed dim", "pio_kv_project_bwd"));
    // PIO_ACC_DX: dx_kv already holds the other consumer's share; LayerNorm's backward adds it in as its residual term
    // (every thread reads the elements it then writes, so the buffer may be both)
    const float* dx_prev = (acc & PIO_ACC_DX) ? dx_kv : nullptr;
    acc &= 3;
    KvSaved s((void*)saved, rows, ckv);
    const int64_t ckp = pad8(ckv);
    const size_t mark = c.W.off;
    const bf16* dkv16 = reinterpret_cast<const bf16*>(dkv);  // bf16 mode: the kv gradient is bf16
    PIO_TAKE_BYTES(float, dh0, c, (size_t)rows * ckv * 4, "pio_kv_project_bwd");
    Side side(c, acc);
    PIO_TRY(side.fork());
    const bool packed_g = g->bv == g->bk + e && g->wv == g->wk + (size_t)e * ckv;
    if (packed_g && side.active() && ckv == 256 && packed16(w->wk, w->wv, e, ckp) && cdiv(2 * e, 128) <= 160) {
        // dW (2E x C_kv) and its bias gradient in one batched launch (pio_dw_batched.cu), the input gradient as 128 x 256 tiles
        DwBatchItem it{dkv16, 2 * e, 2 * e, s.h0, ckp, ckv, g->wk, g->bk};
        PIO_TRY(dw_batched(&it, 1, rows, side.stream()));
        if (!xw_splitk_stores(rows, 2 * e)) PIO_CUDA(cudaMemsetAsync(dh0, 0, (size_t)rows * ckv * sizeof(float), c.st));
        PIO_TRY(xw_splitk(dkv16, 2 * e, w->wk, ckp, dh0, rows, 2 * e, c.st));
        PIO_TRY(pio_layernorm_bwd(dh0, x_kv, s.stats, w->kv_ln_w, dx_prev, dx_kv, g->kv_ln_w, g->kv_ln_b, acc, rows, ckv, c.rest(),
                                  c.rest_bytes(), c.st));
        PIO_TRY(side.join());
        c.W.off = mark;
        return PIO_OK;
    }
    if (g->bv == g->bk + e) {
        PIO_TRY(side_colsum(c, side, dkv16, 2 * e, g->bk, acc, rows, 2 * e));
    } else {
        PIO_TRY(side_colsum(c, side, dkv16, 2 * e, g->bk, acc, rows, e));
        PIO_TRY(side_colsum(c, side, dkv16 + e, 2 * e, g->bv, acc, rows, e));
    }
    if (g->wv == g->wk + (size_t)e * ckv) {
        PIO_TRY(tc_dW(c, dkv16, 2 * e, s.h0, ckp, g->wk, rows, 2 * e, ckv, acc, side.stream()));
    } else {
        PIO_TRY(tc_dW(c, dkv16, 2 * e, s.h0, ckp, g->wk, rows, e, ckv, acc, side.stream()));
        PIO_TRY(tc_dW(c, dkv16 + e, 2 * e, s.h0, ckp, g->wv, rows, e, ckv, acc, side.stream()));
    }
    if (packed16(w->wk, w->wv, e, ckp)) {
        PIO_TRY(tc_dyW(c, dkv16, 2 * e, (const bf16*)w->wk, ckp, rows, 2 * e, ckv, dh0, ckv, nullptr, 0, 0, nullptr, 0, 0));
    } else {
        PIO_TRY(tc_dyW(c, dkv16, 2 * e, (const bf16*)w->wk, ckp, rows, e, ckv, dh0, ckv, nullptr, 0, 0, nullptr, 0, 0));
        PIO_TRY(tc_dyW(c, dkv16 + e, 2 * e, (const bf16*)w->wv, ckp, rows, e, ckv, dh0, ckv, nullptr, 0, 0, nullptr, 0, 1));
    }
    PIO_TRY(pio_layernorm_bwd(dh0, x_kv, s.stats, w->kv_ln_w, dx_prev, dx_kv, g->kv_ln_w, g->kv_ln_b, acc, rows, ckv, c.rest(),
                              c.rest_bytes(), c.st));
    PIO_TRY(side.join());
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- attention helpers
static void fill_attn(pio_attn_desc& a, const bf16* q, int64_t q_rs, const bf16* k, const bf16* v, int64_t kv_rs,
                      const uint8_t* pad, float* o, float* lse, int batch, int heads, int n, int m, int e) {
    // the descriptor's q/k/v are typed float*; AttnBf16Out::in16 tells the kernels they are bf16
    a.q = reinterpret_cast<const float*>(q); a.k = reinterpret_cast<const float*>(k); a.v = reinterpret_cast<const float*>(v);
    a.pad = pad; a.o = o; a.lse = lse;
    a.q_rs = q_rs; a.q_bs = (int64_t)n * q_rs;
    a.k_rs = kv_rs; a.k_bs = (int64_t)m * kv_rs;
    a.v_rs = kv_rs; a.v_bs = (int64_t)m * kv_rs;
    a.o_rs = e; a.o_bs = (int64_t)n * e;
    a.batch = batch; a.heads = heads; a.n = n; a.m = m; a.dh = e / heads;
    a.scale = 1.0f / sqrtf((float)(e / heads));
}

// bf16 q/k/v -> o (fp32 + bf16) and lse: tensor-core flash kernel when the head size allows, else
// the fp32-arithmetic kernel reading the same bf16 operands
static int attention_fwd(const pio_attn_desc& a, bf16* o16, cudaStream_t st, const DropCfg& drop) {
    if (attn_fwd_tc_supported(&a)) return attn_fwd_tc(&a, o16, st, &drop);
    AttnBf16Out ex;
    ex.o16 = o16;
    ex.in16 = 1;
    ex.drop = drop;
    return attn_fwd_simt(&a, st, &ex);
}

// ---------------------------------------------------------------- query side of cross-attention
struct XaSaved {
    float *stats, *o, *lse;
    bf16 *hq, *q, *o16;
    size_t total;
    XaSaved(void* p, int batch, int n, int e, int heads) {
        const size_t r = (size_t)batch * n;
        ByteCarver k{(char*)p, 0};
        stats = k.f32(r * 2);
        hq = k.b16(r * e);
        q = k.b16(r * e);
        o = k.f32(r * e);
        o16 = k.b16(r * e);
        lse = k.f32((size_t)batch * heads * n);
        total = k.off;
    }
};
size_t xa_saved_bytes(int batch, int n, int e, int heads) { return XaSaved(nullptr, batch, n, e, heads).total + 256; }
size_t xa_o16_offset(int batch, int n, int e, int heads) {
    char* const base = reinterpret_cast<char*>((uintptr_t)1 << 20);  // (any aligned non-null address: only the difference is used)
    return (size_t)((char*)XaSaved(base, batch, n, e, heads).o16 - base);
}
bool xa_merges_dkv(int batch, int n, int m, int e, int heads) {  // the shapes whose backward runs on attn_bwd_tc
    if (heads <= 0 || e % heads != 0 || e % 8 != 0) return false;
    pio_attn_desc d{};
    char* const base = reinterpret_cast<char*>((uintptr_t)1 << 20);
    fill_attn(d, (const bf16*)base, e, (const bf16*)base, (const bf16*)base + e, 2 * e, nullptr, (float*)base, (float*)base, batch, heads, n, m, e);
    return attn_fwd_tc_supported(&d);
}
size_t xa_ws_bytes(int batch, int n, int e, int heads) {
    const size_t r = (size_t)batch * n;
    return 2 * align_up(r * e * 2, 256) + 4 * align_up(r * e * 4, 256) + align_up((size_t)batch * heads * n * 4, 256) +
           scratch_bytes(r, e, e);
}

int cross_attend_fwd(Ctx& c, const float* x_q, const float* kv, const uint8_t* pad, float* y, const pio_attn_weights* w,
                     void* saved, int batch, int n, int m, int e, int heads, int residual) {
    PIO_TRY(need_mult8(e, "embed dim", "pio_cross_attend_fwd"));
    XaSaved s(saved, batch, n, e, heads);
    const int64_t rows = (int64_t)batch * n;
    PIO_TRY(layernorm_fwd_bf16(x_q, w->ln_w, w->ln_b, s.hq, e, s.stats, rows, e, LN_EPS, c.st));
    PIO_TRY(tc_xwT(c, s.hq, e, (const bf16*)w->wq, e, rows, e, e, nullptr, 0, s.q, e, w->bq, nullptr, 0, 0, nullptr, 0));
    const bf16* kv16 = reinterpret_cast<const bf16*>(kv);
    pio_attn_desc a{};
    fill_attn(a, s.q, e, kv16, kv16 + e, 2 * e, pad, s.o, s.lse, batch, heads, n, m, e);
    PIO_TRY(attention_fwd(a, s.o16, c.st, c.attn_drop()));
    if (residual == 2) return PIO_OK;  // deferred out-projection: the caller takes the attention output (s.o16) from the saved buffer
    PIO_TRY(tc_xwT(c, s.o16, e, (const bf16*)w->wo, e, rows, e, e, y, e, nullptr, 0, w->bo, residual ? x_q : nullptr, e, 0,
                   nullptr, 0, residual ? c.drop_p : 0.f, c.drop_seed + 1));
    return PIO_OK;
}

int cross_attend_bwd(Ctx& c, const float* dy, const float* x_q, const float* kv, const uint8_t* pad, float* dx_q, float* dkv,
                     const pio_attn_weights* w, const void* saved, const pio_attn_grads* g, int acc, int batch, int n, int m,
                     int e, int heads, int residual) {
    PIO_TRY(need_mult8(e, "embed dim", "pio_cross_attend_bwd"));
    const bool dx_add = (acc & PIO_ACC_DX) != 0, dkv_add = (acc & PIO_ACC_DKV) != 0;  // see pio_b200.h
    acc &= 3;
    XaSaved s((void*)saved, batch, n, e, heads);
    const int64_t rows = (int64_t)batch * n;
    const size_t mark = c.W.off;
    PIO_TAKE_BYTES(bf16, dy16, c, (size_t)rows * e * 2, "pio_cross_attend_bwd");
    PIO_TAKE_BYTES(bf16, dq16, c, (size_t)rows * e * 2, "pio_cross_attend_bwd");
    PIO_TAKE_BYTES(float, d_o, c, (size_t)rows * e * 4, "pio_cross_attend_bwd");
    PIO_TAKE_BYTES(float, dhq, c, (size_t)rows * e * 4, "pio_cross_attend_bwd");
    PIO_TAKE_BYTES(float, delta, c, (size_t)batch * heads * n * 4, "pio_cross_attend_bwd");
    Side side(c, acc);
    const bf16* kv16 = reinterpret_cast<const bf16*>(kv);
    bf16* dkv16 = reinterpret_cast<bf16*>(dkv);
    pio_attn_bwd_desc b{};
    fill_attn(b.fwd, s.q, e, kv16, kv16 + e, 2 * e, pad, s.o, s.lse, batch, heads, n, m, e);
    const DropCfg adrop = c.attn_drop();
    PIO_CHECK_ARG(!dkv_add || attn_fwd_tc_supported(&b.fwd), "pio_cross_attend_bwd: PIO_ACC_DKV needs the tensor-core attention kernels");
    if (residual == 2) {
        // deferred out-projection: `dy` IS the gradient of the attention output, bf16 (rows, e); the out-projection, its
        // parameter gradients and the residual path belong to the caller (pio_latent_block_bwd_pre)
        PIO_CHECK_ARG(attn_fwd_tc_supported(&b.fwd), "pio_cross_attend_bwd: the deferred out-projection needs the tensor-core attention kernels");
        PIO_TRY(attn_bwd_tc(&b.fwd, reinterpret_cast<const bf16*>(dy), delta, dq16, dkv16, dkv16 + e, c.st, &adrop, s.o16, dkv_add));
        PIO_TRY(side.fork());
        PIO_TRY(side_colsum(c, side, dq16, e, g->bq, acc, rows, e));
        PIO_TRY(tc_dW(c, dq16, e, s.hq, e, g->wq, rows, e, e, acc, side.stream()));
        PIO_TRY(tc_dyW(c, dq16, e, (const bf16*)w->wq, e, rows, e, e, dhq, e, nullptr, 0, 0, nullptr, 0, 0));
        PIO_TRY(pio_layernorm_bwd(dhq, x_q, s.stats, w->ln_w, dx_add ? dx_q : nullptr, dx_q, g->ln_w, g->ln_b, acc, rows, e, c.rest(),
                                  c.rest_bytes(), c.st));
        PIO_TRY(side.join());
        c.W.off = mark;
        return PIO_OK;
    }
    const float* dyb;
    PIO_TRY(branch_grad(c, dy, rows * e, residual, &dyb, "pio_cross_attend_bwd"));
    PIO_TRY(cast_colsum_f32(dyb, e, dy16, e, g->bo, acc, rows, e, c.rest(), c.rest_bytes(), c.st));
    PIO_TRY(side.fork());
    PIO_TRY(tc_dW(c, dy16, e, s.o16, e, g->wo, rows, e, e, acc, side.stream()));
    if (attn_fwd_tc_supported(&b.fwd)) {
        bf16* d_o16 = reinterpret_cast<bf16*>(d_o);  // the fp32 scratch is large enough for the bf16 tile
        PIO_TRY(tc_dyW(c, dy16, e, (const bf16*)w->wo, e, rows, e, e, nullptr, 0, d_o16, e, 0, nullptr, 0, 0));
        PIO_TRY(attn_bwd_tc(&b.fwd, d_o16, delta, dq16, dkv16, dkv16 + e, c.st, &adrop, nullptr, dkv_add));
    } else {
        PIO_TRY(tc_dyW(c, dy16, e, (const bf16*)w->wo, e, rows, e, e, d_o, e, nullptr, 0, 0, nullptr, 0, 0));
        b.d_o = d_o; b.delta = delta;
        AttnBf16Out ex;
        ex.in16 = 1;
        ex.drop = adrop;
        ex.dq16 = dq16; ex.dk16 = dkv16; ex.dv16 = dkv16 + e;
        PIO_TRY(attn_bwd_simt(&b, c.st, &ex));
    }
    PIO_TRY(side.fork());
    PIO_TRY(side_colsum(c, side, dq16, e, g->bq, acc, rows, e));
    PIO_TRY(tc_dW(c, dq16, e, s.hq, e, g->wq, rows, e, e, acc, side.stream()));
    PIO_TRY(tc_dyW(c, dq16, e, (const bf16*)w->wq, e, rows, e, e, dhq, e, nullptr, 0, 0, nullptr, 0, 0));
    PIO_TRY(pio_layernorm_bwd(dhq, x_q, s.stats, w->ln_w, residual ? dy : nullptr, dx_q, g->ln_w, g->ln_b, acc, rows, e,
                              c.rest(), c.rest_bytes(), c.st));
    PIO_TRY(side.join());
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- self-attention layer
struct SaSaved {
    float *stats, *o, *lse;
    bf16 *h0, *qkv, *o16;
    size_t total;
    SaSaved(void* p, int batch, int n, int ch, int heads) {
        const size_t r = (size_t)batch * n;
        ByteCarver k{(char*)p, 0};
        stats = k.f32(r * 2);
        h0 = k.b16(r * ch);
        qkv = k.b16(3 * r * ch);
        o = k.f32(r * ch);
        o16 = k.b16(r * ch);
        lse = k.f32((size_t)batch * heads * n);
        total = k.off;
    }
};
size_t sa_saved_bytes(int batch, int n, int ch, int heads) { return SaSaved(nullptr, batch, n, ch, heads).total + 256; }
size_t sa_ws_bytes(int batch, int n, int ch, int heads) {
    const size_t r = (size_t)batch * n;
    return align_up(r * ch * 2, 256) + align_up(3 * r * ch * 2, 256) + 3 * align_up(r * ch * 4, 256) +
           align_up(3 * r * ch * 4, 256) + align_up((size_t)batch * heads * n * 4, 256) + scratch_bytes(r, 3 * ch, ch);
}

static int check_packed_self(const pio_attn_weights* w, int ch, const char* who) {
    const char* q = (const char*)w->wq;
    PIO_CHECK_ARG((const char*)w->wk == q + (size_t)ch * ch * 2 && (const char*)w->wv == q + 2 * (size_t)ch * ch * 2 &&
                      w->bk == w->bq + ch && w->bv == w->bq + 2 * ch,
                  "%s: bf16 mode expects the packed (3C,C) in-projection of a self-attention layer", who);
    return PIO_OK;
}

int self_attn_fwd(Ctx& c, const float* x, float* y, const pio_attn_weights* w, void* saved, int batch, int n, int ch, int heads,
                  int residual) {
    PIO_TRY(need_mult8(ch, "channels", "pio_self_attn_fwd"));
    PIO_TRY(check_packed_self(w, ch, "pio_self_attn_fwd"));
    SaSaved s(saved, batch, n, ch, heads);
    const int64_t rows = (int64_t)batch * n;
    PIO_TRY(layernorm_fwd_bf16(x, w->ln_w, w->ln_b, s.h0, ch, s.stats, rows, ch, LN_EPS, c.st));
    PIO_TRY(tc_xwT(c, s.h0, ch, (const bf16*)w->wq, ch, rows, 3 * ch, ch, nullptr, 0, s.qkv, 3 * ch, w->bq, nullptr, 0, 0,
                   nullptr, 0));
    pio_attn_desc a{};
    fill_attn(a, s.qkv, 3 * ch, s.qkv + ch, s.qkv + 2 * ch, 3 * ch, nullptr, s.o, s.lse, batch, heads, n, n, ch);
    PIO_TRY(attention_fwd(a, s.o16, c.st, c.attn_drop()));
    PIO_TRY(tc_xwT(c, s.o16, ch, (const bf16*)w->wo, ch, rows, ch, ch, y, ch, nullptr, 0, w->bo, residual ? x : nullptr, ch, 0,
                   nullptr, 0, residual ? c.drop_p : 0.f, c.drop_seed + 1));
    return PIO_OK;
}

int self_attn_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_attn_weights* w, const void* saved,
                  const pio_attn_grads* g, int acc, int batch, int n, int ch, int heads, int residual, const bf16* dy16_in,
                  bf16* dx16_out) {
    PIO_TRY(need_mult8(ch, "channels", "pio_self_attn_bwd"));
    PIO_TRY(check_packed_self(w, ch, "pio_self_attn_bwd"));
    PIO_CHECK_ARG(g->wk == g->wq + (size_t)ch * ch && g->wv == g->wq + 2 * (size_t)ch * ch && g->bk == g->bq + ch &&
                      g->bv == g->bq + 2 * ch,
                  "pio_self_attn_bwd: packed (3C,C) gradient storage expected");
    SaSaved s((void*)saved, batch, n, ch, heads);
    const int64_t rows = (int64_t)batch * n;
    const size_t mark = c.W.off;
    PIO_TAKE_BYTES(bf16, dy16_own, c, (size_t)rows * ch * 2, "pio_self_attn_bwd");
    PIO_TAKE_BYTES(bf16, dqkv16, c, (size_t)rows * 3 * ch * 2, "pio_self_attn_bwd");
    PIO_TAKE_BYTES(float, d_o, c, (size_t)rows * ch * 4, "pio_self_attn_bwd");
    PIO_TAKE_BYTES(float, dh0, c, (size_t)rows * ch * 4, "pio_self_attn_bwd");
    PIO_TAKE_BYTES(float, delta, c, (size_t)batch * heads * n * 4, "pio_self_attn_bwd");
    Side side(c, acc);
    const bf16* dy16 = dy16_own;
    if (dy16_in != nullptr && !(residual && c.drop_p > 0.f) && side.active()) {  // see mlp_bwd
        dy16 = dy16_in;
        PIO_TRY(side.fork());
        PIO_TRY(side_colsum(c, side, dy16, ch, g->bo, acc, rows, ch));
    } else {
        const float* dyb;
        PIO_TRY(branch_grad(c, dy, rows * ch, residual, &dyb, "pio_self_attn_bwd"));
        PIO_TRY(cast_colsum_f32(dyb, ch, dy16_own, ch, g->bo, acc, rows, ch, c.rest(), c.rest_bytes(), c.st));
        PIO_TRY(side.fork());
    }
    PIO_TRY(tc_dW(c, dy16, ch, s.o16, ch, g->wo, rows, ch, ch, acc, side.stream()));
    pio_attn_bwd_desc b{};
    fill_attn(b.fwd, s.qkv, 3 * ch, s.qkv + ch, s.qkv + 2 * ch, 3 * ch, nullptr, s.o, s.lse, batch, heads, n, n, ch);
    const DropCfg adrop = c.attn_drop();
    if (attn_fwd_tc_supported(&b.fwd)) {
        bf16* d_o16 = reinterpret_cast<bf16*>(d_o);
        PIO_TRY(tc_dyW(c, dy16, ch, (const bf16*)w->wo, ch, rows, ch, ch, nullptr, 0, d_o16, ch, 0, nullptr, 0, 0));
        PIO_TRY(attn_bwd_tc(&b.fwd, d_o16, delta, dqkv16, dqkv16 + ch, dqkv16 + 2 * ch, c.st, &adrop, s.o16));
    } else {
        PIO_TRY(tc_dyW(c, dy16, ch, (const bf16*)w->wo, ch, rows, ch, ch, d_o, ch, nullptr, 0, 0, nullptr, 0, 0));
        b.d_o = d_o; b.delta = delta;
        AttnBf16Out ex;
        ex.in16 = 1;
        ex.drop = adrop;
        ex.dq16 = dqkv16; ex.dk16 = dqkv16 + ch; ex.dv16 = dqkv16 + 2 * ch;
        PIO_TRY(attn_bwd_simt(&b, c.st, &ex));
    }
    PIO_TRY(side.fork());
    PIO_TRY(side_colsum(c, side, dqkv16, 3 * ch, g->bq, acc, rows, 3 * ch));
    PIO_TRY(tc_dW(c, dqkv16, 3 * ch, s.h0, ch, g->wq, rows, 3 * ch, ch, acc, side.stream()));
    PIO_TRY(tc_dyW(c, dqkv16, 3 * ch, (const bf16*)w->wq, ch, rows, 3 * ch, ch, dh0, ch, nullptr, 0, 0, nullptr, 0, 0));
    PIO_TRY(layernorm_bwd(dh0, x, s.stats, w->ln_w, residual ? dy : nullptr, dx, g->ln_w, g->ln_b, acc, rows, ch, c.rest(),
                          c.rest_bytes(), c.st, nullptr, dx16_out));
    PIO_TRY(side.join());
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- latent block
struct BlockLayout {
    size_t sa, xmid, mlp, xnext, per_layer;  // byte offsets within a layer's slab
    BlockLayout(int batch, int n, int ch, int heads) {
        const size_t r = (size_t)batch * n;
        const size_t act = align_up(r * ch * 4, 256);
        sa = 0;
        xmid = sa + align_up(sa_saved_bytes(batch, n, ch, heads), 256);
        mlp = xmid + act;
        xnext = mlp + align_up(mlp_saved_bytes(r, ch), 256);
        per_layer = xnext + act;
    }
};
size_t block_saved_bytes(int layers, int batch, int n, int ch, int heads) {
    return (size_t)layers * BlockLayout(batch, n, ch, heads).per_layer + 256;
}
// bf16 gradient tensors the fused backward chain leaves per layer for the batched weight-gradient kernel
// (g, dH, g_mid, [dQ | dK | dV]: six [rows][C] tensors)
static size_t block_bwd_layer_bytes(size_t r, int ch) { return 6 * align_up(r * ch * 2, 256); }
size_t block_ws_bytes(int layers, int batch, int n, int ch, int heads) {
    const size_t r = (size_t)batch * n;
    return 2 * align_up(r * ch * 4, 256) + 2 * align_up(r * ch * 2, 256) + std::max(sa_ws_bytes(batch, n, ch, heads), mlp_ws_bytes(r, ch)) +
           (size_t)layers * block_bwd_layer_bytes(r, ch) + 1024;
}

static bool mlp_only_layer(const pio_attn_weights& a) { return a.wq == nullptr; }

int latent_block_fwd(Ctx& c, const float* x, float* y, const pio_attn_weights* attn, const pio_mlp_weights* mlp, void* saved,
                     int layers, int batch, int n, int ch, int heads, const void* pre_o16) {
    BlockLayout L(batch, n, ch, heads);
    const int64_t rows = (int64_t)batch * n;
    const float* cur = x;
    // A first layer without an attention half of its own (attn[0].wq == null): the Residual(mlp) of the cross-attention
    // layer in front of the block and, when attn[0].wo is given, that layer's out-projection + residual
    // (x += pre_o16 Wo^T + bo with pre_o16 the attention output) before it — handed over so that the fused kernels run them
    // as half a layer.  Taken inside the cluster kernels when they run the whole block in one launch; otherwise with the
    // operators of their own, followed by the block.
    const bool pre = mlp_only_layer(attn[0]);
    const bool pre_proj = pre && attn[0].wo != nullptr;
    PIO_CHECK_ARG(!pre_proj || (pre_o16 != nullptr && attn[0].bo != nullptr), "pio_latent_block_fwd: an out-projection first layer needs the attention output");
    bool packed = true;
    for (int s = pre ? 1 : 0; s < layers; ++s) packed = packed && check_packed_self(&attn[s], ch, "pio_latent_block_fwd") == PIO_OK;
    const bool fused = packed && latent_fused_supported(rows, n, ch, heads, c.drop_p);
    int first = 0;
    if (pre && !(fused && latent_cluster_preferred(rows) && layers <= PIO_LATENT_FUSED_MAX_LAYERS)) {
        char* slab = (char*)saved;
        float* out = (layers == 1) ? y : (float*)(slab + L.xnext);
        float* xmid = (float*)(slab + L.xmid);
        const uint64_t seed0 = c.drop_seed;
        if (pre_proj) {  // x_mid = x + O Wo^T + bo
            PIO_TRY(tc_xwT(c, (const bf16*)pre_o16, ch, (const bf16*)attn[0].wo, ch, rows, ch, ch, xmid, ch, nullptr, 0, attn[0].bo, x, ch, 0,
                           nullptr, 0));
        } else {
            // (the fused backward chain reads a layer's x_mid from its slab; for this layer that is the block's input)
            PIO_CUDA(cudaMemcpyAsync(xmid, x, (size_t)rows * ch * sizeof(float), cudaMemcpyDeviceToDevice, c.st));
        }
        c.drop_seed = seed0 + 2;
        PIO_TRY(mlp_fwd(c, xmid, out, &mlp[0], slab + L.mlp, rows, ch, 1));
        c.drop_seed = seed0;
        cur = out;
        first = 1;
        if (layers == 1) return PIO_OK;
    }
    if (fused) {
        // one persistent kernel per (up to 8) layers: pio_latent_cl.cu / pio_latent_tc.cu
        {
            for (int s0 = first; s0 < layers; s0 += PIO_LATENT_FUSED_MAX_LAYERS) {
                const int cnt = std::min(PIO_LATENT_FUSED_MAX_LAYERS, layers - s0);
                LatentFusedLayer fl[PIO_LATENT_FUSED_MAX_LAYERS];
                for (int i = 0; i < cnt; ++i) {
                    const int s = s0 + i;
                    char* slab = (char*)saved + (size_t)s * L.per_layer;
                    SaSaved sa(slab + L.sa, batch, n, ch, heads);
                    MlpSaved ms(slab + L.mlp, rows, ch);
                    LatentFusedLayer& f = fl[i];
                    f.wqkv = attn[s].wq; f.wo = attn[s].wo; f.w1 = mlp[s].w1; f.w2 = mlp[s].w2;
                    f.ln1_w = attn[s].ln_w; f.ln1_b = attn[s].ln_b; f.bqkv = attn[s].bq; f.bo = attn[s].bo;
                    f.ln2_w = mlp[s].ln_w; f.ln2_b = mlp[s].ln_b; f.b1 = mlp[s].b1; f.b2 = mlp[s].b2;
                    f.sa_stats = sa.stats; f.sa_h0 = sa.h0; f.qkv = sa.qkv; f.o32 = sa.o; f.o16 = sa.o16; f.lse = sa.lse;
                    f.xmid = (float*)(slab + L.xmid);
                    f.mlp_stats = ms.stats; f.mlp_h0 = ms.h0; f.mlp_h1 = ms.h1; f.mlp_g = ms.g;
                    f.xout = (s == layers - 1) ? y : (float*)(slab + L.xnext);
                    f.pre_o16 = (s == 0) ? pre_o16 : nullptr;
                }
                PIO_TRY(latent_block_fwd_fused(cur, fl, cnt, batch, n, c.st));
                cur = fl[cnt - 1].xout;
            }
            return PIO_OK;
        }
    }
    for (int s = first; s < layers; ++s) {
        char* slab = (char*)saved + (size_t)s * L.per_layer;
        float* xmid = (float*)(slab + L.xmid);
        float* out = (s == layers - 1) ? y : (float*)(slab + L.xnext);
        const uint64_t seed0 = c.drop_seed;
        c.drop_seed = seed0 + 4 * (uint64_t)s;
        PIO_TRY(self_attn_fwd(c, cur, xmid, &attn[s], slab + L.sa, batch, n, ch, heads, 1));
        c.drop_seed = seed0 + 4 * (uint64_t)s + 2;
        PIO_TRY(mlp_fwd(c, xmid, out, &mlp[s], slab + L.mlp, rows, ch, 1));
        c.drop_seed = seed0;
        cur = out;
    }
    return PIO_OK;
}

// Backward of a block through the fused input-gradient chain (pio_latent_bwd_cl.cu: one launch per up to 8 layers),
// followed by the parameter gradients of every layer on the side streams: four split-K weight-gradient GEMMs, four
// bias column sums and the two LayerNorm parameter-gradient kernels per layer, all reading the bf16 tensors the chain
// kernel left in the workspace.
static int latent_block_bwd_fused(Ctx& c, Side& side, const float* dy, const float* x, float* dx, const pio_attn_weights* attn,
                                  const pio_mlp_weights* mlp, const void* saved, const pio_attn_grads* attn_g,
                                  const pio_mlp_grads* mlp_g, int layers, int batch, int n, int ch, int heads, const void* pre_o16,
                                  void* pre_do16) {
    BlockLayout L(batch, n, ch, heads);
    const int64_t rows = (int64_t)batch * n;
    const size_t mark = c.W.off;
    const size_t one = align_up((size_t)rows * ch * 2, 256);
    PIO_TAKE_BYTES(char, gbuf, c, (size_t)layers * block_bwd_layer_bytes(rows, ch), "pio_latent_block_bwd");
    PIO_TAKE_BYTES(float, dcur, c, (size_t)rows * ch * 4, "pio_latent_block_bwd");
    PIO_TAKE_BYTES(float, dcur2, c, (size_t)rows * ch * 4, "pio_latent_block_bwd");
    std::vector<LatentBwdLayer> bl((size_t)layers);
    for (int s = 0; s < layers; ++s) {
        const char* slab = (const char*)saved + (size_t)s * L.per_layer;
        SaSaved sa((void*)(slab + L.sa), batch, n, ch, heads);
        MlpSaved ms((void*)(slab + L.mlp), rows, ch);
        char* g = gbuf + (size_t)s * block_bwd_layer_bytes(rows, ch);
        LatentBwdLayer& b = bl[(size_t)s];
        b.w2 = mlp[s].w2; b.w1 = mlp[s].w1; b.wo = attn[s].wo; b.wqkv = attn[s].wq;
        b.ln2_w = mlp[s].ln_w; b.ln1_w = attn[s].ln_w;
        b.mlp_stats = ms.stats; b.sa_stats = sa.stats; b.lse = sa.lse;
        b.mlp_h1 = ms.h1; b.xmid = (const float*)(slab + L.xmid); b.xnext = (const float*)(slab + L.xnext); b.qkv = sa.qkv;
        b.gy16 = g; b.gh16 = g + one; b.gmid16 = g + 2 * one; b.gqkv16 = g + 3 * one;
        b.pre_do16 = (s == 0) ? pre_do16 : nullptr;
        b.dln2_w = mlp_g[s].ln_w; b.dln2_b = mlp_g[s].ln_b; b.dln1_w = attn_g[s].ln_w; b.dln1_b = attn_g[s].ln_b;
    }
    // the chain, from the last layers to the first, up to 8 layers per launch
    const float* grad = dy;
    float* bufs[2] = {dcur, dcur2};
    int flip = 0;
    for (int hi = layers; hi > 0;) {
        const int lo = std::max(0, hi - PIO_LATENT_FUSED_MAX_LAYERS);
        const float* x_lo = (lo == 0) ? x : bl[(size_t)lo - 1].xnext;
        float* out = (lo == 0) ? dx : bufs[flip];
        PIO_TRY(latent_block_bwd_cluster(grad, x_lo, out, &bl[(size_t)lo], hi - lo, batch, n, c.st));
        grad = out;
        flip ^= 1;
        hi = lo;
    }
    if (getenv("PIO_LATENT_BWD_CHAIN_ONLY") != nullptr) {  // measurement hook (bench.py times the chain kernel alone)
        c.W.off = mark;
        return PIO_OK;
    }
    // parameter gradients: per (up to 8) layers one batched weight-gradient launch (bias gradients ride on it) and one
    // launch for the LayerNorm parameter gradients, on the two side streams
    PIO_TRY(side.fork());
    for (int hi = layers; hi > 0;) {
        const int lo = std::max(0, hi - PIO_LATENT_FUSED_MAX_LAYERS);
        DwBatchItem dwi[4 * PIO_LATENT_FUSED_MAX_LAYERS];
        int nd = 0;
        for (int s = hi - 1; s >= lo; --s) {
            const char* slab = (const char*)saved + (size_t)s * L.per_layer;
            SaSaved sa((void*)(slab + L.sa), batch, n, ch, heads);
            MlpSaved ms((void*)(slab + L.mlp), rows, ch);
            const LatentBwdLayer& b = bl[(size_t)s];
            dwi[nd++] = DwBatchItem{b.gy16, ch, ch, ms.g, ch, ch, mlp_g[s].w2, mlp_g[s].b2};
            dwi[nd++] = DwBatchItem{b.gh16, ch, ch, ms.h0, ch, ch, mlp_g[s].w1, mlp_g[s].b1};
            if (mlp_only_layer(attn[s])) {
                if (attn[s].wo != nullptr) dwi[nd++] = DwBatchItem{b.gmid16, ch, ch, pre_o16, ch, ch, attn_g[s].wo, attn_g[s].bo};
                continue;
            }
            dwi[nd++] = DwBatchItem{b.gmid16, ch, ch, sa.o16, ch, ch, attn_g[s].wo, attn_g[s].bo};
            dwi[nd++] = DwBatchItem{b.gqkv16, 3 * ch, 3 * ch, sa.h0, ch, ch, attn_g[s].wq, attn_g[s].bq};
        }
        PIO_TRY(dw_batched(dwi, nd, rows, side.stream()));  // (the LayerNorm parameter gradients were added by the chain kernel)
        hi = lo;
    }
    if (!t_defer_side_join) PIO_TRY(side.join());  // (deferred: the caller joins with pio_side_join)
    c.W.off = mark;
    return PIO_OK;
}

int latent_block_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_attn_weights* attn, const pio_mlp_weights* mlp,
                     const void* saved, const pio_attn_grads* attn_g, const pio_mlp_grads* mlp_g, int acc, int layers,
                     int batch, int n, int ch, int heads, const void* pre_o16, void* pre_do16) {
    BlockLayout L(batch, n, ch, heads);
    const int64_t rows = (int64_t)batch * n;
    const bool pre = mlp_only_layer(attn[0]);
    if (acc == 2 && latent_bwd_cluster_supported(rows, n, ch, heads, c.drop_p) && (!pre || layers <= PIO_LATENT_FUSED_MAX_LAYERS)) {
        bool packed = true;
        for (int s = pre ? 1 : 0; s < layers; ++s) packed = packed && check_packed_self(&attn[s], ch, "pio_latent_block_bwd") == PIO_OK;
        Side side(c, acc);
        if (packed && side.active())
            return latent_block_bwd_fused(c, side, dy, x, dx, attn, mlp, saved, attn_g, mlp_g, layers, batch, n, ch, heads, pre_o16, pre_do16);
    }
    PIO_TAKE_BYTES(float, dmid, c, (size_t)rows * ch * 4, "pio_latent_block_bwd");
    PIO_TAKE_BYTES(float, dcur, c, (size_t)rows * ch * 4, "pio_latent_block_bwd");
    // each operator's LayerNorm backward also writes its dx as bf16, which the next operator takes as its dY operand
    PIO_TAKE_BYTES(bf16, dmid16, c, (size_t)rows * ch * 2, "pio_latent_block_bwd");
    PIO_TAKE_BYTES(bf16, dcur16, c, (size_t)rows * ch * 2, "pio_latent_block_bwd");
    const float* grad = dy;
    const bf16* grad16 = nullptr;
    for (int s = layers - 1; s >= 0; --s) {
        const char* slab = (const char*)saved + (size_t)s * L.per_layer;
        const float* x_in = (s == 0) ? x : (const float*)((const char*)saved + (size_t)(s - 1) * L.per_layer + L.xnext);
        const uint64_t seed0 = c.drop_seed;
        if (s == 0 && pre) {  // the first layer without attention of its own
            c.drop_seed = seed0 + 2;
            if (attn[0].wo == nullptr) {  // MLP only: its input is the block's input
                PIO_TRY(mlp_bwd(c, grad, x, dx, &mlp[0], slab + L.mlp, &mlp_g[0], acc, rows, ch, 1, grad16, nullptr));
            } else {
                // out-projection + MLP: the MLP's input is x_mid = x + O Wo^T + bo (in the slab); dx = g_mid, dO = g_mid Wo,
                // dWo += g_mid^T O, dbo += column sums of g_mid
                PIO_CHECK_ARG(pre_o16 != nullptr && pre_do16 != nullptr, "pio_latent_block_bwd: an out-projection first layer needs O and dO");
                PIO_TRY(mlp_bwd(c, grad, (const float*)(slab + L.xmid), dx, &mlp[0], slab + L.mlp, &mlp_g[0], acc, rows, ch, 1, grad16, dmid16));
                Side side(c, acc);
                PIO_TRY(side.fork());
                PIO_TRY(side_colsum(c, side, dmid16, ch, attn_g[0].bo, acc, rows, ch));
                PIO_TRY(tc_dW(c, dmid16, ch, (const bf16*)pre_o16, ch, attn_g[0].wo, rows, ch, ch, acc, side.stream()));
                PIO_TRY(tc_dyW(c, dmid16, ch, (const bf16*)attn[0].wo, ch, rows, ch, ch, nullptr, 0, (bf16*)pre_do16, ch, 0, nullptr, 0, 0));
                PIO_TRY(side.join());
            }
            c.drop_seed = seed0;
            break;
        }
        c.drop_seed = seed0 + 4 * (uint64_t)s + 2;
        PIO_TRY(mlp_bwd(c, grad, (const float*)(slab + L.xmid), dmid, &mlp[s], slab + L.mlp, &mlp_g[s], acc, rows, ch, 1, grad16, dmid16));
        float* out = (s == 0) ? dx : dcur;
        c.drop_seed = seed0 + 4 * (uint64_t)s;
        PIO_TRY(self_attn_bwd(c, dmid, x_in, out, &attn[s], slab + L.sa, &attn_g[s], acc, batch, n, ch, heads, 1, dmid16,
                              s == 0 ? nullptr : dcur16));
        c.drop_seed = seed0;
        grad = out;
        grad16 = (s == 0) ? nullptr : dcur16;
    }
    return PIO_OK;
}

// ---------------------------------------------------------------- output linear
size_t linear_ws_bytes(int64_t rows, int c_in, int64_t c_out) {
    return align_up((size_t)rows * pad8(c_out) * 2, 256) + align_up((size_t)rows * pad8(c_in) * 2, 256) +
           scratch_bytes(rows, c_out, c_in);
}

int linear_fwd(Ctx& c, const float* x, const float* w16, const float* b, float* y, int64_t rows, int c_in, int64_t c_out) {
    const int64_t cip = pad8(c_in);
    const size_t mark = c.W.off;
    PIO_TAKE_BYTES(bf16, x16, c, (size_t)rows * cip * 2, "pio_linear_fwd");
    PIO_TRY(cast_bf16(x, x16, rows, c_in, c_in, cip, c.st));
    PIO_TRY(tc_xwT(c, x16, cip, (const bf16*)w16, cip, rows, c_out, c_in, y, c_out, nullptr, 0, b, nullptr, 0, 0, nullptr, 0));
    c.W.off = mark;
    return PIO_OK;
}

int linear_bwd(Ctx& c, const float* dy, const float* x, const float* w16, float* dx, float* dw, float* db, int acc, int64_t rows,
               int c_in, int64_t c_out) {
    const int64_t cip = pad8(c_in), cop = pad8(c_out);
    const size_t mark = c.W.off;
    PIO_TAKE_BYTES(bf16, dy16, c, (size_t)rows * cop * 2, "pio_linear_bwd");
    PIO_TAKE_BYTES(bf16, x16, c, (size_t)rows * cip * 2, "pio_linear_bwd");
    PIO_TRY(cast_colsum_f32(dy, c_out, dy16, cop, db, acc, rows, (int)c_out, c.rest(), c.rest_bytes(), c.st));
    PIO_TRY(cast_bf16(x, x16, rows, c_in, c_in, cip, c.st));
    if (dx) PIO_TRY(tc_dyW(c, dy16, cop, (const bf16*)w16, cip, rows, c_out, c_in, dx, c_in, nullptr, 0, 0, nullptr, 0, 0));
    PIO_TRY(tc_dW(c, dy16, cop, x16, cip, dw, rows, c_out, c_in, acc));
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- output linear fused with cross-entropy
// loss = mean over valid rows of CE(x W^T + b, labels) and its gradients without the (rows, classes) logits
// tensor ever existing.
//  * Rows whose label is ignored contribute exactly zero to the loss and to every gradient, so they are
//    never projected: the valid rows are compacted on the device (an MLM batch labels ~15 % of its
//    positions), and dx of the ignored rows is written as zeros.  The number of valid rows stays on the
//    device — every kernel below is launched for the worst case (all rows) and skips the tiles beyond
//    the count it reads from memory (pio_gemm_bf16_desc::dyn_rows), so the sequence can be captured in a
//    CUDA graph and replayed on batches with different label counts.
//  * The compacted rows go through in chunks sized so that a chunk's bf16 logits stay in the 126 MB L2:
//    logits GEMM (bf16 out) -> one kernel that turns them in place into bf16 dlogits and per-row losses
//    -> the two gradient GEMMs and the bias column sums (dW / db accumulating across chunks with atomics
//    on the side stream while the main stream forms dx).
struct LceIndex {
    int* count;        // number of valid rows
    int* idx;          // [rows] compact position -> source row
    int* pos;          // [rows] source row -> compact position, -1 for ignored rows
    int64_t* labels;   // [rows] labels of the compacted rows
};

// Stable compaction of the valid rows, 1024 rows per block: pass 1 counts per block, pass 2 turns the block
// counts into offsets (every block re-scans the short count array) and writes the index maps.
__device__ __forceinline__ bool lce_valid(int64_t y, int64_t classes, int64_t ignore) {
    return !(y == ignore || y < 0 || y >= classes);
}
__global__ void __launch_bounds__(1024) lce_count_kernel(const int64_t* __restrict__ labels, int64_t rows, int64_t classes,
                                                         int64_t ignore, int* __restrict__ block_counts) {
    const int64_t r = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    const bool v = r < rows && lce_valid(labels[r], classes, ignore);
    const int n = __syncthreads_count(v);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = n;
}
__global__ void __launch_bounds__(1024) lce_compact_kernel(const int64_t* __restrict__ labels, int64_t rows, int64_t classes,
                                                           int64_t ignore, const int* __restrict__ block_counts,
                                                           float* __restrict__ out3, LceIndex ix) {
    __shared__ int s_red[32];
    __shared__ int s_warp[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // offset of this block = sum of the counts of the blocks before it (and the grand total for block 0's report)
    int before = 0, total = 0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += 1024) {
        const int cnt = block_counts[i];
        total += cnt;
        if (i < (int)blockIdx.x) before += cnt;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        before += __shfl_xor_sync(0xffffffffu, before, o);
        total += __shfl_xor_sync(0xffffffffu, total, o);
    }
    if (lane == 0) {
        s_red[wid] = before;
        s_warp[wid] = total;
    }
    __syncthreads();
    before = 0;
    total = 0;
    for (int w = 0; w < 32; ++w) {
        before += s_red[w];
        total += s_warp[w];
    }
    __syncthreads();
    const int64_t r = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    const int64_t y = r < rows ? labels[r] : ignore;
    const bool v = r < rows && lce_valid(y, classes, ignore);
    const unsigned ballot = __ballot_sync(0xffffffffu, v);
    if (lane == 0) s_warp[wid] = __popc(ballot);
    __syncthreads();
    int warp_base = 0;
    for (int w = 0; w < wid; ++w) warp_base += s_warp[w];
    if (r < rows) {
        if (v) {
            const int at = before + warp_base + __popc(ballot & ((1u << lane) - 1u));
            ix.idx[at] = (int)r;
            ix.labels[at] = y;
            ix.pos[r] = at;
        } else {
            ix.pos[r] = -1;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *ix.count = total;
        out3[1] = (float)total;
    }
}

// x16c[i] = bf16(x[idx[i]]) for i < count; rows up to the next multiple of 128 are zero-filled (they are read
// as operand padding by the partial tiles / k-blocks); nothing beyond is touched.  Warp per row.
__global__ void __launch_bounds__(256) lce_gather_kernel(const float* __restrict__ x, int c_in, bf16* __restrict__ x16c, int64_t ld,
                                                         LceIndex ix, int64_t rows) {
    const int count = *ix.count;
    const int64_t upto = min(rows, ((int64_t)count + 127) / 128 * 128);
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < upto; i += warps) {
        bf16* dst = x16c + i * ld;
        if (i < count) {
            const float* src = x + (int64_t)ix.idx[i] * c_in;
            for (int j = lane; j < (int)ld; j += 32) dst[j] = __float2bfloat16(j < c_in ? src[j] : 0.f);
        } else {
            for (int j = lane; j < (int)ld; j += 32) dst[j] = __float2bfloat16(0.f);
        }
    }
}

// dx[r] = dxc[pos[r]] for valid rows, 0 for ignored rows.  Warp per row.
__global__ void __launch_bounds__(256) lce_scatter_kernel(const float* __restrict__ dxc, int c_in, float* __restrict__ dx,
                                                          LceIndex ix, int64_t rows) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < rows; r += warps) {
        const int p = ix.pos[r];
        float* dst = dx + r * c_in;
        if (p >= 0) {
            const float* src = dxc + (int64_t)p * c_in;
            for (int j = lane; j < c_in; j += 32) dst[j] = src[j];
        } else {
            for (int j = lane; j < c_in; j += 32) dst[j] = 0.f;
        }
    }
}

__device__ __forceinline__ void online_max_sum(float v, float& m, float& s) {
    if (v > m) {
        s = s * __expf(m - v) + 1.0f;
        m = v;
    } else {
        s += __expf(v - m);
    }
}

// One CTA per compacted row of the chunk [r0, r0 + nr): bf16 logits -> logsumexp, row loss, and in place
// dlogits = (softmax - onehot) / n_valid as bf16 (pad columns zeroed).  Rows between the valid count and
// the next multiple of 64 are zero-filled: the dW product reads them as k-block padding.
__global__ void __launch_bounds__(256) ce_chunk_kernel(bf16* __restrict__ lg, int64_t ld, const int64_t* __restrict__ labels_c,
                                                       const float* __restrict__ out3, const int* __restrict__ count, int64_t r0,
                                                       int64_t nr, float* __restrict__ row_loss, int64_t classes) {
    __shared__ float s_m[8], s_s[8];
    __shared__ float s_lse;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t live = max((int64_t)0, min(nr, (int64_t)*count - r0));
    const int64_t upto = min(nr, (live + 63) / 64 * 64);
    const float inv_n = 1.0f / out3[1];
    const int nvec = (int)(ld >> 3);  // ld is a multiple of 8: rows are 16-byte aligned vectors of 8
    for (int64_t r = blockIdx.x; r < upto; r += gridDim.x) {
        uint4* zv = reinterpret_cast<uint4*>(lg + r * ld);
        if (r >= live) {  // block-uniform
            for (int j = threadIdx.x; j < nvec; j += blockDim.x) zv[j] = make_uint4(0u, 0u, 0u, 0u);
            continue;
        }
        const int64_t y = labels_c[r0 + r];
        float m = -INFINITY, s = 0.f;
        for (int j = threadIdx.x; j < nvec; j += blockDim.x) {
            const uint4 u = zv[j];
            const bf16* e = reinterpret_cast<const bf16*>(&u);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if ((int64_t)j * 8 + k < classes) online_max_sum(__bfloat162float(e[k]), m, s);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
            const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
            const float mm = fmaxf(m, m2);
            const float a = (m == -INFINITY) ? 0.f : s * __expf(m - mm);
            const float b2 = (m2 == -INFINITY) ? 0.f : s2 * __expf(m2 - mm);
            m = mm;
            s = a + b2;
        }
        if (lane == 0) {
            s_m[wid] = m;
            s_s[wid] = s;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float mm = -INFINITY;
            for (int w = 0; w < 8; ++w) mm = fmaxf(mm, s_m[w]);
            float ss = 0.f;
            for (int w = 0; w < 8; ++w)
                if (s_m[w] != -INFINITY) ss += s_s[w] * __expf(s_m[w] - mm);
            const float lse = mm + logf(ss);
            s_lse = lse;
            row_loss[r0 + r] = lse - __bfloat162float(lg[r * ld + y]);
        }
        __syncthreads();
        const float lse = s_lse;
        for (int j = threadIdx.x; j < nvec; j += blockDim.x) {
            uint4 u = zv[j];
            bf16* e = reinterpret_cast<bf16*>(&u);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int64_t col = (int64_t)j * 8 + k;
                float g = 0.f;
                if (col < classes) {
                    g = __expf(__bfloat162float(e[k]) - lse);
                    if (col == y) g -= 1.0f;
                    g *= inv_n;
                }
                e[k] = __float2bfloat16(g);
            }
            zv[j] = u;
        }
        __syncthreads();
    }
}

// The same with the row held in registers between the passes (NV vectors of 8 logits per thread, classes <= 2048 NV):
// one read and one write of the row instead of two reads and a write, all loads of a row in flight together, and the
// block-wide reductions finished by every thread from shared memory instead of by thread 0 alone (the two-pass kernel
// above spent 56 us on a 4352 x 10003 chunk that is resident in L2: it ran at the latency of its dependent passes).
template <int NV>
__global__ void __launch_bounds__(256) ce_chunk_reg_kernel(bf16* __restrict__ lg, int64_t ld, const int64_t* __restrict__ labels_c,
                                                           const float* __restrict__ out3, const int* __restrict__ count, int64_t r0,
                                                           int64_t nr, float* __restrict__ row_loss, int64_t classes) {
    __shared__ float s_m[8], s_s[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t live = max((int64_t)0, min(nr, (int64_t)*count - r0));
    const int64_t upto = min(nr, (live + 63) / 64 * 64);
    const float inv_n = 1.0f / out3[1];
    const int nvec = (int)(ld >> 3);
    for (int64_t r = blockIdx.x; r < upto; r += gridDim.x) {
        uint4* zv = reinterpret_cast<uint4*>(lg + r * ld);
        if (r >= live) {  // block-uniform
            for (int j = threadIdx.x; j < nvec; j += blockDim.x) zv[j] = make_uint4(0u, 0u, 0u, 0u);
            continue;
        }
        const int64_t y = labels_c[r0 + r];
        float xy = 0.f;
        if (threadIdx.x == 0) xy = __bfloat162float(lg[r * ld + y]);  // (before anyone overwrites the row)
        uint4 u[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int j = threadIdx.x + i * 256;
            u[i] = j < nvec ? zv[j] : make_uint4(0u, 0u, 0u, 0u);
        }
        float m = -INFINITY;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int64_t c0 = (int64_t)(threadIdx.x + i * 256) * 8;
            const bf16* e = reinterpret_cast<const bf16*>(&u[i]);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (c0 + k < classes) m = fmaxf(m, __bfloat162float(e[k]));
        }
        m = warp_max(m);
        if (lane == 0) s_m[wid] = m;
        __syncthreads();
        float mm = s_m[0];
#pragma unroll
        for (int w = 1; w < 8; ++w) mm = fmaxf(mm, s_m[w]);
        float ssum = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int64_t c0 = (int64_t)(threadIdx.x + i * 256) * 8;
            const bf16* e = reinterpret_cast<const bf16*>(&u[i]);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (c0 + k < classes) ssum += __expf(__bfloat162float(e[k]) - mm);
        }
        ssum = warp_sum(ssum);
        if (lane == 0) s_s[wid] = ssum;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) tot += s_s[w];
        const float lse = mm + logf(tot);
        if (threadIdx.x == 0) row_loss[r0 + r] = lse - xy;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int j = threadIdx.x + i * 256;
            if (j < nvec) {
                bf16* e = reinterpret_cast<bf16*>(&u[i]);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int64_t col = (int64_t)j * 8 + k;
                    float g = 0.f;
                    if (col < classes) {
                        g = __expf(__bfloat162float(e[k]) - lse);
                        if (col == y) g -= 1.0f;
                        g *= inv_n;
                    }
                    e[k] = __float2bfloat16(g);
                }
                zv[j] = u[i];
            }
        }
        __syncthreads();  // s_m / s_s are rewritten for the next row
    }
}

// db[j] += sum over the chunk's live rows of dl[r][j]: block = 128 rows x 256 columns (2 per thread), one atomic
// per column per block.
__global__ void __launch_bounds__(128) lce_colsum_kernel(const bf16* __restrict__ dl, int64_t ld, float* __restrict__ db,
                                                         const int* __restrict__ count, int64_t r0, int64_t nr, int64_t classes) {
    const int64_t live = max((int64_t)0, min(nr, (int64_t)*count - r0));
    const int64_t rb = (int64_t)blockIdx.y * 128, re = min(live, rb + 128);
    const int64_t col = ((int64_t)blockIdx.x * 128 + threadIdx.x) * 2;
    if (rb >= live || col >= classes) return;
    float a0 = 0.f, a1 = 0.f;
    for (int64_t r = rb; r < re; ++r) {
        const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(dl + r * ld + col);
        a0 += __low2float(v);
        a1 += __high2float(v);
    }
    atomicAdd(db + col, a0);
    if (col + 1 < classes) atomicAdd(db + col + 1, a1);
}

__global__ void __launch_bounds__(1024) loss_finalize_kernel(const float* __restrict__ row_loss, const int* __restrict__ count,
                                                             float* __restrict__ out3) {
    __shared__ double s_l[32];
    const int rows = *count;
    double l = 0.0;
    for (int r = threadIdx.x; r < rows; r += blockDim.x) l += (double)row_loss[r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
    if ((threadIdx.x & 31) == 0) s_l[threadIdx.x >> 5] = l;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_l[w];
        out3[2] = (float)t;
        out3[0] = (float)(t / (double)out3[1]);
    }
}

static int64_t lce_chunk_rows(int64_t classes) {
    // bf16 logits (turned into dlogits in place) of a chunk within ~84 MB so they stay L2-resident between kernels
    int64_t rows = (84ll << 20) / (2 * pad8(classes));
    rows = std::max<int64_t>(128, rows / 128 * 128);
    return std::min<int64_t>(rows, 8192);
}

// Chunk schedule: the first chunk is L2-sized; an MLM batch labels ~15 % of its positions, so what follows is almost
// always empty and every further chunk only costs launches of kernels that find nothing to do.  Later chunks are
// therefore three times the previous one (32768 rows: 4352, 13056, 15360 instead of eight chunks); the logits buffer
// has the size of the largest, of which only the live rows are ever written.
static int64_t lce_max_chunk(int64_t rows, int64_t classes) {
    int64_t most = 0;
    for (int64_t r0 = 0, step = std::min(rows, lce_chunk_rows(classes)); r0 < rows; r0 += step, step *= 3) most = std::max(most, std::min(step, rows - r0));
    return std::max<int64_t>(most, 1);
}

size_t linear_ce_ws_bytes(int64_t rows, int c_in, int64_t classes) {
    const int64_t ch = lce_max_chunk(rows, classes);
    return align_up((size_t)rows * pad8(c_in) * 2, 256) + align_up((size_t)ch * pad8(classes) * 2, 256) +
           align_up((size_t)rows * c_in * 4, 256) + 4 * align_up((size_t)rows * 8, 256) + 256 +
           align_up((size_t)cdiv(rows, 1024) * 4, 256) +
           pio_gemm_bf16_ws_bytes(ch, c_in, classes) + 4096;
}

int linear_ce_fwd_bwd(Ctx& c, const float* x, const float* w16, const float* b, const int64_t* labels, float* out3, float* dx,
                      float* dw, float* db, int64_t rows, int c_in, int64_t classes, int64_t ignore) {
    const char* who = "pio_linear_ce_fwd_bwd";
    PIO_CHECK_ARG(rows < (1ll << 31), "%s: more than 2^31 rows", who);
    const int64_t cip = pad8(c_in), cop = pad8(classes);
    const int64_t ch0 = std::min(rows, lce_chunk_rows(classes)), ch = lce_max_chunk(rows, classes);
    const size_t mark = c.W.off;
    PIO_TAKE_BYTES(bf16, x16c, c, (size_t)rows * cip * 2, who);
    PIO_TAKE_BYTES(bf16, lg16, c, (size_t)ch * cop * 2, who);
    PIO_TAKE_BYTES(float, dxc, c, (size_t)rows * c_in * 4, who);
    PIO_TAKE_BYTES(float, row_loss, c, (size_t)rows * 4, who);
    LceIndex ix;
    PIO_TAKE_BYTES(int, idx, c, (size_t)rows * 4, who);
    PIO_TAKE_BYTES(int, pos, c, (size_t)rows * 4, who);
    PIO_TAKE_BYTES(int64_t, labels_c, c, (size_t)rows * 8, who);
    PIO_TAKE_BYTES(int, count, c, 256, who);
    PIO_TAKE_BYTES(int, block_counts, c, (size_t)cdiv(rows, 1024) * 4, who);
    ix.count = count; ix.idx = idx; ix.pos = pos; ix.labels = labels_c;

    const unsigned cblocks = (unsigned)cdiv(rows, 1024);
    lce_count_kernel<<<cblocks, 1024, 0, c.st>>>(labels, rows, classes, ignore, block_counts);
    PIO_LAUNCH_CHECK();
    lce_compact_kernel<<<cblocks, 1024, 0, c.st>>>(labels, rows, classes, ignore, block_counts, out3, ix);
    PIO_LAUNCH_CHECK();
    const int row_blocks = (int)std::min<int64_t>(cdiv(rows, 8), (int64_t)num_sms() * 8);
    lce_gather_kernel<<<row_blocks, 256, 0, c.st>>>(x, c_in, x16c, cip, ix, rows);
    PIO_LAUNCH_CHECK();
    // dW / db accumulate with atomics on the side stream while the main stream forms dx for the same chunk
    // (the destinations are zeroed up front so every chunk can simply add)
    PIO_CUDA(cudaMemsetAsync(dw, 0, (size_t)classes * c_in * sizeof(float), c.st));
    PIO_CUDA(cudaMemsetAsync(db, 0, (size_t)classes * sizeof(float), c.st));
    Side side(c, 2);
    for (int64_t r0 = 0, step = ch0; r0 < rows; r0 += step, step = std::min(3 * step, ch)) {
        const int64_t nr = std::min(step, rows - r0);
        const bf16* xc = x16c + r0 * cip;
        if (logits_persistent_supported(nr, classes, c_in, cip, cip, cop)) {
            // logits (bf16) = x W^T + b for the live rows of this chunk: one CTA per SM over the live tiles only
            PIO_TRY(logits_persistent(xc, cip, w16, cip, b, lg16, cop, nr, classes, c_in, count, (int)r0, c.st));
        } else {
            pio_gemm_bf16_desc d{};
            d.a = xc; d.b = w16; d.m = nr; d.n = classes; d.k = c_in; d.lda = cip; d.ldb = cip; d.a_kmajor = 1; d.b_kmajor = 1;
            d.c16 = lg16; d.ldc16 = cop; d.bias = b;
            d.dyn_rows = count; d.dyn_offset = r0; d.dyn_axis = 0;
            PIO_TRY(tc_gemm(c, d));
        }
        const int blocks = (int)std::min<int64_t>(nr, (int64_t)num_sms() * 8);
        if (cop <= 8 * 256 * 5)
            ce_chunk_reg_kernel<5><<<blocks, 256, 0, c.st>>>(lg16, cop, labels_c, out3, count, r0, nr, row_loss, classes);
        else if (cop <= 8 * 256 * 8)
            ce_chunk_reg_kernel<8><<<blocks, 256, 0, c.st>>>(lg16, cop, labels_c, out3, count, r0, nr, row_loss, classes);
        else
            ce_chunk_kernel<<<blocks, 256, 0, c.st>>>(lg16, cop, labels_c, out3, count, r0, nr, row_loss, classes);
        PIO_LAUNCH_CHECK();
        PIO_TRY(side.fork());
        if (c_in == 256 && side.active() && cdiv(classes, 128) <= 160) {
            // dW += dlogits^T x and db += column sums of dlogits over the live rows, one launch (pio_dw_batched.cu)
            DwBatchItem it{lg16, cop, (int)classes, xc, cip, c_in, dw, db};
            PIO_TRY(dw_batched(&it, 1, nr, side.stream(), count, (int)r0));
        } else {
            {  // dW += dlogits^T x   (contraction over the live rows)
                pio_gemm_bf16_desc d{};
                d.a = lg16; d.b = xc; d.m = classes; d.n = c_in; d.k = nr; d.lda = cop; d.ldb = cip; d.a_kmajor = 0; d.b_kmajor = 0;
                d.c32 = dw; d.ldc32 = c_in; d.accumulate = 2;
                d.dyn_rows = count; d.dyn_offset = r0; d.dyn_axis = 1;
                PIO_TRY(gemm_bf16(&d, nullptr, 0, side.stream()));
            }
            lce_colsum_kernel<<<dim3((unsigned)cdiv(cdiv(classes, 2), 128), (unsigned)cdiv(nr, 128)), 128, 0, side.stream2()>>>(
                lg16, cop, db, count, r0, nr, classes);
            PIO_LAUNCH_CHECK();
        }
        if (xw_splitk_supported(nr, c_in, classes, cop, cip)) {
            // dx (compacted) = dlogits W: 128 x 256 tiles so that dlogits is read once, split over the vocabulary
            if (!xw_splitk_stores(nr, classes)) PIO_CUDA(cudaMemsetAsync(dxc + r0 * c_in, 0, (size_t)nr * c_in * sizeof(float), c.st));
            PIO_TRY(xw_splitk(lg16, cop, w16, cip, dxc + r0 * c_in, nr, classes, c.st, count, (int)r0));
        } else {  // dx (compacted) = dlogits W
            pio_gemm_bf16_desc d{};
            d.a = lg16; d.b = w16; d.m = nr; d.n = c_in; d.k = classes; d.lda = cop; d.ldb = cip; d.a_kmajor = 1; d.b_kmajor = 0;
            d.c32 = dxc + r0 * c_in; d.ldc32 = c_in;
            d.dyn_rows = count; d.dyn_offset = r0; d.dyn_axis = 0;
            PIO_TRY(tc_gemm(c, d));
        }
        PIO_TRY(side.join());  // the next chunk overwrites lg16
    }
    lce_scatter_kernel<<<row_blocks, 256, 0, c.st>>>(dxc, c_in, dx, ix, rows);
    PIO_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 1024, 0, c.st>>>(row_loss, count, out3);
    PIO_LAUNCH_CHECK();
    c.W.off = mark;
    return PIO_OK;
}


// ---------------------------------------------------------------- decoder on the labelled rows only
// PerceiverDecoder's cross_attention_layer (reference model.py:229-237: Residual(CrossAttention) then Residual(mlp))
// for a DYNAMIC subset of the B*K query rows.  Every decoder query is independent of the others (its key / value
// set is the N latents of its own sample), and in the masked-LM step only the positions that carry a label reach
// the loss (an ignored position contributes exactly zero to the loss and to every gradient), so the decoder need
// not touch the other ~85-90 % of the (sample, position) pairs at all.  The labelled pairs are compacted on the
// device (label_compact; their count never reaches the host), and every kernel below is launched for the worst
// case (all B*K rows) and clamps / leaves at the device-side count (Ctx::dyn), so the sequence replays from a
// CUDA graph on any batch.  Row i of every array here is the pair idx[i] = b * l + pos.
//   * the query row is the learned array's row `pos` (batch independent): gathered, never broadcast;
//   * attention: one CTA per (sample, head) keeps that sample's K_h, V_h (N x dh, a few KB) in shared memory and
//     walks the sample's rows (contiguous in the compaction order) one warp per row, fp32 arithmetic — 4*N*E flop
//     per row, far too little for a tensor-core tile;
//   * all products with weights are the tcgen05 GEMM with a device-side row count.
struct DrSaved {
    float *xq, *stats, *lse, *x1;
    bf16 *hq, *q, *o16;
    int* seg;  // [batch + 1] first compact row of every sample (at most rows + 1 entries)
    void* mlp;
    size_t total;
    DrSaved(void* p, int64_t rows, int e, int heads) {
        ByteCarver k{(char*)p, 0};
        xq = k.f32((size_t)rows * e);
        stats = k.f32((size_t)rows * 2);
        hq = k.b16((size_t)rows * e);
        q = k.b16((size_t)rows * e);
        o16 = k.b16((size_t)rows * e);
        lse = k.f32((size_t)rows * heads);
        x1 = k.f32((size_t)rows * e);
        seg = (int*)k.take((size_t)(rows + 1) * 4);
        mlp = k.take(mlp_saved_bytes(rows, e));
        total = k.off;
    }
};
size_t decode_rows_saved_bytes(int64_t max_rows, int e, int heads) { return DrSaved(nullptr, max_rows, e, heads).total + 256; }
size_t decode_rows_ws_bytes(int64_t max_rows, int e, int heads) {
    return 3 * align_up((size_t)max_rows * e * 2, 256) + 3 * align_up((size_t)max_rows * e * 4, 256) + mlp_ws_bytes(max_rows, e) +
           scratch_bytes(max_rows, e, e) + 4096;
}
bool decode_rows_supported(int n, int e, int heads) {
    if (heads <= 0 || e % heads != 0) return false;
    const int dh = e / heads;
    // K_h, V_h as bf16 plus (backward) their fp32 gradients in one CTA's shared memory
    // K_h, V_h (bf16) and the per-warp row buffers of one CTA in shared memory; keys / channels per warp / lane in registers
    return e % 8 == 0 && dh % 8 == 0 && dh <= 128 && n >= 4 && n <= 128 && n % 4 == 0;
}

__global__ void __launch_bounds__(256) dr_pad_labels_kernel(int64_t* __restrict__ labels_c, const int* __restrict__ count, int64_t rows,
                                                             int64_t ignore) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)*count + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < rows; i += stride) labels_c[i] = ignore;
}

// idx (compact -> source row) and the labels of the compacted rows; entries past the count are `ignore`
int label_compact(const int64_t* labels, int64_t rows, int64_t classes, int64_t ignore, int* idx, int64_t* labels_c, int* count,
                  int* pos_scratch, int* block_counts, float* out3_scratch, cudaStream_t st) {
    const unsigned cblocks = (unsigned)cdiv(rows, 1024);
    LceIndex ix;
    ix.count = count; ix.idx = idx; ix.pos = pos_scratch; ix.labels = labels_c;
    lce_count_kernel<<<cblocks, 1024, 0, st>>>(labels, rows, classes, ignore, block_counts);
    PIO_LAUNCH_CHECK();
    lce_compact_kernel<<<cblocks, 1024, 0, st>>>(labels, rows, classes, ignore, block_counts, out3_scratch, ix);
    PIO_LAUNCH_CHECK();
    dr_pad_labels_kernel<<<(unsigned)std::min<int64_t>(cdiv(rows, 256), 1024), 256, 0, st>>>(labels_c, count, rows, ignore);
    PIO_LAUNCH_CHECK();
    return PIO_OK;
}

// xq[i] = queries[idx[i] % l] for i < count (warp per row)
__global__ void __launch_bounds__(256) dr_gather_kernel(const float* __restrict__ queries, const int* __restrict__ idx,
                                                         const int* __restrict__ count, float* __restrict__ xq, int l, int e) {
    const int n = *count, lane = threadIdx.x & 31;
    const int warps = gridDim.x * (blockDim.x >> 5);
    for (int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n; i += warps) {
        const float* src = queries + (int64_t)(idx[i] % l) * e;
        float* dst = xq + (int64_t)i * e;
        for (int j = lane * 4; j < e; j += 128) *reinterpret_cast<float4*>(dst + j) = *reinterpret_cast<const float4*>(src + j);
    }
}
// dqueries[idx[i] % l] += dxq[i] for i < count
__global__ void __launch_bounds__(256) dr_scatter_add_kernel(const float* __restrict__ dxq, const int* __restrict__ idx,
                                                              const int* __restrict__ count, float* __restrict__ dqueries, int l, int e) {
    const int n = *count, lane = threadIdx.x & 31;
    const int warps = gridDim.x * (blockDim.x >> 5);
    for (int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n; i += warps) {
        const float* src = dxq + (int64_t)i * e;
        float* dst = dqueries + (int64_t)(idx[i] % l) * e;
        for (int j = lane; j < e; j += 32) atomicAdd(dst + j, src[j]);
    }
}

__device__ __forceinline__ int dr_lower_bound(const int* __restrict__ idx, int n, int key) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (idx[mid] < key) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
// seg[b] = first compact row of sample b (b = 0 .. batch): one chain of dependent loads per sample here instead of two
// per CTA in the attention kernels
__global__ void __launch_bounds__(256) dr_segments_kernel(const int* __restrict__ idx, const int* __restrict__ count, int* __restrict__ seg,
                                                           int batch, int l) {
    const int cnt = *count;
    for (int b = blockIdx.x * blockDim.x + threadIdx.x; b <= batch; b += gridDim.x * blockDim.x) seg[b] = dr_lower_bound(idx, cnt, b * l);
}

// Shared-memory layout of one (sample, head): K and V as bf16 pairs with a row pitch of dh/2 + 4 words (16-byte
// aligned, and 16-byte reads with lane = key are conflict free: eight lanes x four words cover the 32 banks), per-warp
// q / dO (fp32, dh) and p / ds (fp32, n).  These kernels are bound by shared-memory instruction issue (a row's
// scores are dh-long dot products against every key), hence the 16-byte operand reads everywhere.
struct DrSmem {
    uint32_t *k, *v;
    float* wbuf;
    int kp;  // row pitch of K and V in words
};
__device__ __forceinline__ DrSmem dr_carve(uint8_t* base, int n, int dh) {
    DrSmem s;
    s.kp = dh / 2 + 4;
    s.k = (uint32_t*)base;
    s.v = s.k + (size_t)n * s.kp;
    s.wbuf = (float*)(s.v + (size_t)n * s.kp);  // [warps][2 * n + 2 * dh]
    return s;
}
static size_t dr_smem_bytes(int n, int dh, int warps) {
    return 2 * (size_t)n * (dh / 2 + 4) * 4 + (size_t)warps * (2 * n + 2 * dh) * 4 + 16;
}
__device__ __forceinline__ void dr_load_kv(const DrSmem& s, const bf16* __restrict__ kv, int b, int h, int n, int e, int dh) {
    const int hw = dh / 2;
    for (int t = threadIdx.x; t < n * hw; t += blockDim.x) {
        const int key = t / hw, w = t - key * hw;
        const bf16* row = kv + ((int64_t)b * n + key) * 2 * e + h * dh;
        s.k[key * s.kp + w] = reinterpret_cast<const uint32_t*>(row)[w];
        s.v[key * s.kp + w] = reinterpret_cast<const uint32_t*>(row + e)[w];
    }
}
__device__ __forceinline__ float2 bf2(uint32_t u) {
    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&u);
    return make_float2(__low2float(h), __high2float(h));
}
// acc += sum over 8 channels of x[c] * k[c]: x = two float4 (shared memory, warp-uniform), k = eight bf16 (one 16-byte read)
__device__ __forceinline__ float dot8(const float* x, const uint4& k, float acc) {
    const float4 x0 = *reinterpret_cast<const float4*>(x), x1 = *reinterpret_cast<const float4*>(x + 4);
    const float2 k0 = bf2(k.x), k1 = bf2(k.y), k2 = bf2(k.z), k3 = bf2(k.w);
    acc = fmaf(x0.x, k0.x, acc); acc = fmaf(x0.y, k0.y, acc); acc = fmaf(x0.z, k1.x, acc); acc = fmaf(x0.w, k1.y, acc);
    acc = fmaf(x1.x, k2.x, acc); acc = fmaf(x1.y, k2.y, acc); acc = fmaf(x1.z, k3.x, acc); acc = fmaf(x1.w, k3.y, acc);
    return acc;
}
// (a0, a1) += sum over keys [0, n) of w[key] * tile[key][pair]: w in shared memory (fp32, 16-byte reads), tile rows of pitch kp words
__device__ __forceinline__ void wsum_pairs(const float* w, const uint32_t* tile, int kp, int pair, int n, float& a0, float& a1) {
    float c0 = 0.f, c1 = 0.f;
    int key = 0;
    for (; key + 3 < n; key += 4) {
        const float4 ww = *reinterpret_cast<const float4*>(w + key);
        const float2 v0 = bf2(tile[key * kp + pair]), v1 = bf2(tile[(key + 1) * kp + pair]);
        const float2 v2 = bf2(tile[(key + 2) * kp + pair]), v3 = bf2(tile[(key + 3) * kp + pair]);
        a0 = fmaf(ww.x, v0.x, a0); a1 = fmaf(ww.x, v0.y, a1);
        c0 = fmaf(ww.y, v1.x, c0); c1 = fmaf(ww.y, v1.y, c1);
        a0 = fmaf(ww.z, v2.x, a0); a1 = fmaf(ww.z, v2.y, a1);
        c0 = fmaf(ww.w, v3.x, c0); c1 = fmaf(ww.w, v3.y, c1);
    }
    for (; key < n; ++key) {
        const float2 v0 = bf2(tile[key * kp + pair]);
        a0 = fmaf(w[key], v0.x, a0); a1 = fmaf(w[key], v0.y, a1);
    }
    a0 += c0;
    a1 += c1;
}

constexpr int DR_FWD_WARPS = 8, DR_FWD_SPLIT = 2, DR_BWD_WARPS = 16;

// grid (heads, batch, DR_FWD_SPLIT): o16[i, h*dh ..] and lse[i, h] for every DR_FWD_SPLIT-th batch of rows of sample b
__global__ void __launch_bounds__(DR_FWD_WARPS * 32) dr_attn_fwd_kernel(const bf16* __restrict__ q16, const bf16* __restrict__ kv,
                                                                         const int* __restrict__ seg, const int* __restrict__ count,
                                                                         bf16* __restrict__ o16, float* __restrict__ lse, int l, int n, int e,
                                                                         int heads, float scale, int64_t max_rows) {
    extern __shared__ uint8_t dr_raw[];
    const int h = blockIdx.x, b = blockIdx.y, dh = e / heads, hw = dh / 2;
    const int cnt = *count;
    const int lo = seg[b], hi = seg[b + 1];
    if (h == 0 && b == 0 && blockIdx.z == 0) {  // rows between the count and the next multiple of 128: zero (operand padding of dWo = dY^T O)
        const int64_t upto = min(max_rows, ((int64_t)cnt + 127) / 128 * 128);
        for (int64_t t = (int64_t)cnt * e + threadIdx.x; t < upto * e; t += blockDim.x) o16[t] = __float2bfloat16(0.f);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int first = lo + blockIdx.z * DR_FWD_WARPS + wid;
    if (lo + (int)blockIdx.z * DR_FWD_WARPS >= hi) return;
    const DrSmem s = dr_carve(dr_raw, n, dh);
    dr_load_kv(s, kv, b, h, n, e, dh);
    __syncthreads();
    float* qs = s.wbuf + (size_t)wid * (2 * n + 2 * dh);
    float* ps = qs + 2 * dh;
    for (int i = first; i < hi; i += DR_FWD_WARPS * DR_FWD_SPLIT) {
        const bf16* qrow = q16 + (int64_t)i * e + h * dh;
        for (int d = lane; d < dh; d += 32) qs[d] = __bfloat162float(qrow[d]) * scale;
        __syncwarp();
        float mx = -INFINITY;
        for (int key = lane; key < n; key += 32) {
            float a0 = 0.f, a1 = 0.f;
            const uint4* kr = reinterpret_cast<const uint4*>(s.k + key * s.kp);
            for (int v = 0; v + 1 < dh / 8; v += 2) {
                a0 = dot8(qs + 8 * v, kr[v], a0);
                a1 = dot8(qs + 8 * v + 8, kr[v + 1], a1);
            }
            if ((dh / 8) & 1) a0 = dot8(qs + dh - 8, kr[dh / 8 - 1], a0);
            const float acc = a0 + a1;
            ps[key] = acc;
            mx = fmaxf(mx, acc);
        }
        mx = warp_max(mx);
        float sum = 0.f;
        for (int key = lane; key < n; key += 32) {
            const float pv = __expf(ps[key] - mx);
            ps[key] = pv;
            sum += pv;
        }
        sum = warp_sum(sum);
        __syncwarp();
        const float inv = 1.0f / sum;
        for (int w = lane; w < hw; w += 32) {
            float a0 = 0.f, a1 = 0.f;
            wsum_pairs(ps, s.v, s.kp, w, n, a0, a1);
            __nv_bfloat162 o2 = __floats2bfloat162_rn(a0 * inv, a1 * inv);
            reinterpret_cast<uint32_t*>(o16 + (int64_t)i * e + h * dh)[w] = *reinterpret_cast<uint32_t*>(&o2);
        }
        if (lane == 0) lse[(int64_t)i * heads + h] = mx + __logf(sum);
        __syncwarp();
    }
}

// grid (heads, batch): dq16 rows of sample b, and dK_h / dV_h of (b, h) as bf16 into dkv (B, N, 2E); P recomputed from lse.
// Rows go through in batches of sixteen (one per warp): phase A, warp = row, recomputes the row's probabilities and
// score gradients into the warp's buffer and writes dq; phase B, warp = a slice of KPW keys, lane = CPL channels, adds
// the batch's rank-1 terms into dK / dV accumulators held in REGISTERS for the whole kernel (no two warps own the same
// key: no atomics, no shared-memory round trips — the first version did this with shared-memory atomics from the row
// warps and took 450 us at the benchmark shape, a shared-memory read-modify-write version 297 us).
template <int KPW, int CPL>
__global__ void __launch_bounds__(DR_BWD_WARPS * 32) dr_attn_bwd_kernel(const bf16* __restrict__ q16, const bf16* __restrict__ kv,
                                                                         const bf16* __restrict__ o16, const bf16* __restrict__ do16,
                                                                         const float* __restrict__ lse, const int* __restrict__ seg,
                                                                         const int* __restrict__ count, bf16* __restrict__ dq16,
                                                                         bf16* __restrict__ dkv, int l, int n, int e, int heads, float scale,
                                                                         int64_t max_rows) {
    extern __shared__ uint8_t dr_raw[];
    const int h = blockIdx.x, b = blockIdx.y, dh = e / heads, hw = dh / 2;
    const int cnt = *count;
    const int lo = seg[b], hi = seg[b + 1];
    if (h == 0 && b == 0) {
        const int64_t upto = min(max_rows, ((int64_t)cnt + 127) / 128 * 128);
        for (int64_t t = (int64_t)cnt * e + threadIdx.x; t < upto * e; t += blockDim.x) dq16[t] = __float2bfloat16(0.f);
    }
    const DrSmem s = dr_carve(dr_raw, n, dh);
    if (lo < hi) dr_load_kv(s, kv, b, h, n, e, dh);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int wstride = 2 * n + 2 * dh;
    float* qs = s.wbuf + (size_t)wid * wstride;  // q * scale
    float* gs = qs + dh;                          // dO
    float* ps = qs + 2 * dh;                      // p
    float* dss = ps + n;                          // ds (gradient of the scaled score)
    const int k0 = wid * KPW;                     // keys [k0, k0 + KPW) belong to this warp in phase B
    float dkr[KPW][CPL], dvr[KPW][CPL];
#pragma unroll
    for (int a = 0; a < KPW; ++a)
#pragma unroll
        for (int c = 0; c < CPL; ++c) dkr[a][c] = dvr[a][c] = 0.f;
    for (int base = lo; base < hi; base += DR_BWD_WARPS) {
        const int i = base + wid;
        if (i < hi) {  // ---- phase A (warp-uniform)
            const int64_t off = (int64_t)i * e + h * dh;
            for (int d = lane; d < dh; d += 32) {
                qs[d] = __bfloat162float(q16[off + d]) * scale;
                gs[d] = __bfloat162float(do16[off + d]);
            }
            const float lse_i = lse[(int64_t)i * heads + h];
            __syncwarp();
            // The row term delta = dO . O is formed as sum_k p_k dP_k from the SAME recomputed probabilities that multiply
            // it below: then sum_k dS_k = 0 holds to fp32 rounding.  With delta taken from the bf16 copy of O its rounding
            // error e leaves sum_k dS_k = -e, a component of dq along the sample's mean key that is common to all rows and
            // does not average out — it showed as an O(1) error of the decoder-query LayerNorm gradients.
            float delta = 0.f;
            for (int key = lane; key < n; key += 32) {
                float sc = 0.f, dp = 0.f;
                const uint4* kr = reinterpret_cast<const uint4*>(s.k + key * s.kp);
                const uint4* vr = reinterpret_cast<const uint4*>(s.v + key * s.kp);
                for (int v = 0; v < dh / 8; ++v) {
                    sc = dot8(qs + 8 * v, kr[v], sc);
                    dp = dot8(gs + 8 * v, vr[v], dp);
                }
                const float pv = __expf(sc - lse_i);
                ps[key] = pv;
                dss[key] = dp;
                delta = fmaf(pv, dp, delta);
            }
            delta = warp_sum(delta);
            for (int key = lane; key < n; key += 32) dss[key] = ps[key] * (dss[key] - delta);
            __syncwarp();
            for (int w = lane; w < hw; w += 32) {  // dq = scale * sum_key ds * K[key]
                float a0 = 0.f, a1 = 0.f;
                wsum_pairs(dss, s.k, s.kp, w, n, a0, a1);
                __nv_bfloat162 o2 = __floats2bfloat162_rn(a0 * scale, a1 * scale);
                reinterpret_cast<uint32_t*>(dq16 + off)[w] = *reinterpret_cast<uint32_t*>(&o2);
            }
        }
        __syncthreads();
        // ---- phase B: dV[key] += p dO ; dK[key] += ds * scale * q (qs carries the scale), for this warp's keys
        const int nrows = min(DR_BWD_WARPS, hi - base);
        for (int r = 0; r < nrows; ++r) {
            const float* rq = s.wbuf + (size_t)r * wstride;
            const float* rg = rq + dh;
            const float* rp = rq + 2 * dh;
            const float* rds = rp + n;
            float qv[CPL], gv[CPL];
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
                const int d = lane + 32 * c;
                qv[c] = d < dh ? rq[d] : 0.f;
                gv[c] = d < dh ? rg[d] : 0.f;
            }
#pragma unroll
            for (int a = 0; a < KPW; ++a) {
                const int key = min(k0 + a, n - 1);  // (keys past n are never written out)
                const float pv = rp[key], ds = rds[key];
#pragma unroll
                for (int c = 0; c < CPL; ++c) {
                    dvr[a][c] = fmaf(pv, gv[c], dvr[a][c]);
                    dkr[a][c] = fmaf(ds, qv[c], dkr[a][c]);
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < KPW; ++a) {
        const int key = k0 + a;
        if (key < n) {
            bf16* row = dkv + ((int64_t)b * n + key) * 2 * e + h * dh;
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
                const int d = lane + 32 * c;
                if (d < dh) {
                    row[d] = __float2bfloat16(dkr[a][c]);
                    row[e + d] = __float2bfloat16(dvr[a][c]);
                }
            }
        }
    }
}

typedef void (*DrBwdKernel)(const bf16*, const bf16*, const bf16*, const bf16*, const float*, const int*, const int*, bf16*, bf16*, int, int,
                            int, int, float, int64_t);
static DrBwdKernel dr_bwd_pick(int n, int dh) {
    const int kpw = n <= 2 * DR_BWD_WARPS ? 2 : (n <= 4 * DR_BWD_WARPS ? 4 : 8);
    const int cpl = dh <= 32 ? 1 : (dh <= 64 ? 2 : 4);
#define PIO_DR(K, C) if (kpw == K && cpl == C) return dr_attn_bwd_kernel<K, C>
    PIO_DR(2, 1); PIO_DR(2, 2); PIO_DR(2, 4); PIO_DR(4, 1); PIO_DR(4, 2); PIO_DR(4, 4); PIO_DR(8, 1); PIO_DR(8, 2); PIO_DR(8, 4);
#undef PIO_DR
    return nullptr;
}
static int dr_configure(int n, int dh) {
    const size_t need_f = dr_smem_bytes(n, dh, DR_FWD_WARPS), need_b = dr_smem_bytes(n, dh, DR_BWD_WARPS);
    if (need_f > 48 * 1024) PIO_CUDA(cudaFuncSetAttribute(dr_attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need_f));
    if (need_b > 48 * 1024) PIO_CUDA(cudaFuncSetAttribute(dr_bwd_pick(n, dh), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need_b));
    return PIO_OK;
}

int decode_rows_fwd(Ctx& c, const float* queries, const void* kv, const int* idx, float* y, const pio_attn_weights* aw,
                    const pio_mlp_weights* mw, void* saved, int64_t max_rows, int l, int batch, int n, int e, int heads) {
    const char* who = "pio_decode_rows_fwd";
    PIO_CHECK_ARG(c.dyn != nullptr, "%s: needs the device-side row count", who);
    PIO_CHECK_ARG(decode_rows_supported(n, e, heads), "%s: unsupported shape (n %d, e %d, heads %d)", who, n, e, heads);
    PIO_CHECK_ARG(max_rows == (int64_t)batch * l && max_rows < (1ll << 31), "%s: max_rows must be batch * l", who);
    DrSaved s(saved, max_rows, e, heads);
    const int dh = e / heads;
    PIO_TRY(dr_configure(n, dh));
    const int row_blocks = (int)std::min<int64_t>(cdiv(max_rows, 8), (int64_t)num_sms() * 8);
    dr_gather_kernel<<<row_blocks, 256, 0, c.st>>>(queries, idx, c.dyn, s.xq, l, e);
    PIO_LAUNCH_CHECK();
    dr_segments_kernel<<<(unsigned)cdiv(batch + 1, 256), 256, 0, c.st>>>(idx, c.dyn, s.seg, batch, l);
    PIO_LAUNCH_CHECK();
    PIO_TRY(layernorm_fwd_bf16(s.xq, aw->ln_w, aw->ln_b, s.hq, e, s.stats, max_rows, e, LN_EPS, c.st, c.dyn));
    PIO_TRY(tc_xwT(c, s.hq, e, (const bf16*)aw->wq, e, max_rows, e, e, nullptr, 0, s.q, e, aw->bq, nullptr, 0, 0, nullptr, 0));
    if (dr_attn_bwd_tc_supported(n, e, heads)) {  // tcgen05 version (pio_dr_attn_tc.cu)
        PIO_TRY(dr_attn_fwd_tc(s.q, kv, s.seg, c.dyn, s.o16, s.lse, batch, n, e, heads, max_rows, c.st));
    } else {
        dr_attn_fwd_kernel<<<dim3((unsigned)heads, (unsigned)batch, DR_FWD_SPLIT), DR_FWD_WARPS * 32, dr_smem_bytes(n, dh, DR_FWD_WARPS), c.st>>>(
            s.q, (const bf16*)kv, s.seg, c.dyn, s.o16, s.lse, l, n, e, heads, 1.0f / sqrtf((float)dh), max_rows);
        PIO_LAUNCH_CHECK();
    }
    PIO_TRY(tc_xwT(c, s.o16, e, (const bf16*)aw->wo, e, max_rows, e, e, s.x1, e, nullptr, 0, aw->bo, s.xq, e, 0, nullptr, 0));
    PIO_TRY(mlp_fwd(c, s.x1, y, mw, s.mlp, max_rows, e, 1));
    return PIO_OK;
}

int decode_rows_bwd(Ctx& c, const float* dy, const void* kv, const int* idx, float* dqueries, void* dkv, const pio_attn_weights* aw,
                    const pio_mlp_weights* mw, const void* saved, const pio_attn_grads* ag, const pio_mlp_grads* mg, int acc,
                    int64_t max_rows, int l, int batch, int n, int e, int heads) {
    const char* who = "pio_decode_rows_bwd";
    PIO_CHECK_ARG(c.dyn != nullptr && acc == 2, "%s: needs the device-side row count and atomic gradient accumulation", who);
    PIO_CHECK_ARG(decode_rows_supported(n, e, heads), "%s: unsupported shape (n %d, e %d, heads %d)", who, n, e, heads);
    DrSaved s((void*)saved, max_rows, e, heads);
    const int dh = e / heads;
    PIO_TRY(dr_configure(n, dh));
    const size_t mark = c.W.off;
    PIO_TAKE_BYTES(float, dx1, c, (size_t)max_rows * e * 4, who);
    PIO_TAKE_BYTES(bf16, dy16, c, (size_t)max_rows * e * 2, who);
    PIO_TAKE_BYTES(bf16, d_o16, c, (size_t)max_rows * e * 2, who);
    PIO_TAKE_BYTES(bf16, dq16, c, (size_t)max_rows * e * 2, who);
    PIO_TAKE_BYTES(float, dhq, c, (size_t)max_rows * e * 4, who);
    PIO_TAKE_BYTES(float, dxq, c, (size_t)max_rows * e * 4, who);
    PIO_TRY(mlp_bwd(c, dy, s.x1, dx1, mw, s.mlp, mg, acc, max_rows, e, 1));
    Side side(c, acc);
    PIO_TRY(cast_colsum_f32(dx1, e, dy16, e, ag->bo, acc, max_rows, e, c.rest(), c.rest_bytes(), c.st, c.dyn));
    PIO_TRY(side.fork());
    PIO_TRY(tc_dW(c, dy16, e, s.o16, e, ag->wo, max_rows, e, e, acc, side.stream()));
    PIO_TRY(tc_dyW(c, dy16, e, (const bf16*)aw->wo, e, max_rows, e, e, nullptr, 0, d_o16, e, 0, nullptr, 0, 0));
    if (dr_attn_bwd_tc_supported(n, e, heads)) {  // tcgen05 version (pio_dr_attn_tc.cu): head size 64, up to 128 keys
        PIO_TRY(dr_attn_bwd_tc(s.q, kv, d_o16, s.lse, s.seg, c.dyn, dq16, dkv, batch, n, e, heads, max_rows, c.st));
    } else {
        dr_bwd_pick(n, dh)<<<dim3((unsigned)heads, (unsigned)batch), DR_BWD_WARPS * 32, dr_smem_bytes(n, dh, DR_BWD_WARPS), c.st>>>(
            s.q, (const bf16*)kv, s.o16, d_o16, s.lse, s.seg, c.dyn, dq16, (bf16*)dkv, l, n, e, heads, 1.0f / sqrtf((float)dh), max_rows);
        PIO_LAUNCH_CHECK();
    }
    PIO_TRY(side.fork());
    PIO_TRY(side_colsum(c, side, dq16, e, ag->bq, acc, max_rows, e));
    PIO_TRY(tc_dW(c, dq16, e, s.hq, e, ag->wq, max_rows, e, e, acc, side.stream()));
    PIO_TRY(tc_dyW(c, dq16, e, (const bf16*)aw->wq, e, max_rows, e, e, dhq, e, nullptr, 0, 0, nullptr, 0, 0));
    PIO_TRY(layernorm_bwd(dhq, s.xq, s.stats, aw->ln_w, dx1, dxq, ag->ln_w, ag->ln_b, acc, max_rows, e, c.rest(), c.rest_bytes(), c.st,
                          c.dyn));
    const int row_blocks = (int)std::min<int64_t>(cdiv(max_rows, 8), (int64_t)num_sms() * 8);
    dr_scatter_add_kernel<<<row_blocks, 256, 0, c.st>>>(dxq, idx, c.dyn, dqueries, l, e);
    PIO_LAUNCH_CHECK();
    PIO_TRY(side.join());
    c.W.off = mark;
    return PIO_OK;
}

}  // namespace bf
}  // namespace pio
