// Fused operators: one C-ABI call per reference module (Residual(mlp), the two halves of
// Residual(CrossAttention), the latent self-attention block, the output linear).  Each call
// enqueues its kernel sequence on the caller's stream; saved-for-backward state and scratch are
// carved from caller-owned buffers with a fixed layout that only this file knows.
#include "pio_ops_common.cuh"

namespace pio {

namespace {

// y[rows, n] = act(x[rows,k] W[n,k]^T + bias) (+ residual)
int gemm_xwT(Ctx& c, const float* x, int64_t ldx, const float* w, int64_t ldw, float* y, int64_t ldy, int64_t rows,
             int64_t n, int64_t k, const float* bias, const float* residual, int64_t ldr, int act, float* aux,
             int64_t ldaux, int accumulate = 0, float drop_p = 0.f, uint64_t drop_seed = 0) {
    pio_gemm_desc d{};
    d.dropout_p = drop_p; d.dropout_seed = drop_seed;
    d.a = x; d.b = w; d.c = y;
    d.m = rows; d.n = n; d.k = k;
    d.lda = ldx; d.ldb = ldw; d.ldc = ldy;
    d.a_kmajor = 1; d.b_kmajor = 1;
    d.bias = bias; d.residual = residual; d.ldr = ldr; d.act = act; d.aux = aux; d.ldaux = ldaux;
    d.accumulate = accumulate;
    return pio_gemm_f32(&d, c.rest(), c.rest_bytes(), c.st);
}

// dx[rows, kin] (+)= dy[rows, nout] W[nout, kin]   (optionally multiplied by GELU'(aux))
int gemm_dyW(Ctx& c, const float* dy, int64_t lddy, const float* w, int64_t ldw, float* dx, int64_t lddx, int64_t rows,
             int64_t nout, int64_t kin, int act, float* aux, int64_t ldaux, int accumulate) {
    pio_gemm_desc d{};
    d.a = dy; d.b = w; d.c = dx;
    d.m = rows; d.n = kin; d.k = nout;
    d.lda = lddy; d.ldb = ldw; d.ldc = lddx;
    d.a_kmajor = 1; d.b_kmajor = 0;
    d.act = act; d.aux = aux; d.ldaux = ldaux; d.accumulate = accumulate;
    return pio_gemm_f32(&d, c.rest(), c.rest_bytes(), c.st);
}

// dw[nout, kin] (+)= dy[rows, nout]^T x[rows, kin] ; db[nout] (+)= colsum(dy)
int gemm_dW(Ctx& c, const float* dy, int64_t lddy, const float* x, int64_t ldx, float* dw, float* db, int64_t rows,
            int64_t nout, int64_t kin, int accumulate) {
    pio_gemm_desc d{};
    d.a = dy; d.b = x; d.c = dw;
    d.m = nout; d.n = kin; d.k = rows;
    d.lda = lddy; d.ldb = ldx; d.ldc = kin;
    d.a_kmajor = 0; d.b_kmajor = 0;
    d.accumulate = accumulate;
    d.rowsum_a = db;
    return pio_gemm_f32(&d, c.rest(), c.rest_bytes(), c.st);
}

size_t dW_scratch(int64_t nout, int64_t kin, int64_t rows) { return pio_gemm_f32_ws_bytes(nout, kin, rows); }

// gradient entering the branch of a Residual with dropout: dy masked like the forward output was
int branch_grad(Ctx& c, const float* dy, int64_t count, int residual, const float** out, const char* who) {
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

// ---------------------------------------------------------------- mlp
struct MlpSaved {
    float *stats, *h0, *h1, *g;
    size_t total;
    MlpSaved(void* p, int64_t rows, int c) {
        Carver k{(float*)p, 0};
        stats = k.take((size_t)rows * 2);
        h0 = k.take((size_t)rows * c);
        h1 = k.take((size_t)rows * c);
        g = k.take((size_t)rows * c);
        total = k.off;
    }
    static size_t floats(int64_t rows, int c) { return MlpSaved(nullptr, rows, c).total; }
};

int mlp_fwd(Ctx& c, const float* x, float* y, const pio_mlp_weights* w, void* saved, int64_t rows, int ch,
            int residual = 1) {
    MlpSaved s(saved, rows, ch);
    PIO_TRY(pio_layernorm_fwd(x, w->ln_w, w->ln_b, s.h0, s.stats, rows, ch, LN_EPS, c.st));
    PIO_TRY(gemm_xwT(c, s.h0, ch, w->w1, ch, s.g, ch, rows, ch, ch, w->b1, nullptr, 0, 1, s.h1, ch));
    PIO_TRY(gemm_xwT(c, s.g, ch, w->w2, ch, y, ch, rows, ch, ch, w->b2, residual ? x : nullptr, ch, 0, nullptr, 0, 0,
                     residual ? c.drop_p : 0.f, c.drop_seed + 1));
    return PIO_OK;
}

size_t mlp_bwd_ws(int64_t rows, int ch) {
    return 3 * align_up((size_t)rows * ch * 4, 256) + pio_layernorm_bwd_ws_bytes(rows, ch) + dW_scratch(ch, ch, rows) + 1024;
}

int mlp_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_mlp_weights* w, const void* saved,
            const pio_mlp_grads* g, int acc, int64_t rows, int ch, int residual = 1) {
    MlpSaved s((void*)saved, rows, ch);
    const size_t mark = c.W.off;
    PIO_TAKE(dh1, c, (size_t)rows * ch, "pio_mlp_bwd");
    PIO_TAKE(dh0, c, (size_t)rows * ch, "pio_mlp_bwd");
    const float* dyb;
    PIO_TRY(branch_grad(c, dy, rows * ch, residual, &dyb, "pio_mlp_bwd"));
    PIO_TRY(gemm_dyW(c, dyb, ch, w->w2, ch, dh1, ch, rows, ch, ch, 2, s.h1, ch, 0));
    PIO_TRY(gemm_dW(c, dyb, ch, s.g, ch, g->w2, g->b2, rows, ch, ch, acc));
    PIO_TRY(gemm_dyW(c, dh1, ch, w->w1, ch, dh0, ch, rows, ch, ch, 0, nullptr, 0, 0));
    PIO_TRY(gemm_dW(c, dh1, ch, s.h0, ch, g->w1, g->b1, rows, ch, ch, acc));
    PIO_TRY(pio_layernorm_bwd(dh0, x, s.stats, w->ln_w, residual ? dy : nullptr, dx, g->ln_w, g->ln_b, acc, rows, ch,
                              c.rest(), c.rest_bytes(), c.st));
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- kv projection
struct KvSaved {
    float *stats, *h0;
    size_t total;
    KvSaved(void* p, int64_t rows, int ckv) {
        Carver k{(float*)p, 0};
        stats = k.take((size_t)rows * 2);
        h0 = k.take((size_t)rows * ckv);
        total = k.off;
    }
    static size_t floats(int64_t rows, int ckv) { return KvSaved(nullptr, rows, ckv).total; }
};

bool kv_packed(const float* wk, const float* wv, const float* bk, const float* bv, int e, int ckv) {
    return wv == wk + (size_t)e * ckv && bv == bk + e;
}

int kv_project_fwd(Ctx& c, const float* x_kv, float* kv, const pio_attn_weights* w, void* saved, int64_t rows, int ckv,
                   int e) {
    KvSaved s(saved, rows, ckv);
    PIO_TRY(pio_layernorm_fwd(x_kv, w->kv_ln_w, w->kv_ln_b, s.h0, s.stats, rows, ckv, LN_EPS, c.st));
    if (kv_packed(w->wk, w->wv, w->bk, w->bv, e, ckv)) {
        PIO_TRY(gemm_xwT(c, s.h0, ckv, w->wk, ckv, kv, 2 * e, rows, 2 * e, ckv, w->bk, nullptr, 0, 0, nullptr, 0));
    } else {
        PIO_TRY(gemm_xwT(c, s.h0, ckv, w->wk, ckv, kv, 2 * e, rows, e, ckv, w->bk, nullptr, 0, 0, nullptr, 0));
        PIO_TRY(gemm_xwT(c, s.h0, ckv, w->wv, ckv, kv + e, 2 * e, rows, e, ckv, w->bv, nullptr, 0, 0, nullptr, 0));
    }
    return PIO_OK;
}

size_t kv_project_bwd_ws(int64_t rows, int ckv, int e) {
    return align_up((size_t)rows * ckv * 4, 256) + pio_layernorm_bwd_ws_bytes(rows, ckv) + dW_scratch(2 * e, ckv, rows) + 1024;
}

int kv_project_bwd(Ctx& c, const float* dkv, const float* x_kv, float* dx_kv, const pio_attn_weights* w,
                   const void* saved, const pio_attn_grads* g, int acc, int64_t rows, int ckv, int e) {
    KvSaved s((void*)saved, rows, ckv);
    const size_t mark = c.W.off;
    PIO_TAKE(dh0, c, (size_t)rows * ckv, "pio_kv_project_bwd");
    const bool wpk = kv_packed(w->wk, w->wv, w->bk, w->bv, e, ckv);
    const bool gpk = kv_packed(g->wk, g->wv, g->bk, g->bv, e, ckv);
    if (wpk) {
        PIO_TRY(gemm_dyW(c, dkv, 2 * e, w->wk, ckv, dh0, ckv, rows, 2 * e, ckv, 0, nullptr, 0, 0));
    } else {
        PIO_TRY(gemm_dyW(c, dkv, 2 * e, w->wk, ckv, dh0, ckv, rows, e, ckv, 0, nullptr, 0, 0));
        PIO_TRY(gemm_dyW(c, dkv + e, 2 * e, w->wv, ckv, dh0, ckv, rows, e, ckv, 0, nullptr, 0, 1));
    }
    if (gpk) {
        PIO_TRY(gemm_dW(c, dkv, 2 * e, s.h0, ckv, g->wk, g->bk, rows, 2 * e, ckv, acc));
    } else {
        PIO_TRY(gemm_dW(c, dkv, 2 * e, s.h0, ckv, g->wk, g->bk, rows, e, ckv, acc));
        PIO_TRY(gemm_dW(c, dkv + e, 2 * e, s.h0, ckv, g->wv, g->bv, rows, e, ckv, acc));
    }
    PIO_TRY(pio_layernorm_bwd(dh0, x_kv, s.stats, w->kv_ln_w, nullptr, dx_kv, g->kv_ln_w, g->kv_ln_b, acc, rows, ckv,
                              c.rest(), c.rest_bytes(), c.st));
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- query side of cross-attention
struct XaSaved {
    float *stats, *hq, *q, *o, *lse;
    size_t total;
    XaSaved(void* p, int batch, int n, int e, int heads) {
        const size_t r = (size_t)batch * n;
        Carver k{(float*)p, 0};
        stats = k.take(r * 2);
        hq = k.take(r * e);
        q = k.take(r * e);
        o = k.take(r * e);
        lse = k.take((size_t)batch * heads * n);
        total = k.off;
    }
    static size_t floats(int batch, int n, int e, int heads) { return XaSaved(nullptr, batch, n, e, heads).total; }
};

void fill_attn(pio_attn_desc& a, const float* q, int64_t q_rs, const float* k, const float* v, int64_t kv_rs,
               const uint8_t* pad, float* o, float* lse, int batch, int heads, int n, int m, int e) {
    a.q = q; a.k = k; a.v = v; a.pad = pad; a.o = o; a.lse = lse;
    a.q_rs = q_rs; a.q_bs = (int64_t)n * q_rs;
    a.k_rs = kv_rs; a.k_bs = (int64_t)m * kv_rs;
    a.v_rs = kv_rs; a.v_bs = (int64_t)m * kv_rs;
    a.o_rs = e; a.o_bs = (int64_t)n * e;
    a.batch = batch; a.heads = heads; a.n = n; a.m = m; a.dh = e / heads;
    a.scale = 1.0f / sqrtf((float)(e / heads));
}

int cross_attend_fwd(Ctx& c, const float* x_q, const float* kv, const uint8_t* pad, float* y, const pio_attn_weights* w,
                     void* saved, int batch, int n, int m, int e, int heads, int residual = 1) {
    XaSaved s(saved, batch, n, e, heads);
    const int64_t rows = (int64_t)batch * n;
    PIO_TRY(pio_layernorm_fwd(x_q, w->ln_w, w->ln_b, s.hq, s.stats, rows, e, LN_EPS, c.st));
    PIO_TRY(gemm_xwT(c, s.hq, e, w->wq, e, s.q, e, rows, e, e, w->bq, nullptr, 0, 0, nullptr, 0));
    pio_attn_desc a{};
    fill_attn(a, s.q, e, kv, kv + e, 2 * e, pad, s.o, s.lse, batch, heads, n, m, e);
    AttnBf16Out ex;
    ex.drop = c.attn_drop();
    PIO_TRY(attn_fwd_simt(&a, c.st, &ex));
    PIO_TRY(gemm_xwT(c, s.o, e, w->wo, e, y, e, rows, e, e, w->bo, residual ? x_q : nullptr, e, 0, nullptr, 0, 0,
                     residual ? c.drop_p : 0.f, c.drop_seed + 1));
    return PIO_OK;
}

size_t cross_attend_bwd_ws(int batch, int n, int e, int heads) {
    const size_t r = (size_t)batch * n;
    return 4 * align_up(r * e * 4, 256) + align_up((size_t)batch * heads * n * 4, 256) +
           pio_layernorm_bwd_ws_bytes(r, e) + dW_scratch(e, e, r) + 1024;
}

int cross_attend_bwd(Ctx& c, const float* dy, const float* x_q, const float* kv, const uint8_t* pad, float* dx_q,
                     float* dkv, const pio_attn_weights* w, const void* saved, const pio_attn_grads* g, int acc,
                     int batch, int n, int m, int e, int heads, int residual = 1) {
    XaSaved s((void*)saved, batch, n, e, heads);
    const int64_t rows = (int64_t)batch * n;
    const size_t mark = c.W.off;
    PIO_TAKE(d_o, c, (size_t)rows * e, "pio_cross_attend_bwd");
    PIO_TAKE(dq, c, (size_t)rows * e, "pio_cross_attend_bwd");
    PIO_TAKE(dhq, c, (size_t)rows * e, "pio_cross_attend_bwd");
    PIO_TAKE(delta, c, (size_t)batch * heads * n, "pio_cross_attend_bwd");
    const float* dyb;
    PIO_TRY(branch_grad(c, dy, rows * e, residual, &dyb, "pio_cross_attend_bwd"));
    PIO_TRY(gemm_dyW(c, dyb, e, w->wo, e, d_o, e, rows, e, e, 0, nullptr, 0, 0));
    PIO_TRY(gemm_dW(c, dyb, e, s.o, e, g->wo, g->bo, rows, e, e, acc));
    pio_attn_bwd_desc b{};
    fill_attn(b.fwd, s.q, e, kv, kv + e, 2 * e, pad, s.o, s.lse, batch, heads, n, m, e);
    b.d_o = d_o; b.dq = dq; b.dk = dkv; b.dv = dkv + e; b.delta = delta;
    AttnBf16Out exb;
    exb.drop = c.attn_drop();
    PIO_TRY(attn_bwd_simt(&b, c.st, &exb));
    PIO_TRY(gemm_dyW(c, dq, e, w->wq, e, dhq, e, rows, e, e, 0, nullptr, 0, 0));
    PIO_TRY(gemm_dW(c, dq, e, s.hq, e, g->wq, g->bq, rows, e, e, acc));
    PIO_TRY(pio_layernorm_bwd(dhq, x_q, s.stats, w->ln_w, residual ? dy : nullptr, dx_q, g->ln_w, g->ln_b, acc, rows, e,
                              c.rest(), c.rest_bytes(), c.st));
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- self-attention layer
struct SaSaved {
    float *stats, *h0, *qkv, *o, *lse;
    size_t total;
    SaSaved(void* p, int batch, int n, int ch, int heads) {
        const size_t r = (size_t)batch * n;
        Carver k{(float*)p, 0};
        stats = k.take(r * 2);
        h0 = k.take(r * ch);
        qkv = k.take(3 * r * ch);
        o = k.take(r * ch);
        lse = k.take((size_t)batch * heads * n);
        total = k.off;
    }
    static size_t floats(int batch, int n, int ch, int heads) { return SaSaved(nullptr, batch, n, ch, heads).total; }
};

bool qkv_packed(const float* wq, const float* wk, const float* wv, const float* bq, const float* bk, const float* bv,
                int ch) {
    return wk == wq + (size_t)ch * ch && wv == wq + 2 * (size_t)ch * ch && bk == bq + ch && bv == bq + 2 * ch;
}

int self_attn_fwd(Ctx& c, const float* x, float* y, const pio_attn_weights* w, void* saved, int batch, int n, int ch,
                  int heads, int residual = 1) {
    SaSaved s(saved, batch, n, ch, heads);
    const int64_t rows = (int64_t)batch * n;
    PIO_TRY(pio_layernorm_fwd(x, w->ln_w, w->ln_b, s.h0, s.stats, rows, ch, LN_EPS, c.st));
    if (qkv_packed(w->wq, w->wk, w->wv, w->bq, w->bk, w->bv, ch)) {
        PIO_TRY(gemm_xwT(c, s.h0, ch, w->wq, ch, s.qkv, 3 * ch, rows, 3 * ch, ch, w->bq, nullptr, 0, 0, nullptr, 0));
    } else {
        const float* ws3[3] = {w->wq, w->wk, w->wv};
        const float* bs3[3] = {w->bq, w->bk, w->bv};
        for (int i = 0; i < 3; ++i)
            PIO_TRY(gemm_xwT(c, s.h0, ch, ws3[i], ch, s.qkv + i * ch, 3 * ch, rows, ch, ch, bs3[i], nullptr, 0, 0, nullptr, 0));
    }
    pio_attn_desc a{};
    fill_attn(a, s.qkv, 3 * ch, s.qkv + ch, s.qkv + 2 * ch, 3 * ch, nullptr, s.o, s.lse, batch, heads, n, n, ch);
    AttnBf16Out ex;
    ex.drop = c.attn_drop();
    PIO_TRY(attn_fwd_simt(&a, c.st, &ex));
    PIO_TRY(gemm_xwT(c, s.o, ch, w->wo, ch, y, ch, rows, ch, ch, w->bo, residual ? x : nullptr, ch, 0, nullptr, 0, 0,
                     residual ? c.drop_p : 0.f, c.drop_seed + 1));
    return PIO_OK;
}

size_t self_attn_bwd_ws(int batch, int n, int ch, int heads) {
    const size_t r = (size_t)batch * n;
    return 3 * align_up(r * ch * 4, 256) + align_up(3 * r * ch * 4, 256) + align_up((size_t)batch * heads * n * 4, 256) +
           pio_layernorm_bwd_ws_bytes(r, ch) + dW_scratch(3 * ch, ch, r) + 1024;
}

int self_attn_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_attn_weights* w, const void* saved,
                  const pio_attn_grads* g, int acc, int batch, int n, int ch, int heads, int residual = 1) {
    SaSaved s((void*)saved, batch, n, ch, heads);
    const int64_t rows = (int64_t)batch * n;
    const size_t mark = c.W.off;
    PIO_TAKE(d_o, c, (size_t)rows * ch, "pio_self_attn_bwd");
    PIO_TAKE(dqkv, c, (size_t)rows * 3 * ch, "pio_self_attn_bwd");
    PIO_TAKE(dh0, c, (size_t)rows * ch, "pio_self_attn_bwd");
    PIO_TAKE(delta, c, (size_t)batch * heads * n, "pio_self_attn_bwd");
    const float* dyb;
    PIO_TRY(branch_grad(c, dy, rows * ch, residual, &dyb, "pio_self_attn_bwd"));
    PIO_TRY(gemm_dyW(c, dyb, ch, w->wo, ch, d_o, ch, rows, ch, ch, 0, nullptr, 0, 0));
    PIO_TRY(gemm_dW(c, dyb, ch, s.o, ch, g->wo, g->bo, rows, ch, ch, acc));
    pio_attn_bwd_desc b{};
    fill_attn(b.fwd, s.qkv, 3 * ch, s.qkv + ch, s.qkv + 2 * ch, 3 * ch, nullptr, s.o, s.lse, batch, heads, n, n, ch);
    b.d_o = d_o; b.dq = dqkv; b.dk = dqkv + ch; b.dv = dqkv + 2 * ch; b.delta = delta;
    AttnBf16Out exb;
    exb.drop = c.attn_drop();
    PIO_TRY(attn_bwd_simt(&b, c.st, &exb));
    const bool wpk = qkv_packed(w->wq, w->wk, w->wv, w->bq, w->bk, w->bv, ch);
    const bool gpk = qkv_packed(g->wq, g->wk, g->wv, g->bq, g->bk, g->bv, ch);
    if (wpk) {
        PIO_TRY(gemm_dyW(c, dqkv, 3 * ch, w->wq, ch, dh0, ch, rows, 3 * ch, ch, 0, nullptr, 0, 0));
    } else {
        const float* ws3[3] = {w->wq, w->wk, w->wv};
        for (int i = 0; i < 3; ++i)
            PIO_TRY(gemm_dyW(c, dqkv + i * ch, 3 * ch, ws3[i], ch, dh0, ch, rows, ch, ch, 0, nullptr, 0, i > 0));
    }
    if (gpk) {
        PIO_TRY(gemm_dW(c, dqkv, 3 * ch, s.h0, ch, g->wq, g->bq, rows, 3 * ch, ch, acc));
    } else {
        float* gw3[3] = {g->wq, g->wk, g->wv};
        float* gb3[3] = {g->bq, g->bk, g->bv};
        for (int i = 0; i < 3; ++i)
            PIO_TRY(gemm_dW(c, dqkv + i * ch, 3 * ch, s.h0, ch, gw3[i], gb3[i], rows, ch, ch, acc));
    }
    PIO_TRY(pio_layernorm_bwd(dh0, x, s.stats, w->ln_w, residual ? dy : nullptr, dx, g->ln_w, g->ln_b, acc, rows, ch,
                              c.rest(), c.rest_bytes(), c.st));
    c.W.off = mark;
    return PIO_OK;
}

// ---------------------------------------------------------------- latent block layout
struct BlockLayout {
    size_t sa, xmid, mlp, xnext, per_layer;  // float offsets within a layer's slab
    BlockLayout(int batch, int n, int ch, int heads) {
        const size_t r = (size_t)batch * n;
        sa = 0;
        const size_t act = (r * ch + 63) / 64 * 64;
        xmid = sa + SaSaved::floats(batch, n, ch, heads);
        mlp = xmid + act;
        xnext = mlp + MlpSaved::floats(r, ch);
        per_layer = xnext + act;
    }
};

int check_mode(int mode, const char* who) {
    if (mode != PIO_MODE_FP32 && mode != PIO_MODE_BF16) {
        set_error("%s: unknown mode %d", who, mode);
        return PIO_ERR_UNSUPPORTED;
    }
    return PIO_OK;
}

}  // namespace
}  // namespace pio

using namespace pio;

extern "C" {

// ---- mlp
size_t pio_mlp_saved_bytes(int64_t rows, int c, int mode) {
    return mode == PIO_MODE_BF16 ? bf::mlp_saved_bytes(rows, c) : MlpSaved::floats(rows, c) * 4 + 256;
}
size_t pio_mlp_ws_bytes(int64_t rows, int c, int mode) { return mode == PIO_MODE_BF16 ? bf::mlp_ws_bytes(rows, c) : mlp_bwd_ws(rows, c); }

int pio_mlp_fwd(const float* x, float* y, const pio_mlp_weights* w, void* saved, int64_t rows, int c, int residual,
                const pio_dropout* drop, int mode, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_mlp_fwd"));
    PIO_CHECK_ARG(x && y && w && saved && rows >= 0 && c > 0, "pio_mlp_fwd: bad arguments");
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16) return bf::mlp_fwd(ctx, x, y, w, saved, rows, c, residual);
    return mlp_fwd(ctx, x, y, w, saved, rows, c, residual);
}

int pio_mlp_bwd(const float* dy, const float* x, float* dx, const pio_mlp_weights* w, const void* saved,
                const pio_mlp_grads* g, int accumulate, int64_t rows, int c, int residual, const pio_dropout* drop, int mode,
                void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_mlp_bwd"));
    PIO_CHECK_ARG(dy && x && dx && w && saved && g && rows >= 0 && c > 0, "pio_mlp_bwd: bad arguments");
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16) return bf::mlp_bwd(ctx, dy, x, dx, w, saved, g, accumulate, rows, c, residual);
    return mlp_bwd(ctx, dy, x, dx, w, saved, g, accumulate, rows, c, residual);
}

// ---- decoder on the labelled rows of a masked-LM batch (PIO_MODE_BF16)
size_t pio_label_compact_ws_bytes(int64_t rows) {
    return align_up((size_t)std::max<int64_t>(rows, 1) * 4, 256) + align_up((size_t)cdiv(std::max<int64_t>(rows, 1), 1024) * 4, 256) + 512;
}
int pio_label_compact(const int64_t* labels, int64_t rows, int64_t classes, int64_t ignore, int32_t* idx, int64_t* labels_c,
                      int32_t* count, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_CHECK_ARG(labels && idx && labels_c && count && rows >= 1 && rows < (1ll << 31), "pio_label_compact: bad arguments");
    Workspace W(ws, ws_bytes);
    int* pos = (int*)W.take((size_t)rows * 4);
    int* block_counts = (int*)W.take((size_t)cdiv(rows, 1024) * 4);
    float* out3 = W.takef(4);
    if (!pos || !block_counts || !out3) {
        set_error("pio_label_compact: workspace too small (%zu bytes given)", ws_bytes);
        return PIO_ERR_WORKSPACE;
    }
    return bf::label_compact(labels, rows, classes, ignore, idx, labels_c, count, pos, block_counts, out3, (cudaStream_t)stream);
}
int pio_decode_rows_supported(int n, int e, int heads) { return bf::decode_rows_supported(n, e, heads) ? 1 : 0; }
size_t pio_decode_rows_saved_bytes(int64_t max_rows, int e, int heads) { return bf::decode_rows_saved_bytes(max_rows, e, heads); }
size_t pio_decode_rows_ws_bytes(int64_t max_rows, int e, int heads) { return bf::decode_rows_ws_bytes(max_rows, e, heads); }

int pio_decode_rows_fwd(const float* queries, const void* kv, const int32_t* idx, const int32_t* count, float* y,
                        const pio_attn_weights* attn, const pio_mlp_weights* mlp, void* saved, int64_t max_rows, int l, int batch,
                        int n, int e, int heads, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_CHECK_ARG(queries && kv && idx && count && y && attn && mlp && saved && max_rows >= 1 && l >= 1 && batch >= 1,
                  "pio_decode_rows_fwd: bad arguments");
    PIO_CHECK_ARG(heads >= 1 && e % heads == 0, "pio_decode_rows_fwd: embed dim %d not divisible by %d heads", e, heads);
    Ctx ctx(ws, ws_bytes, stream);
    ctx.dyn = count;
    return bf::decode_rows_fwd(ctx, queries, kv, idx, y, attn, mlp, saved, max_rows, l, batch, n, e, heads);
}

int pio_decode_rows_bwd(const float* dy, const void* kv, const int32_t* idx, const int32_t* count, float* dqueries, void* dkv,
                        const pio_attn_weights* attn, const pio_mlp_weights* mlp, const void* saved, const pio_attn_grads* attn_g,
                        const pio_mlp_grads* mlp_g, int accumulate, int64_t max_rows, int l, int batch, int n, int e, int heads,
                        void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_CHECK_ARG(dy && kv && idx && count && dqueries && dkv && attn && mlp && saved && attn_g && mlp_g && max_rows >= 1,
                  "pio_decode_rows_bwd: bad arguments");
    PIO_CHECK_ARG(heads >= 1 && e % heads == 0, "pio_decode_rows_bwd: embed dim %d not divisible by %d heads", e, heads);
    Ctx ctx(ws, ws_bytes, stream);
    ctx.dyn = count;
    return bf::decode_rows_bwd(ctx, dy, kv, idx, dqueries, dkv, attn, mlp, saved, attn_g, mlp_g, accumulate, max_rows, l, batch, n, e,
                               heads);
}

// ---- kv projection
size_t pio_kv_project_saved_bytes(int64_t rows, int c_kv, int e, int mode) {
    (void)e;
    return mode == PIO_MODE_BF16 ? bf::kv_saved_bytes(rows, c_kv) : KvSaved::floats(rows, c_kv) * 4 + 256;
}
size_t pio_kv_project_ws_bytes(int64_t rows, int c_kv, int e, int mode) {
    return mode == PIO_MODE_BF16 ? bf::kv_ws_bytes(rows, c_kv, e) : kv_project_bwd_ws(rows, c_kv, e);
}

int pio_kv_project_fwd(const float* x_kv, float* kv, const pio_attn_weights* w, void* saved, int64_t rows, int c_kv, int e,
                       int mode, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_kv_project_fwd"));
    PIO_CHECK_ARG(x_kv && kv && w && saved && rows >= 0 && c_kv > 0 && e > 0, "pio_kv_project_fwd: bad arguments");
    PIO_CHECK_ARG(w->kv_ln_w && w->kv_ln_b && w->wk && w->wv && w->bk && w->bv, "pio_kv_project_fwd: missing weights");
    Ctx ctx(ws, ws_bytes, stream);
    if (mode == PIO_MODE_BF16) return bf::kv_project_fwd(ctx, x_kv, kv, w, saved, rows, c_kv, e);
    return kv_project_fwd(ctx, x_kv, kv, w, saved, rows, c_kv, e);
}

int pio_kv_project_bwd(const float* dkv, const float* x_kv, float* dx_kv, const pio_attn_weights* w, const void* saved,
                       const pio_attn_grads* g, int accumulate, int64_t rows, int c_kv, int e, int mode, void* ws,
                       size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_kv_project_bwd"));
    PIO_CHECK_ARG(dkv && x_kv && w && saved && g && rows >= 0 && c_kv > 0 && e > 0, "pio_kv_project_bwd: bad arguments");
    PIO_CHECK_ARG(g->kv_ln_w && g->kv_ln_b && g->wk && g->wv && g->bk && g->bv, "pio_kv_project_bwd: missing grads");
    PIO_CHECK_ARG((accumulate & ~3) == 0 || (mode == PIO_MODE_BF16 && accumulate == (2 | PIO_ACC_DX) && dx_kv),
                  "pio_kv_project_bwd: PIO_ACC_DX goes with PIO_MODE_BF16, accumulate = 2 and a dx_kv buffer");
    Ctx ctx(ws, ws_bytes, stream);
    if (mode == PIO_MODE_BF16) return bf::kv_project_bwd(ctx, dkv, x_kv, dx_kv, w, saved, g, accumulate, rows, c_kv, e);
    return kv_project_bwd(ctx, dkv, x_kv, dx_kv, w, saved, g, accumulate, rows, c_kv, e);
}

// ---- query side of cross attention
size_t pio_cross_attend_saved_bytes(int batch, int n, int m, int e, int heads, int mode) {
    (void)m;
    return mode == PIO_MODE_BF16 ? bf::xa_saved_bytes(batch, n, e, heads) : XaSaved::floats(batch, n, e, heads) * 4 + 256;
}
size_t pio_cross_attend_ws_bytes(int batch, int n, int m, int e, int heads, int mode) {
    (void)m;
    return mode == PIO_MODE_BF16 ? bf::xa_ws_bytes(batch, n, e, heads) : cross_attend_bwd_ws(batch, n, e, heads);
}

static int check_heads(int e, int heads, const char* who) {
    PIO_CHECK_ARG(heads > 0 && e > 0 && e % heads == 0, "%s: embed dim %d not divisible by heads %d", who, e, heads);
    return PIO_OK;
}

int pio_cross_attend_fwd(const float* x_q, const float* kv, const uint8_t* pad, float* y, const pio_attn_weights* w,
                         void* saved, int batch, int n, int m, int e, int heads, int residual, const pio_dropout* drop,
                         int mode, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_cross_attend_fwd"));
    PIO_CHECK_ARG(x_q && kv && (y || residual == 2) && w && saved && batch >= 0 && n >= 0 && m >= 0, "pio_cross_attend_fwd: bad arguments");
    PIO_CHECK_ARG(residual != 2 || mode == PIO_MODE_BF16, "pio_cross_attend_fwd: the deferred out-projection (residual = 2) needs PIO_MODE_BF16");
    PIO_TRY(check_heads(e, heads, "pio_cross_attend_fwd"));
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16) return bf::cross_attend_fwd(ctx, x_q, kv, pad, y, w, saved, batch, n, m, e, heads, residual);
    return cross_attend_fwd(ctx, x_q, kv, pad, y, w, saved, batch, n, m, e, heads, residual);
}

int pio_cross_attend_bwd(const float* dy, const float* x_q, const float* kv, const uint8_t* pad, float* dx_q, float* dkv,
                         const pio_attn_weights* w, const void* saved, const pio_attn_grads* g, int accumulate, int batch,
                         int n, int m, int e, int heads, int residual, const pio_dropout* drop, int mode, void* ws,
                         size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_cross_attend_bwd"));
    PIO_CHECK_ARG(dy && x_q && kv && dx_q && dkv && w && saved && g, "pio_cross_attend_bwd: bad arguments");
    PIO_CHECK_ARG(residual != 2 || mode == PIO_MODE_BF16, "pio_cross_attend_bwd: the deferred out-projection (residual = 2) needs PIO_MODE_BF16");
    PIO_CHECK_ARG((accumulate & ~3) == 0 || (mode == PIO_MODE_BF16 && (accumulate & ~(PIO_ACC_DX | PIO_ACC_DKV)) == 2),
                  "pio_cross_attend_bwd: PIO_ACC_DX / PIO_ACC_DKV go with PIO_MODE_BF16 and accumulate = 2");
    PIO_CHECK_ARG(!(accumulate & PIO_ACC_DX) || residual == 2, "pio_cross_attend_bwd: PIO_ACC_DX needs the deferred out-projection (residual = 2)");
    PIO_TRY(check_heads(e, heads, "pio_cross_attend_bwd"));
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16)
        return bf::cross_attend_bwd(ctx, dy, x_q, kv, pad, dx_q, dkv, w, saved, g, accumulate, batch, n, m, e, heads, residual);
    return cross_attend_bwd(ctx, dy, x_q, kv, pad, dx_q, dkv, w, saved, g, accumulate, batch, n, m, e, heads, residual);
}

// ---- one self-attention layer
size_t pio_self_attn_saved_bytes(int batch, int n, int c, int heads, int mode) {
    return mode == PIO_MODE_BF16 ? bf::sa_saved_bytes(batch, n, c, heads) : SaSaved::floats(batch, n, c, heads) * 4 + 256;
}
size_t pio_self_attn_ws_bytes(int batch, int n, int c, int heads, int mode) {
    return mode == PIO_MODE_BF16 ? bf::sa_ws_bytes(batch, n, c, heads) : self_attn_bwd_ws(batch, n, c, heads);
}

int pio_self_attn_fwd(const float* x, float* y, const pio_attn_weights* w, void* saved, int batch, int n, int c, int heads,
                      int residual, const pio_dropout* drop, int mode, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_self_attn_fwd"));
    PIO_CHECK_ARG(x && y && w && saved && batch >= 0 && n >= 0, "pio_self_attn_fwd: bad arguments");
    PIO_TRY(check_heads(c, heads, "pio_self_attn_fwd"));
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16) return bf::self_attn_fwd(ctx, x, y, w, saved, batch, n, c, heads, residual);
    return self_attn_fwd(ctx, x, y, w, saved, batch, n, c, heads, residual);
}

int pio_self_attn_bwd(const float* dy, const float* x, float* dx, const pio_attn_weights* w, const void* saved,
                      const pio_attn_grads* g, int accumulate, int batch, int n, int c, int heads, int residual,
                      const pio_dropout* drop, int mode, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_self_attn_bwd"));
    PIO_CHECK_ARG(dy && x && dx && w && saved && g, "pio_self_attn_bwd: bad arguments");
    PIO_TRY(check_heads(c, heads, "pio_self_attn_bwd"));
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16) return bf::self_attn_bwd(ctx, dy, x, dx, w, saved, g, accumulate, batch, n, c, heads, residual);
    return self_attn_bwd(ctx, dy, x, dx, w, saved, g, accumulate, batch, n, c, heads, residual);
}

// ---- latent block
size_t pio_cross_attend_o16_offset(int batch, int n, int e, int heads) { return bf::xa_o16_offset(batch, n, e, heads); }
int pio_cross_attend_merges_dkv(int batch, int n, int m, int e, int heads) { return bf::xa_merges_dkv(batch, n, m, e, heads) ? 1 : 0; }

int pio_latent_block_cluster_ok(int layers, int batch, int n, int c, int heads) {
    const int64_t rows = (int64_t)batch * n;
    return layers >= 1 && layers <= PIO_LATENT_FUSED_MAX_LAYERS && latent_fused_supported(rows, n, c, heads, 0.f) && latent_cluster_preferred(rows) &&
                   latent_bwd_cluster_supported(rows, n, c, heads, 0.f)
               ? 1
               : 0;
}

int pio_set_side_join_deferred(int on) {
    bf::set_side_join_deferred(on);
    return PIO_OK;
}

int pio_side_join(pio_stream_t stream) { return bf::side_join((cudaStream_t)stream); }

int pio_logits_bf16_supported(int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldw, int64_t ldc) {
    return logits_persistent_supported(m, n, k, lda, ldw, ldc) ? 1 : 0;
}

int pio_logits_bf16(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, void* c, int64_t ldc, int64_t m,
                    int64_t n, int64_t k, const int32_t* dyn_rows, int dyn_off, pio_stream_t stream) {
    return logits_persistent(a, lda, w, ldw, bias, c, ldc, m, n, k, dyn_rows, dyn_off, (cudaStream_t)stream);
}

int pio_dw_batched(const pio_dw_item* items, int count, int64_t rows, pio_stream_t stream) {
    static_assert(sizeof(pio_dw_item) == sizeof(DwBatchItem), "pio_dw_item and DwBatchItem must have the same layout");
    return dw_batched(reinterpret_cast<const DwBatchItem*>(items), count, rows, (cudaStream_t)stream);
}

int pio_ln_param_grads(const pio_ln_grad_item* items, int count, int64_t rows, int c, pio_stream_t stream) {
    static_assert(sizeof(pio_ln_grad_item) == sizeof(LnGradItem), "pio_ln_grad_item and LnGradItem must have the same layout");
    return ln_param_grads(reinterpret_cast<const LnGradItem*>(items), count, rows, c, (cudaStream_t)stream);
}

size_t pio_latent_block_saved_bytes(int layers, int batch, int n, int c, int heads, int mode) {
    if (mode == PIO_MODE_BF16) return bf::block_saved_bytes(layers, batch, n, c, heads);
    BlockLayout L(batch, n, c, heads);
    return (size_t)layers * L.per_layer * 4 + 256;
}
size_t pio_latent_block_ws_bytes(int layers, int batch, int n, int c, int heads, int mode) {
    (void)layers;
    if (mode == PIO_MODE_BF16) return bf::block_ws_bytes(layers, batch, n, c, heads);
    const size_t r = (size_t)batch * n;
    return 2 * align_up(r * c * 4, 256) + std::max(self_attn_bwd_ws(batch, n, c, heads), mlp_bwd_ws(r, c)) + 1024;
}

int pio_latent_block_fwd(const float* x, float* y, const pio_attn_weights* attn, const pio_mlp_weights* mlp, void* saved,
                         int layers, int batch, int n, int c, int heads, const pio_dropout* drop, int mode, void* ws,
                         size_t ws_bytes, pio_stream_t stream) {
    return pio_latent_block_fwd_pre(x, y, attn, mlp, saved, layers, batch, n, c, heads, drop, mode, ws, ws_bytes, stream, nullptr);
}

int pio_latent_block_fwd_pre(const float* x, float* y, const pio_attn_weights* attn, const pio_mlp_weights* mlp, void* saved,
                             int layers, int batch, int n, int c, int heads, const pio_dropout* drop, int mode, void* ws,
                             size_t ws_bytes, pio_stream_t stream, const void* pre_o16) {
    PIO_TRY(check_mode(mode, "pio_latent_block_fwd"));
    PIO_CHECK_ARG(x && y && attn && mlp && saved && layers >= 1 && batch >= 0 && n >= 0, "pio_latent_block_fwd: bad arguments");
    PIO_TRY(check_heads(c, heads, "pio_latent_block_fwd"));
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16) return bf::latent_block_fwd(ctx, x, y, attn, mlp, saved, layers, batch, n, c, heads, pre_o16);
    PIO_CHECK_ARG(attn[0].wq != nullptr, "pio_latent_block_fwd: a first layer without attention weights needs PIO_MODE_BF16");
    BlockLayout L(batch, n, c, heads);
    const int64_t rows = (int64_t)batch * n;
    const float* cur = x;
    for (int s = 0; s < layers; ++s) {
        float* slab = (float*)saved + (size_t)s * L.per_layer;
        float* xmid = slab + L.xmid;
        float* out = (s == layers - 1) ? y : slab + L.xnext;
        const uint64_t seed0 = ctx.drop_seed;
        ctx.drop_seed = seed0 + 4 * (uint64_t)s;
        PIO_TRY(self_attn_fwd(ctx, cur, xmid, &attn[s], slab + L.sa, batch, n, c, heads));
        ctx.drop_seed = seed0 + 4 * (uint64_t)s + 2;
        PIO_TRY(mlp_fwd(ctx, xmid, out, &mlp[s], slab + L.mlp, rows, c));
        ctx.drop_seed = seed0;
        cur = out;
    }
    return PIO_OK;
}

int pio_latent_block_bwd(const float* dy, const float* x, float* dx, const pio_attn_weights* attn, const pio_mlp_weights* mlp,
                         const void* saved, const pio_attn_grads* attn_g, const pio_mlp_grads* mlp_g, int accumulate,
                         int layers, int batch, int n, int c, int heads, const pio_dropout* drop, int mode, void* ws,
                         size_t ws_bytes, pio_stream_t stream) {
    return pio_latent_block_bwd_pre(dy, x, dx, attn, mlp, saved, attn_g, mlp_g, accumulate, layers, batch, n, c, heads, drop, mode, ws,
                                    ws_bytes, stream, nullptr, nullptr);
}

int pio_latent_block_bwd_pre(const float* dy, const float* x, float* dx, const pio_attn_weights* attn, const pio_mlp_weights* mlp,
                             const void* saved, const pio_attn_grads* attn_g, const pio_mlp_grads* mlp_g, int accumulate,
                             int layers, int batch, int n, int c, int heads, const pio_dropout* drop, int mode, void* ws,
                             size_t ws_bytes, pio_stream_t stream, const void* pre_o16, void* pre_do16) {
    PIO_TRY(check_mode(mode, "pio_latent_block_bwd"));
    PIO_CHECK_ARG(dy && x && dx && attn && mlp && saved && attn_g && mlp_g && layers >= 1, "pio_latent_block_bwd: bad arguments");
    PIO_TRY(check_heads(c, heads, "pio_latent_block_bwd"));
    Ctx ctx(ws, ws_bytes, stream, drop);
    if (mode == PIO_MODE_BF16)
        return bf::latent_block_bwd(ctx, dy, x, dx, attn, mlp, saved, attn_g, mlp_g, accumulate, layers, batch, n, c, heads, pre_o16,
                                    pre_do16);
    PIO_CHECK_ARG(attn[0].wq != nullptr, "pio_latent_block_bwd: a first layer without attention weights needs PIO_MODE_BF16");
    BlockLayout L(batch, n, c, heads);
    const int64_t rows = (int64_t)batch * n;
    PIO_TAKE(dmid, ctx, (size_t)rows * c, "pio_latent_block_bwd");
    PIO_TAKE(dcur, ctx, (size_t)rows * c, "pio_latent_block_bwd");
    const float* grad = dy;
    for (int s = layers - 1; s >= 0; --s) {
        const float* slab = (const float*)saved + (size_t)s * L.per_layer;
        const float* x_in = (s == 0) ? x : ((const float*)saved + (size_t)(s - 1) * L.per_layer + L.xnext);
        const uint64_t seed0 = ctx.drop_seed;
        ctx.drop_seed = seed0 + 4 * (uint64_t)s + 2;
        PIO_TRY(mlp_bwd(ctx, grad, slab + L.xmid, dmid, &mlp[s], slab + L.mlp, &mlp_g[s], accumulate, rows, c));
        float* out = (s == 0) ? dx : dcur;
        ctx.drop_seed = seed0 + 4 * (uint64_t)s;
        PIO_TRY(self_attn_bwd(ctx, dmid, x_in, out, &attn[s], slab + L.sa, &attn_g[s], accumulate, batch, n, c, heads));
        ctx.drop_seed = seed0;
        grad = out;
    }
    return PIO_OK;
}

// ---- output linear
size_t pio_linear_ws_bytes(int64_t rows, int c_in, int64_t c_out, int mode) {
    if (mode == PIO_MODE_BF16) return bf::linear_ws_bytes(rows, c_in, c_out);
    return dW_scratch(c_out, c_in, rows) + 1024;
}

int pio_linear_fwd(const float* x, const float* w, const float* b, float* y, int64_t rows, int c_in, int64_t c_out, int mode,
                   void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_linear_fwd"));
    PIO_CHECK_ARG(x && w && y && rows >= 0 && c_in > 0 && c_out > 0, "pio_linear_fwd: bad arguments");
    Ctx ctx(ws, ws_bytes, stream);
    if (mode == PIO_MODE_BF16) return bf::linear_fwd(ctx, x, w, b, y, rows, c_in, c_out);
    return gemm_xwT(ctx, x, c_in, w, c_in, y, c_out, rows, c_out, c_in, b, nullptr, 0, 0, nullptr, 0);
}

int pio_linear_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, float* db, int accumulate,
                   int64_t rows, int c_in, int64_t c_out, int mode, void* ws, size_t ws_bytes, pio_stream_t stream) {
    PIO_TRY(check_mode(mode, "pio_linear_bwd"));
    PIO_CHECK_ARG(dy && x && w && dw && rows >= 0 && c_in > 0 && c_out > 0, "pio_linear_bwd: bad arguments");
    Ctx ctx(ws, ws_bytes, stream);
    if (mode == PIO_MODE_BF16) return bf::linear_bwd(ctx, dy, x, w, dx, dw, db, accumulate, rows, c_in, c_out);
    if (dx) PIO_TRY(gemm_dyW(ctx, dy, c_out, w, c_in, dx, c_in, rows, c_out, c_in, 0, nullptr, 0, 0));
    return gemm_dW(ctx, dy, c_out, x, c_in, dw, db, rows, c_out, c_in, accumulate);
}

// ---- output linear fused with cross-entropy (training)
size_t pio_linear_ce_ws_bytes(int64_t rows, int c_in, int64_t classes, int mode) {
    (void)mode;
    return bf::linear_ce_ws_bytes(rows, c_in, classes);
}

int pio_linear_ce_fwd_bwd(const float* x, const float* w, const float* b, const int64_t* labels, float* out3, float* dx, float* dw,
                          float* db, int64_t rows, int c_in, int64_t classes, int64_t ignore_index, int mode, void* ws,
                          size_t ws_bytes, pio_stream_t stream) {
    if (mode != PIO_MODE_BF16) {
        set_error("pio_linear_ce_fwd_bwd: only PIO_MODE_BF16 has the fused path (compose pio_linear_* and pio_cross_entropy_* in fp32)");
        return PIO_ERR_UNSUPPORTED;
    }
    PIO_CHECK_ARG(x && w && b && labels && out3 && dx && dw && db && rows > 0 && c_in > 0 && classes > 0,
                  "pio_linear_ce_fwd_bwd: bad arguments");
    Ctx ctx(ws, ws_bytes, stream);
    return bf::linear_ce_fwd_bwd(ctx, x, w, b, labels, out3, dx, dw, db, rows, c_in, classes, ignore_index);
}

}  // extern "C"
