// Shared plumbing of the fused operators (fp32 and bf16 pipelines).
#pragma once
#include "pio_common.cuh"

namespace pio {

// optional bf16 copies written by the fp32 attention kernels (same element strides as the fp32 tensors)
struct AttnBf16Out {
    __nv_bfloat16* o16 = nullptr;
    __nv_bfloat16* dq16 = nullptr;
    __nv_bfloat16* dk16 = nullptr;
    __nv_bfloat16* dv16 = nullptr;
    int in16 = 0;  // q, k, v of the descriptor point at bf16 data (same element strides)
    DropCfg drop = DropCfg{0, 0, 1.f};  // attention-probability dropout (fp32-arithmetic kernels only)
};
int attn_fwd_simt(const pio_attn_desc* d, cudaStream_t st, const AttnBf16Out* extra = nullptr);
int attn_bwd_simt(const pio_attn_bwd_desc* d, cudaStream_t st, const AttnBf16Out* extra = nullptr);

// tcgen05 flash attention forward (pio_attn_tc.cu): q, k, v of the descriptor are bf16
bool attn_fwd_tc_supported(const pio_attn_desc* d);
int attn_fwd_tc(const pio_attn_desc* d, __nv_bfloat16* o16, cudaStream_t st, const DropCfg* drop = nullptr);
// tcgen05 backward: dO bf16 (strides of o), gradients bf16 with the strides of q / k / v
// (o16 given: the row term delta = dO . O is formed from the bf16 copy of the output instead of d->o;
//  dkv_add: dK and dV are added to what dk16 / dv16 hold — fp32 sum, one rounding)
int attn_bwd_tc(const pio_attn_desc* d, const __nv_bfloat16* do16, float* delta, __nv_bfloat16* dq16, __nv_bfloat16* dk16,
                __nv_bfloat16* dv16, cudaStream_t st, const DropCfg* drop = nullptr, const __nv_bfloat16* o16 = nullptr,
                bool dkv_add = false);

// every query of a sample in one 128-row tile (the latent array against the input), head size 64, no attention dropout:
// one kernel, CTA = (sample, head), nothing recomputed (pio_xattn_tc.cu); attn_bwd_tc dispatches to it
bool xattn_bwd_tc_supported(const pio_attn_desc* d, const DropCfg* drop);
int xattn_bwd_tc(const pio_attn_desc* d, const __nv_bfloat16* do16, __nv_bfloat16* dq16, __nv_bfloat16* dk16, __nv_bfloat16* dv16,
                 cudaStream_t st, const __nv_bfloat16* o16, bool dkv_add);

int gemm_bf16(const pio_gemm_bf16_desc* d, void* ws, size_t ws_bytes, cudaStream_t st);
int cast_bf16(const float* src, void* dst, int64_t rows, int cols, int64_t ld_src, int64_t ld_dst, cudaStream_t st);

// bf16 helpers (pio_bf16_aux.cu)
// (dyn_rows: optional DEVICE-side row count <= rows; rows between it and the next multiple of 128 of every bf16 output
//  are zero-filled so they can serve as operand padding of a product over the row axis)
int layernorm_fwd_bf16(const float* x, const float* w, const float* b, __nv_bfloat16* y, int64_t ldy, float* stats,
                       int64_t rows, int c, float eps, cudaStream_t st, const int* dyn_rows = nullptr);
int layernorm_bwd(const float* dy, const float* x, const float* stats, const float* w, const float* dres, float* dx, float* dw,
                  float* db, int accumulate, int64_t rows, int c, void* ws, size_t ws_bytes, cudaStream_t st,
                  const int* dyn_rows = nullptr, __nv_bfloat16* dx16 = nullptr);  // dx16: optional bf16 copy of dx
// dst16 (optional) = bf16(src); colsum (optional, c floats) (+)= column sums of src.  scratch: colsum_ws_bytes
int cast_colsum_f32(const float* src, int64_t ld_src, __nv_bfloat16* dst16, int64_t ld_dst, float* colsum, int accumulate,
                    int64_t rows, int c, void* ws, size_t ws_bytes, cudaStream_t st, const int* dyn_rows = nullptr);
int colsum_bf16(const __nv_bfloat16* src, int64_t ld_src, float* colsum, int accumulate, int64_t rows, int c, void* ws,
                size_t ws_bytes, cudaStream_t st, const int* dyn_rows = nullptr);
size_t colsum_ws_bytes(int64_t rows, int c);

constexpr float LN_EPS = 1e-5f;  // nn.LayerNorm default, used everywhere in the reference

struct Ctx {
    cudaStream_t st;
    Workspace W;
    // dropout of the operator being run (0 = off): site seed+0 = attention probabilities, seed+1 = Residual
    float drop_p = 0.f;
    uint64_t drop_seed = 0;
    // optional DEVICE-side row count of the operator's activations (<= the row count passed by the host): every kernel is
    // launched for the host's worst case and leaves / clamps at the device count (the labelled-row decoder of the MLM step)
    const int* dyn = nullptr;
    Ctx(void* ws, size_t n, pio_stream_t s) : st((cudaStream_t)s), W(ws, n) {}
    Ctx(void* ws, size_t n, pio_stream_t s, const pio_dropout* d) : st((cudaStream_t)s), W(ws, n) {
        if (d != nullptr && d->p > 0.f) {
            drop_p = d->p;
            drop_seed = d->seed;
        }
    }
    DropCfg attn_drop() const { return make_drop(drop_p, drop_seed); }
    DropCfg res_drop(int residual) const { return make_drop(residual ? drop_p : 0.f, drop_seed + 1); }
    void* rest() { return W.base ? W.base + align_up(W.off, 256) : nullptr; }
    size_t rest_bytes() {
        const size_t o = align_up(W.off, 256);
        return (W.base && o < W.size) ? W.size - o : 0;
    }
};

#define PIO_TAKE(var, ctx, count, who)                                                     \
    float* var = (ctx).W.takef((size_t)(count));                                           \
    if (!var) {                                                                            \
        set_error("%s: workspace too small (%zu bytes given)", who, (ctx).W.size);        \
        return PIO_ERR_WORKSPACE;                                                          \
    }

// carve 256-byte aligned arrays out of an opaque saved buffer (nullptr base = size query)
struct Carver {
    float* base;
    size_t off;
    float* take(size_t n) {
        float* r = base ? base + off : nullptr;
        off += (n + 63) / 64 * 64;
        return r;
    }
};


static inline int64_t pad8(int64_t x) { return (x + 7) / 8 * 8; }

// Fused latent-block forward (pio_latent_tc.cu): one launch for up to PIO_LATENT_FUSED_MAX_LAYERS layers of a
// self-attention block.  Per layer: bf16 weight shadows, fp32 LayerNorm parameters / biases, and where each piece
// of saved-for-backward state goes (the layout of the unfused bf16 operators, so their backward runs on it).
#define PIO_LATENT_FUSED_MAX_LAYERS 8
struct LatentFusedLayer {
    const void *wqkv, *wo, *w1, *w2;                                    // bf16, row pitch C
    const float *ln1_w, *ln1_b, *bqkv, *bo, *ln2_w, *ln2_b, *b1, *b2;  // fp32
    float* sa_stats; void* sa_h0; void* qkv; float* o32; void* o16; float* lse;
    float* xmid;
    float* mlp_stats; void* mlp_h0; void* mlp_h1; void* mlp_g;
    float* xout;
    const void* pre_o16;  // layer 0 as an out-projection + MLP first layer: the attention output (rows x C, bf16)
};
// one layer of the fused backward input-gradient chain (pio_latent_bwd_cl.cu): weights (bf16 shadows), saved forward
// state, and the bf16 gradient tensors the kernel leaves for the batched weight-gradient kernel (all [rows][C] except
// gqkv16 [rows][3C]; equally spaced over the layers, like the saved slabs)
struct LatentBwdLayer {
    const void *w2, *w1, *wo, *wqkv;
    const float *ln2_w, *ln1_w;
    const float *mlp_stats, *sa_stats, *lse;
    const void* mlp_h1;
    const float* xmid;
    const float* xnext;  // the layer's output (the LayerNorm-1 input of the NEXT layer)
    const void* qkv;
    void *gy16, *gh16, *gmid16, *gqkv16;
    void* pre_do16;  // layer 0 as an out-projection + MLP first layer: the gradient of the attention output (rows x C, bf16), written
    float *dln2_w, *dln2_b, *dln1_w, *dln1_b;  // LayerNorm parameter gradients: added into by the chain kernel itself
};
bool latent_bwd_cluster_supported(int64_t rows, int n, int ch, int heads, float drop_p);
int latent_block_bwd_cluster(const float* dy, const float* x, float* dx, const LatentBwdLayer* layers, int nlayers, int batch, int n,
                             cudaStream_t st);
#define PIO_LN_GRAD_MAX_ITEMS 16
struct LnGradItem {
    const void* dn16;
    const float* x;
    const float* stats;
    float* dgamma;
    float* dbeta;
};
int ln_param_grads(const LnGradItem* items, int count, int64_t rows, int c, cudaStream_t st);
// one launch for a table of weight-gradient products over the same rows (pio_dw_batched.cu)
struct DwBatchItem {
    const void* dy16;
    int64_t lddy;
    int m;
    const void* x16;
    int64_t ldx;
    int n;  // 256
    float* dw;
    float* db;
};
int dw_batched(const DwBatchItem* items, int count, int64_t rows, cudaStream_t st, const int* dyn_rows = nullptr, int dyn_off = 0);
bool xw_splitk_supported(int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldb);
bool xw_splitk_stores(int64_t m, int64_t k);
int xw_splitk(const void* a16, int64_t lda, const void* b16, int64_t ldb, float* c32, int64_t m, int64_t k, cudaStream_t st,
              const int* dyn_rows = nullptr, int dyn_off = 0);
// persistent vocabulary-logits product (pio_logits_tc.cu)
bool logits_persistent_supported(int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldb, int64_t ldc);
int logits_persistent(const void* a, int64_t lda, const void* w, int64_t ldb, const float* bias, void* c, int64_t ldc, int64_t m,
                      int64_t n, int64_t k, const int* dyn_rows, int dyn_off, cudaStream_t st);
bool dr_attn_bwd_tc_supported(int n, int e, int heads);
int dr_attn_bwd_tc(const void* q16, const void* kv, const void* do16, const float* lse, const int* seg, const int* count, void* dq16,
                   void* dkv, int batch, int n, int e, int heads, int64_t max_rows, cudaStream_t st);
int dr_attn_fwd_tc(const void* q16, const void* kv, const int* seg, const int* count, void* o16, float* lse, int batch, int n, int e,
                   int heads, int64_t max_rows, cudaStream_t st);
bool ln_linear_supported(int64_t m, int64_t n, int64_t k, int64_t ldw, int64_t ldc, int64_t ldh);
int ln_linear(const float* x, const float* gamma, const float* beta, float eps, const void* w, int64_t ldw, const float* bias, void* c,
              int64_t ldc, void* h0, int64_t ldh, float* stats, int64_t m, int64_t n, cudaStream_t st);
bool latent_cluster_preferred(int64_t rows);
bool latent_fused_supported(int64_t rows, int n, int ch, int heads, float drop_p);
int latent_block_fwd_fused(const float* x, const LatentFusedLayer* layers, int nlayers, int batch, int n, cudaStream_t st);

// PIO_MODE_BF16 pipelines (pio_ops_bf16.cu)
namespace bf {
size_t mlp_saved_bytes(int64_t rows, int ch);
size_t mlp_ws_bytes(int64_t rows, int ch);
int mlp_fwd(Ctx& c, const float* x, float* y, const pio_mlp_weights* w, void* saved, int64_t rows, int ch, int residual);
// dy16_in: dy already as bf16 (skips the cast; the bias-gradient column sums then run on the side stream);
// dx16_out: also write dx as bf16 (for the next operator's dy16_in)
int mlp_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_mlp_weights* w, const void* saved,
            const pio_mlp_grads* g, int acc, int64_t rows, int ch, int residual, const __nv_bfloat16* dy16_in = nullptr,
            __nv_bfloat16* dx16_out = nullptr);
size_t kv_saved_bytes(int64_t rows, int ckv);
size_t kv_ws_bytes(int64_t rows, int ckv, int e);
int kv_project_fwd(Ctx& c, const float* x_kv, float* kv, const pio_attn_weights* w, void* saved, int64_t rows, int ckv, int e);
int kv_project_bwd(Ctx& c, const float* dkv, const float* x_kv, float* dx_kv, const pio_attn_weights* w, const void* saved,
                   const pio_attn_grads* g, int acc, int64_t rows, int ckv, int e);
size_t xa_saved_bytes(int batch, int n, int e, int heads);
size_t xa_ws_bytes(int batch, int n, int e, int heads);
int cross_attend_fwd(Ctx& c, const float* x_q, const float* kv, const uint8_t* pad, float* y, const pio_attn_weights* w,
                     void* saved, int batch, int n, int m, int e, int heads, int residual);
int cross_attend_bwd(Ctx& c, const float* dy, const float* x_q, const float* kv, const uint8_t* pad, float* dx_q, float* dkv,
                     const pio_attn_weights* w, const void* saved, const pio_attn_grads* g, int acc, int batch, int n, int m,
                     int e, int heads, int residual);
size_t sa_saved_bytes(int batch, int n, int ch, int heads);
size_t sa_ws_bytes(int batch, int n, int ch, int heads);
int self_attn_fwd(Ctx& c, const float* x, float* y, const pio_attn_weights* w, void* saved, int batch, int n, int ch, int heads,
                  int residual);
int self_attn_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_attn_weights* w, const void* saved,
                  const pio_attn_grads* g, int acc, int batch, int n, int ch, int heads, int residual,
                  const __nv_bfloat16* dy16_in = nullptr, __nv_bfloat16* dx16_out = nullptr);
size_t xa_o16_offset(int batch, int n, int e, int heads);
bool xa_merges_dkv(int batch, int n, int m, int e, int heads);
void set_side_join_deferred(int on);
int side_join(cudaStream_t st);
size_t block_saved_bytes(int layers, int batch, int n, int ch, int heads);
size_t block_ws_bytes(int layers, int batch, int n, int ch, int heads);
int latent_block_fwd(Ctx& c, const float* x, float* y, const pio_attn_weights* attn, const pio_mlp_weights* mlp, void* saved,
                     int layers, int batch, int n, int ch, int heads, const void* pre_o16 = nullptr);
int latent_block_bwd(Ctx& c, const float* dy, const float* x, float* dx, const pio_attn_weights* attn, const pio_mlp_weights* mlp,
                     const void* saved, const pio_attn_grads* attn_g, const pio_mlp_grads* mlp_g, int acc, int layers,
                     int batch, int n, int ch, int heads, const void* pre_o16 = nullptr, void* pre_do16 = nullptr);
size_t linear_ws_bytes(int64_t rows, int c_in, int64_t c_out);
int linear_fwd(Ctx& c, const float* x, const float* w16, const float* b, float* y, int64_t rows, int c_in, int64_t c_out);
int linear_bwd(Ctx& c, const float* dy, const float* x, const float* w16, float* dx, float* dw, float* db, int acc, int64_t rows,
               int c_in, int64_t c_out);
size_t decode_rows_saved_bytes(int64_t max_rows, int e, int heads);
size_t decode_rows_ws_bytes(int64_t max_rows, int e, int heads);
bool decode_rows_supported(int n, int e, int heads);
int label_compact(const int64_t* labels, int64_t rows, int64_t classes, int64_t ignore, int* idx, int64_t* labels_c, int* count,
                  int* pos_scratch, int* block_counts, float* out3_scratch, cudaStream_t st);
int decode_rows_fwd(Ctx& c, const float* queries, const void* kv, const int* idx, float* y, const pio_attn_weights* aw,
                    const pio_mlp_weights* mw, void* saved, int64_t max_rows, int l, int batch, int n, int e, int heads);
int decode_rows_bwd(Ctx& c, const float* dy, const void* kv, const int* idx, float* dqueries, void* dkv, const pio_attn_weights* aw,
                    const pio_mlp_weights* mw, const void* saved, const pio_attn_grads* ag, const pio_mlp_grads* mg, int acc,
                    int64_t max_rows, int l, int batch, int n, int e, int heads);
size_t linear_ce_ws_bytes(int64_t rows, int c_in, int64_t classes);
int linear_ce_fwd_bwd(Ctx& c, const float* x, const float* w16, const float* b, const int64_t* labels, float* out3, float* dx,
                      float* dw, float* db, int64_t rows, int c_in, int64_t classes, int64_t ignore);
}  // namespace bf

}  // namespace pio
