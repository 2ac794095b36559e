/* pio_b200 — C ABI of the B200-native Perceiver IO hot path (libpio_b200.so).
 *
 * The reference (felixyu7/perceiver-io-1) is pure Python and has no FFI of its own: its
 * boundary for this path is the Python class API of perceiver/model.py + perceiver/adapter.py
 * (SURVEY.md §8b).  Each entry point below names the reference code whose arithmetic it
 * replaces; the drop-in Python classes in perceiver-io-1_b200/perceiver/ bind these through
 * ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated otherwise; tensors are dense row-major
 *     fp32 (int64 for token ids / labels, uint8 for boolean masks) unless a stride is given;
 *   - nothing here allocates device memory: scratch comes from the caller (`ws`, `ws_bytes`),
 *     saved-for-backward state lives in an opaque caller-owned buffer (`saved`) whose size is
 *     given by the matching pio_*_saved_bytes();
 *   - every launch goes to the caller's stream (`cudaStream_t` passed as void*); calls are
 *     asynchronous, re-entrant, and may come from any host thread (autograd's worker thread);
 *   - return value: PIO_OK or a PIO_ERR_* code; pio_last_error() gives the text (thread-local).
 *     No C++ exception crosses this boundary.  There is no CPU fallback anywhere.
 *   - `accumulate` (backward entry points): 0 = gradients are overwritten, 1 = added to (deterministic
 *     partial sums + reduction pass), 2 = added to with fp32 atomics straight from the producing kernels
 *     (destinations must hold zeros or a running sum; no reduction passes; summation order, hence the last
 *     bits, vary run to run).  2 is honoured by the PIO_MODE_BF16 pipelines and treated as 1 elsewhere.
 *     pio_kv_project_bwd and pio_cross_attend_bwd (PIO_MODE_BF16) also take, OR-ed in, PIO_ACC_DX and PIO_ACC_DKV:
 *     the ACTIVATION gradient is added into what its buffer already holds instead of overwriting it (an input that a
 *     second operator consumes too — the caller then has nothing left to sum).
 *   - `mode`: PIO_MODE_FP32 = fp32 FFMA arithmetic end to end (parity mode, rel 1e-4);
 *             PIO_MODE_BF16 = bf16 tcgen05 tensor-core contractions, fp32 accumulate/residual
 *             stream (performance mode, rel 1e-2).
 */
#ifndef PIO_B200_H
#define PIO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PIO_OK 0
#define PIO_ERR_ARG 1
#define PIO_ERR_CUDA 2
#define PIO_ERR_WORKSPACE 3
#define PIO_ERR_UNSUPPORTED 4

#define PIO_MODE_FP32 0
#define PIO_MODE_BF16 1

typedef void* pio_stream_t; /* cudaStream_t */

const char* pio_last_error(void);
int pio_version(void);
/* number of CUDA kernels this library has launched in this process (bench.py's gpu_launches) */
long long pio_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Input adapters
 * ---------------------------------------------------------------------------------------- */

/* ImageInputAdapter.forward, adapter.py:99-109: out[b,m,:] = concat(x[b,m,0:c_img], pos[m,0:c_pos]).
 * x (batch, m, c_img) is the image with its spatial dims flattened; out (batch, m, c_img+c_pos). */
int pio_image_encode(const float* x, const float* pos_enc, float* out, int64_t batch, int64_t m,
                     int c_img, int c_pos, pio_stream_t stream);

/* ImageInputAdapter._positions/_position_encodings, adapter.py:53-94, generated on the device:
 * out (prod(shape), ndim*(2*bands+1)); `shape` is a HOST pointer to ndim ints (ndim <= 4). */
int pio_fourier_table(float* out, const int* shape, int ndim, int bands, pio_stream_t stream);

/* TextInputAdapter.forward, adapter.py:127-133: out[b,i,:] = emb[ids[b,i],:]*scale + pos[i,:]. */
int pio_text_embed_fwd(const int64_t* ids, const float* emb, const float* pos, float* out,
                       int64_t batch, int64_t l, int c, int64_t vocab, float scale, pio_stream_t stream);
/* its backward: demb (vocab,c) must be zero-filled by the caller (rows are accumulated with
 * atomics); dpos (l,c) is overwritten with the sum over the batch. */
int pio_text_embed_bwd(const int64_t* ids, const float* dout, float* demb, float* dpos,
                       int64_t batch, int64_t l, int c, int64_t vocab, float scale, pio_stream_t stream);

/* TextMasking.forward, model.py:265-293, split at its one data-dependent point (the size of the
 * randint draw).  Elements are processed in blocks of PIO_MASK_BLOCK.
 *   select: flags[i] bit0 = selected (label kept), bit1 = replace by [MASK], bit2 = replace by a
 *           random token; block_counts[j] = number of bit2 elements in block j.
 *   apply : ids[i] <- mask_token_id / random_ids[rank of i among bit2 elements] ; labels[i] <-
 *           original id if selected else -100.  In place on ids like the reference. */
#define PIO_MASK_BLOCK 1024
int pio_text_mask_select(const int64_t* ids, const uint8_t* pad, const float* r0, const float* r1,
                         const float* r2, uint8_t* flags, int32_t* block_counts, int64_t n,
                         int64_t unk_token_id, float mask_p, pio_stream_t stream);
int pio_text_mask_apply(int64_t* ids, int64_t* labels, const uint8_t* flags, const int32_t* block_counts,
                        const int64_t* random_ids, int64_t n_random, int64_t n, int64_t mask_token_id,
                        pio_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Primitives (torch ops the reference delegates to; exported for the parity tests)
 * ---------------------------------------------------------------------------------------- */

/* nn.LayerNorm (model.py:22,89,90,108), eps 1e-5 in the reference. stats (rows,2) = mean, rstd. */
int pio_layernorm_fwd(const float* x, const float* w, const float* b, float* y, float* stats,
                      int64_t rows, int c, float eps, pio_stream_t stream);
/* dx = LN'(dy) (+ dres if non-null; dx itself may be null when the input needs no gradient);
 * dw/db overwritten (accumulate=0) or added to (1). */
int pio_layernorm_bwd(const float* dy, const float* x, const float* stats, const float* w,
                      const float* dres, float* dx, float* dw, float* db, int accumulate,
                      int64_t rows, int c, void* ws, size_t ws_bytes, pio_stream_t stream);
size_t pio_layernorm_bwd_ws_bytes(int64_t rows, int c);

/* No counterpart in the reference (which is Python over torch's allocator): pio_layernorm_bwd for
 * callers that cannot allocate device memory — code running under stream capture or in hooks where
 * cudaMalloc / torch's allocator is off limits, plugin hosts that pass only inputs and outputs,
 * embedders with a fixed device-memory budget.  Identical arguments and results, but the scratch for
 * the dgamma / dbeta partial sums is a static device array inside the library (1 MiB per GPU that
 * loads it, held until exit) instead of a workspace.  c must be <= PIO_LN_NOALLOC_MAX_C (PIO_ERR_ARG
 * otherwise).  The first call on a device looks up the array's address and should be made before any
 * stream capture.  All calls on a device share the array: calls that can overlap on the GPU race, so
 * serialise them; stale contents from the previous call are harmless (every partial row read is
 * written first by the same call).  Calls no cudaMalloc / cudaFree / allocating routine. */
#define PIO_LN_NOALLOC_MAX_C 1024
int pio_layernorm_bwd_noalloc(const float* dy, const float* x, const float* stats, const float* w,
                              const float* dres, float* dx, float* dw, float* db, int accumulate,
                              int64_t rows, int c, pio_stream_t stream);

/* C[m,n] = sum_k A(m,k) B(n,k) with a fused epilogue — nn.Linear forward / backward
 * (model.py:23,25; adapter.py:146-149) and the MHA projections (model.py:66-74).
 *   a_kmajor: A(m,k) = a[m*lda+k] else a[k*lda+m];  b_kmajor: B(n,k) = b[n*ldb+k] else b[k*ldb+n].
 *   act: 0 none | 1 GELU(erf) (pre-activation stored to aux if non-null) | 2 multiply by GELU'(aux).
 *   rowsum_a (m) optional: sum_k A(m,k) (bias gradient riding on a weight-gradient GEMM). */
typedef struct {
    const float* a;
    const float* b;
    float* c;
    int64_t m, n, k;
    int64_t lda, ldb, ldc;
    int a_kmajor, b_kmajor;
    const float* bias;
    const float* residual;
    int64_t ldr;
    float* aux;
    int64_t ldaux;
    int act;
    int accumulate;
    float* rowsum_a;
    float dropout_p;       /* Residual's dropout (model.py:51,56): applied to the value before `+ residual` */
    uint64_t dropout_seed; /* element (m, n) is kept iff hash(seed, m*n_cols + n) >= p * 2^32, scaled by 1/(1-p) */
} pio_gemm_desc;
int pio_gemm_f32(const pio_gemm_desc* d, void* ws, size_t ws_bytes, pio_stream_t stream);
size_t pio_gemm_f32_ws_bytes(int64_t m, int64_t n, int64_t k);

/* The same product on the bf16 tensor cores (tcgen05.mma, TMA-staged operands, fp32 accumulator in
 * TMEM) for PIO_MODE_BF16.  a / b are bf16; leading dimensions must be multiples of 8 elements and
 * bases 16-byte aligned (TMA).  Outputs: fp32 (c32) and/or bf16 (c16).  aux16 (bf16) carries the GELU
 * derivative between the passes: act 1 stores GELU'(pre-activation) to it (one erf serves both values),
 * act 2 multiplies by it — the backward epilogue is a plain product.  No rowsum: bias gradients are
 * reduced by pio_colsum_bf16. */
typedef struct {
    const void* a;
    const void* b;
    int64_t m, n, k;
    int64_t lda, ldb;
    int a_kmajor, b_kmajor;
    float* c32;
    int64_t ldc32;
    void* c16;
    int64_t ldc16;
    const float* bias;
    const float* residual;
    int64_t ldr;
    void* aux16;
    int64_t ldaux;
    int act;
    int accumulate;
    float dropout_p;       /* as in pio_gemm_desc */
    uint64_t dropout_seed;
    /* Optional device-side row count (NULL = off): v = clamp(*dyn_rows - dyn_offset, 0, m or k) replaces
     * m (dyn_axis 0: output tiles at or beyond row v are skipped, rows >= v are not written) or k
     * (dyn_axis 1: 64-row k-blocks at or beyond v are skipped; needs accumulate == 2, and operand rows
     * between v and the next multiple of 64 must be finite with one side zero).  Lets a captured graph
     * run a product whose row count is only known on the device (rows with a label). */
    const int32_t* dyn_rows;
    int64_t dyn_offset;
    int dyn_axis;
} pio_gemm_bf16_desc;
int pio_gemm_bf16(const pio_gemm_bf16_desc* d, void* ws, size_t ws_bytes, pio_stream_t stream);
size_t pio_gemm_bf16_ws_bytes(int64_t m, int64_t n, int64_t k);
/* fp32 (rows, cols; pitch ld_src) -> bf16 with pitch ld_dst >= cols (pad columns zero-filled). */
int pio_cast_bf16(const float* src, void* dst, int64_t rows, int cols, int64_t ld_src, int64_t ld_dst,
                  pio_stream_t stream);
/* The same cast for a whole set of weight matrices in ONE launch: `items` is a table of `count`
 * entries in DEVICE memory (src is dense, pitch = cols; only the cols real columns of dst are
 * written, so dst's pad columns must have been zeroed once, e.g. by pio_cast_bf16).  Refreshes the
 * bf16 weight shadows after an optimizer step (reference lightning.py:67-72 steps torch optimizers,
 * some of which update parameters without moving their version counters). */
typedef struct {
    const float* src;
    void* dst;
    int64_t rows;
    int32_t cols;
    int32_t ld_dst;
} pio_cast_item;
int pio_cast_bf16_multi(const pio_cast_item* items, int count, pio_stream_t stream);

/* AdamW (decoupled weight decay) over a whole parameter set in one launch — the optimizer step of the
 * reference's LightningModules (torch.optim.AdamW, lightning.py:67-72).  Same update as
 * torch.optim.AdamW:  p *= 1 - lr*wd;  m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;
 * p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps), with t = hyper[5] + 1.
 * `tensors` and `chunks` live in DEVICE memory: one entry per parameter tensor, and a list cutting the tensors
 * into pieces of at most PIO_ADAMW_CHUNK elements, one CTA each.  When shadow16 is non-null the updated
 * parameter is also written as bf16 at row pitch ld16 (the weight shadow the bf16 GEMMs read), so no separate
 * cast pass follows the step.  `hyper` (DEVICE memory, 6 floats) = {lr, beta1, beta2, eps, weight_decay,
 * steps_done}: the kernel reads it, so a captured launch follows schedulers that rewrite it; pio_adamw_tick
 * adds 1 to steps_done afterwards. */
#define PIO_ADAMW_CHUNK 16384
typedef struct {
    float* p;
    const float* g;
    float* m;
    float* v;
    void* shadow16;   /* bf16 shadow of p (rows x ld16), or NULL */
    int32_t cols;     /* row length of p (for the shadow's pitch); 0 when shadow16 is NULL */
    int32_t ld16;
} pio_adamw_tensor;
typedef struct {
    int32_t tensor;   /* index into `tensors` */
    int32_t count;    /* elements in this piece (<= PIO_ADAMW_CHUNK) */
    int64_t offset;   /* first element of this piece within the tensor */
} pio_adamw_chunk;
int pio_adamw_step(const pio_adamw_tensor* tensors, const pio_adamw_chunk* chunks, int n_chunks, const float* hyper,
                   pio_stream_t stream);
int pio_adamw_tick(float* hyper, pio_stream_t stream);

/* Gather a set of fp32 tensors into one flat buffer (to_flat = 1) or scatter the flat buffer back into them,
 * multiplied by `scale` (to_flat = 0), in one launch — the flatten / unflatten around the single gradient
 * all-reduce of data-parallel training (reference: Lightning DDP, scripts/trainer.yaml:47).  `tensors[i]` =
 * {base pointer, offset of the tensor inside the flat buffer (a multiple of 4 elements)}; `chunks` as in
 * pio_adamw_step.  Both tables live in DEVICE memory. */
typedef struct {
    float* tensor;
    int64_t flat_offset;
} pio_flat_tensor;
int pio_flat_copy_f32(float* flat, const pio_flat_tensor* tensors, const pio_adamw_chunk* chunks, int n_chunks, int to_flat,
                      float scale, pio_stream_t stream);

/* In-place a *= *scale, b *= *scale, c *= *scale with the scale read from DEVICE memory; a scale of exactly 1 makes
 * the launch a no-op.  Backward of the fused linear + cross-entropy operator: its gradients are formed in the
 * forward for d(loss) = 1 and only need the incoming d(loss) applied (reference: autograd through
 * nn.CrossEntropyLoss, lightning.py:225-226).  Null pointers are allowed with a count of 0. */
int pio_scale3_f32(float* a, int64_t na, float* b, int64_t nb, float* c, int64_t nc, const float* scale, pio_stream_t stream);

/* 1 if pio_latent_block_fwd / _bwd (bf16 mode, dropout 0) would run a block of `layers` layers on these shapes with the
 * cluster kernels, one launch each way (C = 256, 4 heads, n in {32, 64, 128}, at most 8 layers and 66 row tiles; the
 * environment switches PIO_NO_FUSED_LATENT / PIO_LATENT_CLUSTER / PIO_LATENT_BWD_CLUSTER are honoured).  A caller may then
 * hand the Residual(mlp) of the cross-attention layer in front of the block (reference model.py:29-33, :36-44) to the
 * same call as an MLP-ONLY FIRST LAYER: attn[0] all null pointers, mlp[0] that MLP's weights, layers = 1 + the block's.
 * pio_latent_block_fwd / _bwd accept such a first layer in bf16 mode whether or not the cluster kernels take it (otherwise
 * it runs as pio_mlp_fwd / _bwd would); the saved buffer is sized for `layers` layers as usual. */
int pio_latent_block_cluster_ok(int layers, int batch, int n, int c, int heads);

/* Deferred join of the side streams.  The bf16 backward operators run their parameter-gradient kernels on two side
 * streams per (device, caller stream) and make the caller's stream wait for them before they return.  After
 * pio_set_side_join_deferred(1) (per host thread) pio_latent_block_bwd returns WITHOUT that wait when all it has left
 * on the side streams are parameter-gradient kernels, so they overlap with whatever the caller enqueues next; the
 * caller must then call pio_side_join(stream) before anything on `stream` reads those gradients, and must keep the
 * operator's workspace, saved buffer and inputs alive and unmodified until then.  Every other backward operator
 * still joins before it returns (which also covers work deferred earlier on the same stream). */
int pio_set_side_join_deferred(int on);
int pio_side_join(pio_stream_t stream);

/* Vocabulary logits of the text output adapter (reference adapter.py:101-109, nn.Linear) as the fused loss head forms
 * them: c (m x n, bf16, pitch ldc >= n rounded up to 8) = a (m x k, bf16, pitch lda) w (n x k, bf16, pitch ldw)^T + bias
 * (fp32, nullable), for the first *dyn_rows - dyn_off rows only when dyn_rows (DEVICE memory) is given; rows from there to
 * the end of the last live 128-row tile are written as zeros, later rows are not touched.  Persistent kernel, one CTA per
 * SM over the live tiles; k in {64, 128, 192, 256}, n >= 128, pitches multiples of 8.  Returns PIO_ERR_ARG for other
 * shapes (pio_logits_bf16_supported tells beforehand). */
int pio_logits_bf16_supported(int64_t m, int64_t n, int64_t k, int64_t lda, int64_t ldw, int64_t ldc);
int pio_logits_bf16(const void* a, int64_t lda, const void* w, int64_t ldw, const float* bias, void* c, int64_t ldc, int64_t m,
                    int64_t n, int64_t k, const int32_t* dyn_rows, int dyn_off, pio_stream_t stream);

/* Batched weight-gradient products of Linear layers that share their rows (reference: autograd through the nn.Linear /
 * nn.MultiheadAttention projections of self_attention_layer, model.py:20-26, :102-116): for every item
 *   dw (m x 256, row-major) += dy16^T x16,   db (m; nullable) += column sums of dy16
 * with dy16 (rows x m, bf16, pitch lddy elements) and x16 (rows x 256, bf16, pitch ldx).  One launch for up to 32 items
 * (at most 160 output tiles of 128 rows); n must be 256, m a multiple of 8, pitches multiples of 8, dw 16-byte aligned.
 * The sums are added with fp32 atomics: the caller zero-fills dw / db or accumulates into running gradients. */
typedef struct {
    const void* dy16;
    int64_t lddy;
    int m;
    const void* x16;
    int64_t ldx;
    int n;
    float* dw;
    float* db;
} pio_dw_item;
int pio_dw_batched(const pio_dw_item* items, int count, int64_t rows, pio_stream_t stream);

/* LayerNorm parameter gradients for a table of up to 16 LayerNorms over the same (rows, c): dgamma += sum_rows dn * xhat,
 * dbeta += sum_rows dn with xhat = (x - mean) * rstd, dn16 (rows x c, bf16) the gradient of the normalised output BEFORE
 * the affine scale's backward, stats (rows x 2) = (mean, rstd); c a multiple of 4.  fp32 atomics into dgamma / dbeta. */
typedef struct {
    const void* dn16;
    const float* x;
    const float* stats;
    float* dgamma;
    float* dbeta;
} pio_ln_grad_item;
int pio_ln_param_grads(const pio_ln_grad_item* items, int count, int64_t rows, int c, pio_stream_t stream);

/* Attention core of nn.MultiheadAttention as called at model.py:74 (dropout 0):
 * S = (q dh^-1/2) k^T, padded keys -> -inf, P = softmax(S), O = P v.  Flash-style: P is never
 * stored; lse (batch,heads,n) = logsumexp of the scaled scores, saved for the backward, which
 * recomputes P.  Head h is columns [h*dh,(h+1)*dh) of each row; *_bs / *_rs are batch / row
 * strides in elements (so a packed (rows,3E) qkv buffer is addressed without a copy).
 * A query whose keys are all padded yields NaN, like the reference.
 * mode = PIO_MODE_BF16: q, k, v point at bf16 data (same element strides); the forward then runs the
 * tcgen05 flash kernel (TMA-staged K/V tiles, QK^T and PV on the tensor cores, accumulators in TMEM)
 * when dh is 32, 64 or 128 and the fp32-arithmetic kernel otherwise; o, lse and the gradients stay fp32. */
typedef struct {
    const float* q;
    const float* k;
    const float* v;
    const uint8_t* pad; /* (batch, m) 1 = ignore key; nullable */
    float* o;
    float* lse;
    int64_t q_bs, q_rs, k_bs, k_rs, v_bs, v_rs, o_bs, o_rs;
    int batch, heads, n, m, dh;
    float scale;
} pio_attn_desc;
int pio_attn_fwd(const pio_attn_desc* d, int mode, pio_stream_t stream);

typedef struct {
    pio_attn_desc fwd;  /* q,k,v,pad,o(read),lse(read) as in the forward */
    const float* d_o;   /* same strides as o */
    float* dq;          /* strides as q */
    float* dk;          /* strides as k */
    float* dv;          /* strides as v */
    float* delta;       /* scratch (batch,heads,n) */
} pio_attn_bwd_desc;
int pio_attn_bwd(const pio_attn_bwd_desc* d, int mode, pio_stream_t stream);
/* The all-bf16 tensor-core backward used inside the PIO_MODE_BF16 operators (dh 32, 64 or 128, else
 * PIO_ERR_UNSUPPORTED): q, k, v of `fwd`, d_o (strides of o) and the outputs dq / dk / dv (strides of
 * q / k / v) are bf16; o and lse fp32.  Two kernels, both recomputing P from lse with QK^T, dO V^T and
 * the three gradient products on tcgen05: one per 128 query rows (dq, delta), one per 128 key rows
 * (dk, dv).  delta: scratch (batch, heads, n). */
int pio_attn_bwd_bf16(const pio_attn_desc* fwd, const void* d_o16, void* dq16, void* dk16, void* dv16,
                      float* delta, pio_stream_t stream);

/* nn.CrossEntropyLoss (lightning.py:61,184,225-226): mean over targets != ignore_index of
 * logsumexp(logits) - logits[target].  out3 = {loss, n_valid, loss_sum}; row_lse holds 2*rows floats: the per-row
 * logsumexp (saved for the backward) followed by per-row losses (scratch). */
int pio_cross_entropy_fwd(const float* logits, const int64_t* labels, float* out3, float* row_lse,
                          int64_t rows, int64_t classes, int64_t ld, int64_t ignore_index,
                          pio_stream_t stream);
int pio_cross_entropy_bwd(const float* logits, const int64_t* labels, const float* out3,
                          const float* row_lse, const float* dloss, float* dlogits, int64_t rows,
                          int64_t classes, int64_t ld, int64_t ignore_index, pio_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Fused operators (one call per reference module)
 * ---------------------------------------------------------------------------------------- */

/* Dropout of the reference's training mode (nn.MultiheadAttention(dropout=p) on the attention
 * probabilities, model.py:70; Residual's nn.Dropout on the branch output, model.py:51,56).  NULL or
 * p == 0 = off.  Masks are a counter-based hash of (seed, element index): the backward call must get
 * the same struct as the forward call it differentiates; nothing is stored.  Each operator derives its
 * sites from `seed` (seed+0 attention probabilities, seed+1 residual; the latent block uses seed+4*layer
 * for the attention layer and seed+4*layer+2 for the MLP).
 * torch's own Philox stream cannot be reproduced, so parity with the reference at p > 0 is statistical. */
typedef struct {
    float p;
    uint64_t seed;
} pio_dropout;

/* Residual(mlp(C)), model.py:20-26,47-56: y = Linear2(GELU(Linear1(LN(x)))) + x.
 * `residual` = 0 gives the bare module (mlp / CrossAttention / SelfAttention .forward) without
 * the `+ x` of Residual.forward (model.py:54-56); same flag on the attention operators below. */
typedef struct {
    const float *ln_w, *ln_b, *w1, *b1, *w2, *b2;
} pio_mlp_weights;
typedef struct {
    float *ln_w, *ln_b, *w1, *b1, *w2, *b2;
} pio_mlp_grads;
size_t pio_mlp_saved_bytes(int64_t rows, int c, int mode);
size_t pio_mlp_ws_bytes(int64_t rows, int c, int mode);
int pio_mlp_fwd(const float* x, float* y, const pio_mlp_weights* w, void* saved, int64_t rows, int c,
                int residual, const pio_dropout* drop, int mode, void* ws, size_t ws_bytes, pio_stream_t stream);
int pio_mlp_bwd(const float* dy, const float* x, float* dx, const pio_mlp_weights* w, const void* saved,
                const pio_mlp_grads* g, int accumulate, int64_t rows, int c, int residual,
                const pio_dropout* drop, int mode, void* ws, size_t ws_bytes, pio_stream_t stream);

/* Attention weights as torch.nn.MultiheadAttention stores them (model.py:66-71): wq (E,E),
 * wk / wv (E,C_kv); they may alias one packed (3E,E) in_proj_weight.  Biases likewise. */
typedef struct {
    const float *ln_w, *ln_b;       /* q_norm (cross) or norm (self) */
    const float *wq, *wk, *wv;
    const float *bq, *bk, *bv;
    const float *wo, *bo;
    const float *kv_ln_w, *kv_ln_b; /* kv_norm (cross-attention only) */
} pio_attn_weights;
typedef struct {
    float *ln_w, *ln_b;
    float *wq, *wk, *wv;
    float *bq, *bk, *bv;
    float *wo, *bo;
    float *kv_ln_w, *kv_ln_b;
} pio_attn_grads;

/* The key/value side of CrossAttention, model.py:90,98 + the k/v in-projection of :74:
 * kv[b,j,0:E] = LN_kv(x_kv[b,j]) Wk^T + bk,  kv[b,j,E:2E] = ... Wv^T + bv.
 * Computed once per distinct layer and shared by the repeats of layer_n (model.py:186-187).
 * In PIO_MODE_BF16 the kv tensor and its gradient dkv are bf16 (B, M, 2E) behind the float* type,
 * here and in pio_cross_attend_*. */
size_t pio_kv_project_saved_bytes(int64_t rows, int c_kv, int e, int mode);
size_t pio_kv_project_ws_bytes(int64_t rows, int c_kv, int e, int mode);
int pio_kv_project_fwd(const float* x_kv, float* kv, const pio_attn_weights* w, void* saved,
                       int64_t rows, int c_kv, int e, int mode, void* ws, size_t ws_bytes,
                       pio_stream_t stream);
/* dx_kv may be null (input needs no gradient, e.g. images).  accumulate | PIO_ACC_DX (bf16 mode): dx_kv += instead of =. */
#define PIO_ACC_DX 4   /* the input gradient (dx_kv, dx_q) is added into its buffer */
#define PIO_ACC_DKV 8  /* pio_cross_attend_bwd: the key/value gradient (bf16) is added into dkv */
int pio_kv_project_bwd(const float* dkv, const float* x_kv, float* dx_kv, const pio_attn_weights* w,
                       const void* saved, const pio_attn_grads* g, int accumulate, int64_t rows,
                       int c_kv, int e, int mode, void* ws, size_t ws_bytes, pio_stream_t stream);

/* The query side of Residual(CrossAttention), model.py:77-99 + :54-56:
 * y = out_proj(attn(LN_q(x_q) Wq^T + bq, K, V, pad)) + x_q  with kv from pio_kv_project_fwd. */
size_t pio_cross_attend_saved_bytes(int batch, int n, int m, int e, int heads, int mode);
size_t pio_cross_attend_ws_bytes(int batch, int n, int m, int e, int heads, int mode);
int pio_cross_attend_fwd(const float* x_q, const float* kv, const uint8_t* pad, float* y,
                         const pio_attn_weights* w, void* saved, int batch, int n, int m, int e,
                         int heads, int residual, const pio_dropout* drop, int mode, void* ws, size_t ws_bytes,
                         pio_stream_t stream);
int pio_cross_attend_bwd(const float* dy, const float* x_q, const float* kv, const uint8_t* pad,
                         float* dx_q, float* dkv, const pio_attn_weights* w, const void* saved,
                         const pio_attn_grads* g, int accumulate, int batch, int n, int m, int e,
                         int heads, int residual, const pio_dropout* drop, int mode, void* ws, size_t ws_bytes,
                         pio_stream_t stream);
/* accumulate | PIO_ACC_DKV needs the tensor-core attention kernels (1 here: the shape has them); accumulate | PIO_ACC_DX
 * needs residual = 2 (below), where dx_q carries no residual term of its own. */
int pio_cross_attend_merges_dkv(int batch, int n, int m, int e, int heads);

/* Residual(SelfAttention), model.py:102-116 + :54-56, one layer: y = out_proj(attn(qkv(LN(x)))) + x. */
size_t pio_self_attn_saved_bytes(int batch, int n, int c, int heads, int mode);
size_t pio_self_attn_ws_bytes(int batch, int n, int c, int heads, int mode);
int pio_self_attn_fwd(const float* x, float* y, const pio_attn_weights* w, void* saved, int batch, int n,
                      int c, int heads, int residual, const pio_dropout* drop, int mode, void* ws,
                      size_t ws_bytes, pio_stream_t stream);
int pio_self_attn_bwd(const float* dy, const float* x, float* dx, const pio_attn_weights* w,
                      const void* saved, const pio_attn_grads* g, int accumulate, int batch, int n, int c,
                      int heads, int residual, const pio_dropout* drop, int mode, void* ws, size_t ws_bytes,
                      pio_stream_t stream);

/* self_attention_block, model.py:36-44,102-116: `layers` x [Residual(SelfAttention) ->
 * Residual(mlp)] applied to x (batch,n,c).  `attn`/`mlp` are arrays of `layers` structs. */
size_t pio_latent_block_saved_bytes(int layers, int batch, int n, int c, int heads, int mode);
size_t pio_latent_block_ws_bytes(int layers, int batch, int n, int c, int heads, int mode);
int pio_latent_block_fwd(const float* x, float* y, const pio_attn_weights* attn, const pio_mlp_weights* mlp,
                         void* saved, int layers, int batch, int n, int c, int heads, const pio_dropout* drop,
                         int mode, void* ws, size_t ws_bytes, pio_stream_t stream);
int pio_latent_block_bwd(const float* dy, const float* x, float* dx, const pio_attn_weights* attn,
                         const pio_mlp_weights* mlp, const void* saved, const pio_attn_grads* attn_g,
                         const pio_mlp_grads* mlp_g, int accumulate, int layers, int batch, int n, int c,
                         int heads, const pio_dropout* drop, int mode, void* ws, size_t ws_bytes,
                         pio_stream_t stream);

/* pio_latent_block_fwd / _bwd with a FIRST LAYER THAT HAS NO ATTENTION OF ITS OWN (bf16 mode): attn[0].wq == NULL.
 *   attn[0] all null: layer 0 is only mlp[0] with its residual — the Residual(mlp) of the cross-attention layer in front
 *     of the block (reference model.py:29-33).
 *   attn[0].wo / .bo given: layer 0 first adds that cross-attention layer's out-projection and residual,
 *     x += pre_o16 Wo^T + bo, where pre_o16 (rows x c, bf16) is the attention output (pio_cross_attend_fwd with
 *     residual = 2 leaves it in its saved buffer at pio_cross_attend_o16_offset), then runs mlp[0].
 * Backward: dx is the gradient of the block's input x; attn_g[0].wo / .bo receive the out-projection's gradients and
 * pre_do16 (rows x c, bf16, written) the gradient of the attention output, to be handed to pio_cross_attend_bwd
 * (residual = 2) as its `dy`.  Shapes the cluster kernels do not take run the same arithmetic with the per-operator
 * kernels.  The plain entry points are these with pre_o16 = pre_do16 = NULL. */
int pio_latent_block_fwd_pre(const float* x, float* y, const pio_attn_weights* attn, const pio_mlp_weights* mlp, void* saved,
                             int layers, int batch, int n, int c, int heads, const pio_dropout* drop, int mode, void* ws,
                             size_t ws_bytes, pio_stream_t stream, const void* pre_o16);
int pio_latent_block_bwd_pre(const float* dy, const float* x, float* dx, const pio_attn_weights* attn, const pio_mlp_weights* mlp,
                             const void* saved, const pio_attn_grads* attn_g, const pio_mlp_grads* mlp_g, int accumulate,
                             int layers, int batch, int n, int c, int heads, const pio_dropout* drop, int mode, void* ws,
                             size_t ws_bytes, pio_stream_t stream, const void* pre_o16, void* pre_do16);
/* Byte offset of the attention output (batch*n x e, bf16) inside the saved buffer of pio_cross_attend_fwd (bf16 mode).
 * With residual = 2, pio_cross_attend_fwd stops after the attention (y is not written) and pio_cross_attend_bwd takes as
 * `dy` the bf16 gradient of that attention output, leaves wo / bo gradients alone and returns in dx_q only the gradient
 * that reaches x_q through q_norm and the Q projection. */
size_t pio_cross_attend_o16_offset(int batch, int n, int e, int heads);


/* nn.Linear of Classification/TextOutputAdapter, adapter.py:146-149: y = x W^T + b. */
size_t pio_linear_ws_bytes(int64_t rows, int c_in, int64_t c_out, int mode);
int pio_linear_fwd(const float* x, const float* w, const float* b, float* y, int64_t rows, int c_in,
                   int64_t c_out, int mode, void* ws, size_t ws_bytes, pio_stream_t stream);
/* dx nullable. dw/db overwritten or accumulated. */
int pio_linear_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, float* db,
                   int accumulate, int64_t rows, int c_in, int64_t c_out, int mode, void* ws,
                   size_t ws_bytes, pio_stream_t stream);

/* PerceiverDecoder's cross_attention_layer (model.py:203-206, :229-237: Residual(CrossAttention(q = learned output array,
 * kv = latents)) then Residual(mlp)) evaluated ONLY for the (sample, position) pairs that carry a label in a masked-LM
 * batch (PerceiverMLM with the loss of LitMaskedLanguageModel.step, model.py:306-318, lightning.py:223-226).  Each decoder
 * query is independent of the others and an ignored position adds exactly zero to the loss and to every gradient, so the
 * result (loss, gradients) is that of decoding every position.  PIO_MODE_BF16 only.
 *   pio_label_compact: idx[i] = flat position b*l + pos of the i-th labelled pair (stable order), labels_c[i] its label
 *     (`ignore` for i >= count), *count their number — all in DEVICE memory; the host never learns the count.
 *   pio_decode_rows_fwd/bwd: row i of y / dy (max_rows = batch*l rows allocated, the first *count used) belongs to pair
 *     idx[i].  queries (K >= l, e) is the decoder's learned output array, kv the bf16 (batch, n, 2e) tensor of
 *     pio_kv_project_fwd on the latents.  attn: ln_w/ln_b = q_norm, wq/bq, wo/bo (bf16 matrices); mlp as pio_mlp_fwd.
 *     Backward (accumulate must be 2): parameter gradients are ADDED into zero-filled / running buffers, dqueries (K, e)
 *     likewise; dkv (bf16, batch x n x 2e) is overwritten.  Every launch is sized for max_rows and clamps at *count, so
 *     the sequence can be captured in a CUDA graph and replayed on batches with other label counts. */
size_t pio_label_compact_ws_bytes(int64_t rows);
int pio_label_compact(const int64_t* labels, int64_t rows, int64_t classes, int64_t ignore, int32_t* idx, int64_t* labels_c,
                      int32_t* count, void* ws, size_t ws_bytes, pio_stream_t stream);
int pio_decode_rows_supported(int n, int e, int heads);
size_t pio_decode_rows_saved_bytes(int64_t max_rows, int e, int heads);
size_t pio_decode_rows_ws_bytes(int64_t max_rows, int e, int heads);
int pio_decode_rows_fwd(const float* queries, const void* kv, const int32_t* idx, const int32_t* count, float* y,
                        const pio_attn_weights* attn, const pio_mlp_weights* mlp, void* saved, int64_t max_rows, int l, int batch,
                        int n, int e, int heads, void* ws, size_t ws_bytes, pio_stream_t stream);
int pio_decode_rows_bwd(const float* dy, const void* kv, const int32_t* idx, const int32_t* count, float* dqueries, void* dkv,
                        const pio_attn_weights* attn, const pio_mlp_weights* mlp, const void* saved, const pio_attn_grads* attn_g,
                        const pio_mlp_grads* mlp_g, int accumulate, int64_t max_rows, int l, int batch, int n, int e, int heads,
                        void* ws, size_t ws_bytes, pio_stream_t stream);

/* Output-adapter Linear fused with the cross-entropy loss for training (TextOutputAdapter.linear,
 * adapter.py:146-149,166-173 + nn.CrossEntropyLoss, lightning.py:184,225-226).
 * PIO_MODE_BF16 only (`w` = bf16 shadow).  Rows with an ignored label contribute exactly zero to the loss
 * and to every gradient, so only the valid rows are projected: they are compacted on the device (the count
 * never reaches the host — kernels are launched for the worst case and skip tiles past the count, so the call
 * can be captured in a CUDA graph and replayed on other labels) and go through in L2-sized chunks: logits
 * GEMM (bf16), one kernel producing per-row losses and bf16 dlogits in place, the two gradient GEMMs and the
 * bias column sums.  No logit is stored in HBM.  out3 = {loss, n_valid, loss_sum}; dx (rows,c_in; zero rows
 * for ignored labels), dw (classes,c_in), db (classes) are the gradients for d(loss) = 1 (scale them by the
 * incoming gradient). */
size_t pio_linear_ce_ws_bytes(int64_t rows, int c_in, int64_t classes, int mode);
int pio_linear_ce_fwd_bwd(const float* x, const float* w, const float* b, const int64_t* labels, float* out3,
                          float* dx, float* dw, float* db, int64_t rows, int c_in, int64_t classes,
                          int64_t ignore_index, int mode, void* ws, size_t ws_bytes, pio_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PIO_B200_H */
