/* C ABI of libftl_b200.so — the hand-written sm_100a kernels of the B200 training path.
 *
 * Conventions: every pointer is a device pointer unless named host_*; every bf16 tensor is row-major
 * and 16-byte aligned; every function enqueues work on `stream` and returns 0 or a cudaError_t value
 * (no function synchronises, and none allocates memory unless its comment says so).  The reference
 * (danilodjor/fault-tolerant-llm-training) has no native code; each entry names the reference lines
 * whose PyTorch ops it replaces.
 */
#ifndef FTL_B200_H
#define FTL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* ftl_stream_t;

/* ---- RMSNorm (+ residual add)  — reference: model.py:43-48, 310-311 ---- */
/* y = bf16(x*rstd)*w; if delta != NULL first h_out = x + delta (bf16 add) and the norm is taken of h_out.
 * rstd (fp32 [rows]) may be NULL. */
int ftl_rmsnorm_fwd(const void* x, const void* delta, void* h_out, const void* w, void* y, float* rstd,
                    int rows, int dim, float eps, ftl_stream_t stream);
/* rows of the fp32 scratch the backward needs: scratch is [ftl_rmsnorm_partials(rows), dim] */
int ftl_rmsnorm_partials(int rows);
/* dx = d(norm)/dx (+ dres if not NULL, the residual-path gradient); dw = bf16 weight gradient. */
int ftl_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                    void* dx, void* dw, float* scratch, int rows, int dim, ftl_stream_t stream);

/* Same backward, for callers that cannot supply scratch (no counterpart in the reference, which has no
 * native code and lets PyTorch allocate).  Meant for the recovery path after a CUDA out-of-memory error and
 * for host programs that sit next to a caching allocator owning the whole device: there every cudaMalloc
 * fails, so the work area is a 4 849 664-byte array inside the library image (held by each GPU that loads
 * the library until exit).  Limit: dim <= 4096; larger requests return cudaErrorInvalidValue and set
 * ftl_last_error().  Allocates nothing.  The first call on each device resolves the array's device address
 * (do it outside stream capture).  Calls share the array: calls that may overlap on one GPU race. */
int ftl_rmsnorm_bwd_static(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                           void* dx, void* dw, int rows, int dim, ftl_stream_t stream);
int64_t ftl_rmsnorm_bwd_static_capacity(void);   /* bytes of the built-in work area */
const char* ftl_last_error(void);                /* message of the last refused call on this thread ("" if none) */

/* ---- RoPE, in place on the q|k column block of a fused qkv buffer — reference: model.py:100-126 ---- */
int ftl_rope_inplace(void* qkv, const float* cos_t, const float* sin_t, int64_t tokens, int seq_len,
                     int row_stride, int rot_cols, int head_dim, int inverse, ftl_stream_t stream);

/* ---- SwiGLU — reference: model.py:254 ---- */
int ftl_swiglu_fwd(const void* gate, const void* up, void* act, int64_t n, ftl_stream_t stream);
int ftl_swiglu_bwd(const void* dact, const void* gate, const void* up, void* dgate, void* dup, int64_t n,
                   ftl_stream_t stream);

/* ---- Embedding — reference: model.py:373 ---- */
int ftl_embedding_fwd(const int64_t* tokens, const void* table, void* out, int64_t ntok, int dim, int vocab,
                      ftl_stream_t stream);
/* deterministic: sorted_tokens/perm = stable ascending sort of the token ids; dtable must be zero */
int ftl_embedding_bwd(const int64_t* sorted_tokens, const int64_t* perm, const void* dout, void* dtable,
                      int64_t ntok, int dim, int vocab, ftl_stream_t stream);

/* ---- Cross-entropy forward+backward in place over bf16 logits — reference: train.py:101-104 ---- */
int ftl_cross_entropy(void* logits, const int64_t* labels, const float* inv_n, float* row_loss, float* row_lse,
                      float* loss_out, int64_t rows, int vocab, int64_t ld, int write_grad, ftl_stream_t stream);

/* ---- Gradient norm + fused AdamW over flat bf16 buffers — reference: utils.py:58-63, train.py:68,107-109 ----
 * hyper (fp32[8]) = lr, beta1, beta2, eps, weight_decay, 1-beta1^t, sqrt(1-beta2^t), max_norm
 * stats (fp32[4]) = grad_norm, clip_coef, finite, sum of squares */
int ftl_grad_norm_partials(void);
int ftl_grad_norm(const void* g, int64_t n, const float* hyper, float* scratch, float* stats, int clip,
                  ftl_stream_t stream);
int ftl_adamw_step(void* p, const void* g, void* m, void* v, int64_t n, const float* hyper, const float* stats,
                   ftl_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* FTL_B200_H */
