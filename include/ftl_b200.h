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
/* Same, and sumsq (fp32 [ftl_rmsnorm_sumsq_slots(dim)], may be NULL) receives sum(dw^2) in pieces: the gradient-norm
 * partial of this weight, produced where the gradient is written (reference: utils.py:58-63 re-reads every gradient). */
int ftl_rmsnorm_sumsq_slots(int dim);
int ftl_rmsnorm_bwd_ex(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                       void* dx, void* dw, float* scratch, float* sumsq, int rows, int dim, ftl_stream_t stream);

/* Same backward, for callers that cannot supply scratch (no counterpart in the reference, which has no
 * native code and lets PyTorch allocate).  Meant for the recovery path after a CUDA out-of-memory error and
 * for host programs that sit next to a caching allocator owning the whole device: there every cudaMalloc
 * fails, so the work area is a 4 849 664-byte array inside the library image (held by each GPU that loads
 * the library until exit).  Limit: dim <= 4096; larger requests return cudaErrorInvalidValue and set
 * ftl_last_error().  Allocates nothing.  The first call on each device resolves the array's device address
 * (do it outside stream capture) and stores it in a per-device table without taking a lock: two host threads
 * that make their first call on the same device at the same moment both store the same value, which is harmless.
 * Calls share the array: calls that may overlap on one GPU race. */
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
/* deterministic and sort-free: the CTA of the first position holding an id sums that id's dout rows in position order.
 * prev_tokens (may be NULL): ids whose rows the previous call wrote; rows this call does not write are cleared, so the
 * 1 GB table gradient never needs a memset.  sumsq (fp32 [ntok], may be NULL): sum of squares per written row. */
int ftl_embedding_bwd_direct(const int64_t* tokens, const int64_t* prev_tokens, const void* dout, void* dtable,
                             int64_t ntok, int dim, int vocab, float* sumsq, ftl_stream_t stream);

/* ---- Cross-entropy forward+backward in place over bf16 logits — reference: train.py:101-104 ---- */
int ftl_cross_entropy(void* logits, const int64_t* labels, const float* inv_n, float* row_loss, float* row_lse,
                      float* loss_out, int64_t rows, int vocab, int64_t ld, int write_grad, ftl_stream_t stream);

/* Vocabulary-chunked LM head + cross-entropy (the [T, V] logits are never held; reference: model.py:379 +
 * train.py:100-103).  Per chunk c of columns [col0, col0 + width): run the head GEMM into a [rows, width] buffer, then
 * ftl_ce_chunk_stats; after the last chunk ftl_ce_combine gives lse / row losses / loss.  Backward: recompute the chunk,
 * ftl_ce_chunk_grad turns it into d(logits) in place, and the two gradient GEMMs read it (ftl_gemm_bf16_acc32 keeps the
 * running d(hidden) in fp32 across chunks). */
int ftl_ce_chunk_stats(const void* chunk, int64_t ld, const int64_t* labels, int col0, int width, int n_chunks, int c,
                       float* cmax, float* csum, float* picked, int64_t rows, ftl_stream_t stream);
int ftl_ce_combine(const float* cmax, const float* csum, const float* picked, const int64_t* labels, int n_chunks,
                   int64_t rows, int vocab, const float* inv_n, float* lse, float* row_loss, float* loss_out,
                   ftl_stream_t stream);
int ftl_ce_chunk_grad(void* chunk, int64_t ld, const int64_t* labels, int col0, int width, int vocab, const float* lse,
                      const float* inv_n, int64_t rows, ftl_stream_t stream);

/* ---- Gradient norm + fused AdamW over flat bf16 buffers — reference: utils.py:58-63, train.py:68,107-109 ----
 * hyper (fp32[8]) = lr, beta1, beta2, eps, weight_decay, 1-beta1^t, sqrt(1-beta2^t), max_norm
 * stats (fp32[4]) = grad_norm, clip_coef, finite, sum of squares */
int ftl_grad_norm_partials(void);
int ftl_grad_norm(const void* g, int64_t n, const float* hyper, float* scratch, float* stats, int clip,
                  ftl_stream_t stream);
int ftl_adamw_step(void* p, const void* g, void* m, void* v, int64_t n, const float* hyper, const float* stats,
                   ftl_stream_t stream);

/* ---- tcgen05 GEMMs (csrc/gemm_tc.cu), bf16 in, fp32 accumulate, bf16 out -------------------------------------------
 * C[M,N] = op(A) op(B)^T (+ op(A2) op(B2)^T when a2 != NULL: the split reduction used for dgate@W1 + dup@W3).
 * a_mn / b_mn = 1: the operand is stored with its non-reduction dimension contiguous (no transposes for dgrad/wgrad).
 * max_ctas > 0 caps the persistent grid (lends SMs to a concurrent communication kernel). */
int ftl_gemm_bf16(const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb, int b_mn, void* c, int64_t ldc,
                  int M, int N, int K, const void* a2, int64_t lda2, const void* b2, int64_t ldb2, int K2, int max_ctas,
                  ftl_stream_t stream);
/* Same; sumsq (fp32 [ftl_gemm_sumsq_slots()], may be NULL) receives per-(CTA, epilogue warp) sums of squares of the
 * stored bf16 values: the weight-gradient GEMMs emit their gradient-norm partials (reference: utils.py:58-63). */
int ftl_gemm_sumsq_slots(void);
int ftl_gemm_bf16_ex(const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb, int b_mn, void* c, int64_t ldc,
                     int M, int N, int K, const void* a2, int64_t lda2, const void* b2, int64_t ldb2, int K2,
                     float* sumsq, int max_ctas, ftl_stream_t stream);
/* c32 (+)= a [M,K] @ b [K,N] with an fp32 running sum across calls.  mode bit 0: add to what c32 holds (else overwrite);
 * bit 1: also write the bf16 result to c (c may be NULL otherwise).  c32: fp32 [M, N], leading dimension ldc, 32-byte
 * aligned.  No counterpart in the reference (autograd accumulates the head's input gradient in one GEMM). */
int ftl_gemm_bf16_acc32(const void* a, int64_t lda, const void* b, int64_t ldb, void* c, float* c32, int64_t ldc, int M,
                        int N, int K, int mode, int max_ctas, ftl_stream_t stream);
/* SwiGLU backward in the epilogue of the w2 dgrad GEMM: dact = dx [M,K] @ w2 [K,H] stays in TMEM; reads the saved
 * gate / up, writes dgate / dup (all [M,H], leading dimension ldo; H, ldo multiples of 16) — autograd of model.py:254 */
int ftl_gemm_swiglu_bwd(const void* dx, int64_t lddx, const void* w2, int64_t ldw, const void* gate, const void* up,
                        void* dgate, void* dup, int64_t ldo, int M, int H, int K, int max_ctas, ftl_stream_t stream);
/* act = silu(x W1^T) * (x W3^T); gate / up (pre-activations, kept for the backward) may be NULL */
int ftl_gemm_swiglu(const void* x, int64_t ldx, const void* w1, const void* w3, int64_t ldw, void* act, void* gate,
                    void* up, int64_t ldo, int M, int H, int K, int max_ctas, ftl_stream_t stream);
/* qkv = x Wqkv^T with the rotary embedding applied to the first rot_cols output columns in the epilogue */
int ftl_gemm_qkv_rope(const void* x, int64_t ldx, const void* wqkv, int64_t ldw, void* qkv, int64_t ldo,
                      const float* cos_t, const float* sin_t, int M, int N, int K, int seq_len, int head_dim,
                      int rot_cols, int max_ctas, ftl_stream_t stream);
/* Same two GEMMs with the weight all-gather fused in (fully-sharded mode): the epilogue warps pull 16-row slices of the
 * peers' shards into wqkv / w1 / w3 while the producer waits per 128-row block.  scratch: int[2 * blocks], zeroed. */
int ftl_gemm_qkv_rope_gather(const void* x, int64_t ldx, void* wqkv, int64_t ldw, void* qkv, int64_t ldo,
                             const float* cos_t, const float* sin_t, int M, int N, int K, int seq_len, int head_dim,
                             int rot_cols, const void* const* peer_shards, int64_t shard_off, int rows_per_rank,
                             int rank, int world, int* scratch, ftl_stream_t stream);
int ftl_gemm_swiglu_gather(const void* x, int64_t ldx, void* w1, void* w3, int64_t ldw, void* act, void* gate, void* up,
                           int64_t ldo, int M, int H, int K, const void* const* peer_shards, int64_t shard_off_w1,
                           int64_t shard_off_w3, int rows_per_rank, int rank, int world, int* scratch,
                           ftl_stream_t stream);

/* ---- tcgen05 flash attention (csrc/attn_tc.cu): causal, grouped-query, reads q|k|v in place from the fused projection
 * qkv: [B*S, ld_qkv] = q heads | k heads | v heads (RoPE already applied); out: [B*S, ld_out]; lse: fp32 [B, H, S]
 * (log2 domain).  S % 128 == 0, hd in {64, 128}.  The backward writes dq | dk | dv into dqkv ([B*S, ld_qkv]) and, when
 * cos_t / sin_t are given, applies the inverse rotary embedding to dq and dk.  delta: fp32 [B, H, S] work area;
 * scratch: ftl_attn_bwd_scratch_bytes() bytes.  Forward and backward are bitwise reproducible.  Work items are handed
 * out through per-device queue words inside the library: keep the launches of one GPU on one stream. */
int ftl_attn_fwd(const void* qkv, int64_t ld_qkv, void* out, int64_t ld_out, float* lse, int B, int S, int H, int KV,
                 int hd, ftl_stream_t stream);
int64_t ftl_attn_bwd_scratch_bytes(int B, int S, int H, int KV, int hd);
int ftl_attn_bwd(const void* qkv, int64_t ld_qkv, const void* out, const void* d_out, int64_t ld_out, const float* lse,
                 float* delta, void* dqkv, const float* cos_t, const float* sin_t, float* scratch, int B, int S, int H,
                 int KV, int hd, ftl_stream_t stream);

/* ---- Symmetric heap on the CUDA virtual-memory API + NVSwitch multicast (csrc/nvls.cu) --------------------------------
 * No counterpart in the reference (single process, no communication: train.py:16 is its only rank-related line).
 * Every rank allocates with ftl_vmm_alloc and passes the returned file descriptor to its peers (unix socket,
 * SCM_RIGHTS); peers map it with ftl_vmm_import.  Sizes must be rounded with ftl_vmm_round first.  Return codes
 * >= 1000 are 1000 + CUresult.  A multicast object (one rank creates, all import, all add their device, then all bind
 * their own allocation and map) yields an address at which multimem.st writes every rank's copy and multimem.ld_reduce
 * returns the sum of all copies, computed in the switch. */
int ftl_vmm_supported(int* vmm, int* multicast);
int64_t ftl_vmm_round(int64_t bytes, int world);
int ftl_vmm_alloc(int64_t bytes, void** ptr, uint64_t* handle, int* fd);   /* set-up time only: allocates */
int ftl_vmm_import(int fd, int64_t bytes, void** ptr, uint64_t* handle);
int ftl_vmm_unmap(void* ptr, int64_t bytes, uint64_t handle);
int ftl_fd_close(int fd);
int ftl_mc_create(int world, int64_t bytes, uint64_t* mc, int* fd);
int ftl_mc_import(int fd, uint64_t* mc);
int ftl_mc_add_device(uint64_t mc);
int ftl_mc_bind_map(uint64_t mc, uint64_t mem, int64_t bytes, void** ptr);
/* NVLS variants of the two fused collectives below: mc_g / mc_p are the multicast addresses of the gradient / parameter
 * buffers.  The reduce-scatter reads shard [lo, lo+n) with multimem.ld_reduce (fp32 accumulation in the switch) into
 * g_local and emits the sum-of-squares partials; the update stores each new 16-byte parameter vector once with
 * multimem.st.  Summation order is the switch's: use the peer-load kernels when bitwise reproducibility matters. */
int ftl_nvls_reduce_scatter(const void* mc_g, void* g_local, int64_t lo, int64_t n, float* partial, const void* status,
                            int max_ctas, ftl_stream_t stream);
int ftl_nvls_adamw_allgather(void* mc_p, const void* p_local, const void* g, void* m, void* v, int64_t lo, int64_t n,
                             const float* hyper, const float* stats, const void* status, void* commits, int max_ctas,
                             ftl_stream_t stream);

/* ---- NVLink data plane (csrc/comm.cu): symmetric buffers over CUDA IPC, device barriers, fused collectives -----------
 * peer_* arguments are device arrays of `world` pointers (this rank's own buffer at index `rank`).
 * status: device uint32, 0 ok / 1 barrier timed out / 2 a peer raised the abort word; sticky, every kernel below becomes
 * a no-op once it is non-zero.  commits: device uint64 counting applied updates (may be NULL). */
int ftl_symm_alloc(void** ptr, int64_t bytes);                  /* set-up time only: allocates */
int ftl_symm_free(void* ptr);
int ftl_ipc_export(void* ptr, void* handle64);                  /* 64-byte handle to send to the peers */
int ftl_ipc_import(const void* handle64, void** ptr);
int ftl_ipc_close(void* ptr);
int ftl_comm_max_ranks(void);
int ftl_comm_flag_words(void);                                  /* uint32 words of a rank's flag area */
int ftl_barrier(void* const* peer_flags, int rank, int world, uint32_t epoch, void* status, int64_t timeout_ms,
                ftl_stream_t stream);
int ftl_raise_abort(void* const* peer_flags, int world, ftl_stream_t stream);
/* Embedding-table gradient: publish this rank's token ids to all peers, then reduce only the rows that occur in any
 * rank's list over this rank's element range [lo, hi) of the flat gradient (bit-identical to the dense fixed-order sum;
 * rows of the previous step's lists that are not rewritten are cleared).  No counterpart in the reference. */
int ftl_emb_tokens_push(const int64_t* tokens, void* const* peer_toks, int64_t slot_off, int rank, int world, int T,
                        int cap, ftl_stream_t stream);
int ftl_emb_sparse_reduce(void* const* peer_g, int rank, int world, const int64_t* toks, const int64_t* toks_prev, int T,
                          int cap, int64_t emb_off, int dim, int vocab, int64_t lo, int64_t hi, float* partial,
                          const void* status, ftl_stream_t stream);
int ftl_dp_partials(int max_ctas);                              /* floats of `partial` one reduce-scatter launch writes */
/* gradient range [lo, lo+n) of every rank summed in rank order into this rank's buffer, with the sum of squares of the
 * result in `partial` (reduce-scatter fused with the norm) */
int ftl_dp_reduce_scatter_to(void* const* peer_g, int rank, int world, int64_t lo, int64_t n, void* dst, float* partial,
                             const void* status, int max_ctas, ftl_stream_t stream);   /* sum goes to dst[0, n) */
int ftl_dp_reduce_scatter(void* const* peer_g, int rank, int world, int64_t lo, int64_t n, float* partial,
                          const void* status, int max_ctas, ftl_stream_t stream);
int ftl_dp_publish(const float* partial, int nparts, const float* loss_local, void* const* peer_slots, int rank,
                   int world, ftl_stream_t stream);
int ftl_dp_finalize(const float* slots, int world, const float* hyper, float* stats, float* loss_out, int clip,
                    void* status /* may be NULL; set to 3 when the norm is not finite */,
                    float* result /* may be NULL; fp32[5] = loss, norm, clip coefficient, finite, status */,
                    ftl_stream_t stream);
/* clip + AdamW on parameter range [lo, lo+n) with every updated vector stored to all ranks' copies (all-gather fused) */
int ftl_dp_adamw_allgather(void* const* peer_p, const void* g, void* m, void* v, int rank, int world, int64_t lo,
                           int64_t n, const float* hyper, const float* stats, const void* status, void* commits,
                           int max_ctas, ftl_stream_t stream);
int ftl_fsdp_gather(const void* const* peer_shards, int64_t off, void* dst, int64_t n_per_rank, int rank, int world,
                    const void* status, int max_ctas, ftl_stream_t stream);
int ftl_adamw_shard(void* p, const void* g, void* m, void* v, int64_t n, const float* hyper, const float* stats,
                    const void* status, void* commits, int max_ctas, ftl_stream_t stream);

/* ---- checkpoint snapshot engine (csrc/ckpt_io.cu): device segments <-> one file through pinned double buffers ---------
 * timings (double[4], may be NULL) = d2h (h2d), write (read), fsync, total seconds.  Returns 0 or -1 (ftl_ckpt_last_error). */
int ftl_ckpt_init(int64_t stage_bytes);
int ftl_ckpt_write(const char* path, int nseg, const void* const* dev_ptrs, const int64_t* sizes,
                   const int64_t* file_offsets, int nthreads, int do_fsync, double* timings);
int ftl_ckpt_read(const char* path, int nseg, void* const* dev_ptrs, const int64_t* sizes, const int64_t* file_offsets,
                  int nthreads, double* timings);
const char* ftl_ckpt_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* FTL_B200_H */
