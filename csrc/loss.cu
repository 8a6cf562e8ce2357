// Fused softmax cross-entropy forward+backward over bf16 logits (replaces K13).
//
// The reference casts the [T, V] bf16 logits to fp32 (a 1 GiB copy at V=131072), runs log_softmax and
// nll_loss(sum), divides by the token count, and autograd later produces an fp32 gradient that is cast
// back to bf16 (/root/reference/train.py:101-104).  Here one kernel reads each bf16 row twice (the second
// read hits L2: a 256 KB row per CTA), does the arithmetic in fp32, and overwrites the row in place
// with the bf16 gradient (softmax - onehot) / n_items.  Per-row losses go to a small fp32 buffer that a
// single CTA then folds in fixed order, so the loss is deterministic.  ignore_index = -100.
#include "common/util.cuh"

namespace ftl {

constexpr int kCeThreads = 1024;

__global__ void __launch_bounds__(kCeThreads)
ce_fwd_bwd_kernel(__nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels,
                  const float* __restrict__ inv_n_ptr, float* __restrict__ row_loss, float* __restrict__ row_lse,
                  int vocab, int64_t ld, int write_grad) {
    __shared__ float red[33];
    const int64_t row = blockIdx.x;
    __nv_bfloat16* lr = logits + row * ld;
    const int64_t label = labels[row];
    const bool valid = label >= 0 && label < vocab;
    const int nvec = vocab >> 3;

    // pass 1: online max / sum-exp per thread, then combine
    float m = -INFINITY, s = 0.f;
    for (int i = threadIdx.x; i < nvec; i += kCeThreads) {
        float f[8];
        unpack8(ld8(lr + i * 8), f);
        float mx = f[0];
#pragma unroll
        for (int j = 1; j < 8; ++j) mx = fmaxf(mx, f[j]);
        const float nm = fmaxf(m, mx);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += __expf(f[j] - nm);
        s = s * __expf(m - nm) + acc;
        m = nm;
    }
    const float gmax = block_max(m, red);
    const float gsum = block_sum(s * __expf(m - gmax), red);  // m=-inf, s=0 -> 0 * exp(-inf) = 0
    const float lse = gmax + __logf(gsum);
    if (threadIdx.x == 0) {
        const float picked = valid ? __bfloat162float(lr[label]) : 0.f;
        row_loss[row] = valid ? (lse - picked) : 0.f;
        if (row_lse) row_lse[row] = lse;
    }
    if (!write_grad) return;
    __syncthreads();  // the label logit is read above before anyone overwrites it
    const float scale = valid ? *inv_n_ptr : 0.f;
    for (int i = threadIdx.x; i < nvec; i += kCeThreads) {
        float f[8];
        unpack8(ld8(lr + i * 8), f);
        const int base = i << 3;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float p = __expf(f[j] - lse);
            if (base + j == label) p -= 1.f;
            f[j] = p * scale;
        }
        st8(lr + i * 8, pack8(f));
    }
}

// out[0] = inv_n * sum(row_loss) folded by one CTA in a fixed order
__global__ void loss_reduce_kernel(const float* __restrict__ row_loss, int64_t rows, const float* __restrict__ inv_n_ptr,
                                   float* __restrict__ out) {
    __shared__ float red[33];
    float s = 0.f;
    for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) s += row_loss[i];
    const float total = block_sum(s, red);
    if (threadIdx.x == 0) out[0] = total * inv_n_ptr[0];
}

// ---- vocabulary-chunked LM head + cross-entropy (never holds [T, V]) ---------------------------------------------------
// The head GEMM is run one vocabulary chunk at a time into a small [T, Vc] buffer that stays in L2.  Pass 1 keeps only
// three numbers per (row, chunk): max, sum of exp relative to that max, and the label's logit when it falls into the
// chunk.  After the combine the row log-sum-exp is known, and pass 2 recomputes each chunk's logits, turns them into
// d(logits) in place and feeds the two gradient GEMMs from the chunk buffer.  Costs one extra head GEMM; saves the
// T x V logits (512 MiB at seq 2048, 2 GiB at 8192) — the long-context memory lever of SURVEY §5.7.
__global__ void __launch_bounds__(256)
ce_chunk_stats_kernel(const __nv_bfloat16* __restrict__ chunk, int64_t ld, const int64_t* __restrict__ labels, int col0,
                      int width, int n_chunks, int c, float* __restrict__ cmax, float* __restrict__ csum,
                      float* __restrict__ picked) {
    __shared__ float red[33];
    const int64_t row = blockIdx.x;
    const __nv_bfloat16* lr = chunk + row * ld;
    float m = -INFINITY, s = 0.f;
    for (int i = threadIdx.x; i < (width >> 3); i += blockDim.x) {
        float f[8];
        unpack8(ld8(lr + i * 8), f);
        float mx = f[0];
#pragma unroll
        for (int j = 1; j < 8; ++j) mx = fmaxf(mx, f[j]);
        const float nm = fmaxf(m, mx);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += __expf(f[j] - nm);
        s = s * __expf(m - nm) + acc;
        m = nm;
    }
    const float gmax = block_max(m, red);
    const float gsum = block_sum(s * __expf(m - gmax), red);
    if (threadIdx.x == 0) {
        cmax[row * n_chunks + c] = gmax;
        csum[row * n_chunks + c] = gsum;
        const int64_t label = labels[row];
        if (label >= col0 && label < col0 + width) picked[row] = __bfloat162float(lr[label - col0]);
    }
}

__global__ void ce_combine_kernel(const float* __restrict__ cmax, const float* __restrict__ csum,
                                  const float* __restrict__ picked, const int64_t* __restrict__ labels, int n_chunks,
                                  int64_t rows, int vocab, float* __restrict__ lse_out, float* __restrict__ row_loss) {
    const int64_t row = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (row >= rows) return;
    float M = -INFINITY;
    for (int c = 0; c < n_chunks; ++c) M = fmaxf(M, cmax[row * n_chunks + c]);
    float S = 0.f;
    for (int c = 0; c < n_chunks; ++c) S += csum[row * n_chunks + c] * __expf(cmax[row * n_chunks + c] - M);   // chunk order
    const float lse = M + __logf(S);
    const int64_t label = labels[row];
    const bool valid = label >= 0 && label < vocab;
    lse_out[row] = lse;
    row_loss[row] = valid ? lse - picked[row] : 0.f;
}

// chunk (logits of columns [col0, col0 + width)) -> d(loss)/d(logits) = (softmax - onehot) * inv_n, in place
__global__ void __launch_bounds__(256)
ce_chunk_grad_kernel(__nv_bfloat16* __restrict__ chunk, int64_t ld, const int64_t* __restrict__ labels, int col0, int width,
                     int vocab, const float* __restrict__ lse, const float* __restrict__ inv_n_ptr) {
    const int64_t row = blockIdx.x;
    __nv_bfloat16* lr = chunk + row * ld;
    const int64_t label = labels[row];
    const bool valid = label >= 0 && label < vocab;
    const float scale = valid ? *inv_n_ptr : 0.f, l = lse[row];
    const int64_t hot = label - col0;
    for (int i = threadIdx.x; i < (width >> 3); i += blockDim.x) {
        float f[8];
        unpack8(ld8(lr + i * 8), f);
        const int base = i << 3;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float p = __expf(f[j] - l);
            if (base + j == hot) p -= 1.f;
            f[j] = p * scale;
        }
        st8(lr + i * 8, pack8(f));
    }
}

}  // namespace ftl

using namespace ftl;

// Pass 1 of the chunked head, chunk c of n_chunks: chunk = logits of columns [col0, col0 + width) for all rows.
// cmax / csum: fp32 [rows, n_chunks]; picked: fp32 [rows] (written for the rows whose label lies in this chunk).
FTL_API int ftl_ce_chunk_stats(const void* chunk, int64_t ld, const int64_t* labels, int col0, int width, int n_chunks,
                               int c, float* cmax, float* csum, float* picked, int64_t rows, cudaStream_t stream) {
    if (width % 8 || ld % 8 || rows <= 0 || c < 0 || c >= n_chunks) return (int)cudaErrorInvalidValue;
    ce_chunk_stats_kernel<<<(unsigned)rows, 256, 0, stream>>>((const __nv_bfloat16*)chunk, ld, labels, col0, width,
                                                              n_chunks, c, cmax, csum, picked);
    FTL_LAUNCH_CHECK();
    return 0;
}

// After all chunks: row log-sum-exp, row losses and (loss_out != NULL) loss = sum(row_loss) * inv_n.
FTL_API int ftl_ce_combine(const float* cmax, const float* csum, const float* picked, const int64_t* labels, int n_chunks,
                           int64_t rows, int vocab, const float* inv_n, float* lse, float* row_loss, float* loss_out,
                           cudaStream_t stream) {
    if (rows <= 0 || n_chunks <= 0) return (int)cudaErrorInvalidValue;
    ce_combine_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, stream>>>(cmax, csum, picked, labels, n_chunks, rows, vocab,
                                                                          lse, row_loss);
    FTL_LAUNCH_CHECK();
    if (loss_out) {
        loss_reduce_kernel<<<1, 1024, 0, stream>>>(row_loss, rows, inv_n, loss_out);
        FTL_LAUNCH_CHECK();
    }
    return 0;
}

// Pass 2: the recomputed chunk of logits becomes d(logits) in place.
FTL_API int ftl_ce_chunk_grad(void* chunk, int64_t ld, const int64_t* labels, int col0, int width, int vocab,
                              const float* lse, const float* inv_n, int64_t rows, cudaStream_t stream) {
    if (width % 8 || ld % 8 || rows <= 0) return (int)cudaErrorInvalidValue;
    ce_chunk_grad_kernel<<<(unsigned)rows, 256, 0, stream>>>((__nv_bfloat16*)chunk, ld, labels, col0, width, vocab, lse, inv_n);
    FTL_LAUNCH_CHECK();
    return 0;
}

// logits: bf16 [rows, ld] (first `vocab` columns used), overwritten with d(loss)/d(logits) when write_grad.
// row_loss: fp32 [rows] scratch; loss_out: fp32 [1] = sum(row_loss) * inv_n.
FTL_API int ftl_cross_entropy(void* logits, const int64_t* labels, const float* inv_n, float* row_loss, float* row_lse,
                              float* loss_out, int64_t rows, int vocab, int64_t ld, int write_grad,
                              cudaStream_t stream) {
    if (vocab % 8 || ld % 8 || rows <= 0) return (int)cudaErrorInvalidValue;
    ce_fwd_bwd_kernel<<<(unsigned)rows, kCeThreads, 0, stream>>>((__nv_bfloat16*)logits, labels, inv_n, row_loss,
                                                                row_lse, vocab, ld, write_grad);
    FTL_LAUNCH_CHECK();
    if (loss_out) {
        loss_reduce_kernel<<<1, 1024, 0, stream>>>(row_loss, rows, inv_n, loss_out);
        FTL_LAUNCH_CHECK();
    }
    return 0;
}
