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

}  // namespace ftl

using namespace ftl;

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
