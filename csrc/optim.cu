// Gradient norm + fused AdamW over flat bf16 buffers (replaces K16 + K18).
//
// The reference runs torch's foreach AdamW over 291 bf16 tensors — about ten elementwise passes, ≈320 GB
// of HBM traffic per step for the 8 B model — after `get_total_norm` (utils.py:58-63, train.py:107-109).
// Here parameters, gradients and both moments live in four flat bf16 buffers and one kernel makes a
// single pass: read p, g, m, v; fp32 math; write p, m, v (7 x 2 B per element = 113 GB).  The update
// rule and its defaults are torch.optim.AdamW's (decoupled weight decay 0.01 on every parameter, betas
// (0.9, 0.999), eps 1e-8: train.py:68); moments stay bf16 like the reference's.
//
// Hyper-parameters and the clip coefficient are read from device memory so that the launch sequence
// is identical every step (CUDA-graph friendly) and so that no host sync is needed between the norm
// and the update:
//   hyper[0..7] = lr, beta1, beta2, eps, weight_decay, 1-beta1^t, sqrt(1-beta2^t), max_norm
//   stats[0..3] = grad_norm, clip_coef, finite (1/0), sum of squares
// A non-finite norm turns the update into a no-op (state stays clean for the checkpoint that follows;
// the reference raises from get_total_norm(error_if_nonfinite=True) instead).
#include "common/util.cuh"

namespace ftl {

constexpr int kNormSqThreads = 512;

// partial[b] = sum of squares of a contiguous chunk; chunking and in-block order are fixed.
__global__ void __launch_bounds__(kNormSqThreads)
grad_sumsq_kernel(const __nv_bfloat16* __restrict__ g, int64_t nvec, float* __restrict__ partial) {
    __shared__ float red[33];
    const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
    const int64_t lo = blockIdx.x * per;
    const int64_t hi = lo + per < nvec ? lo + per : nvec;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    int64_t i = lo + threadIdx.x;
    for (; i + 3 * kNormSqThreads < hi; i += 4 * kNormSqThreads) {  // 4 independent 16-byte loads in flight
        bf16x8 a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) a[u] = ld8_stream(g + (i + u * kNormSqThreads) * 8);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            float f[8];
            unpack8(a[u], f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[u] += f[j] * f[j];
        }
    }
    for (; i < hi; i += kNormSqThreads) {
        float f[8];
        unpack8(ld8_stream(g + i * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[0] += f[j] * f[j];
    }
    const float total = block_sum((acc[0] + acc[1]) + (acc[2] + acc[3]), red);
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}

// stats <- norm, clip coefficient (torch.nn.utils.clip_grads_with_norm_: max_norm / (norm + 1e-6), capped at 1),
// finite flag.  `extra_sumsq` (optional) adds other ranks' / buckets' contributions.
__global__ void grad_norm_finalize_kernel(const float* __restrict__ partial, int nparts, const float* __restrict__ hyper,
                                          float* __restrict__ stats, int clip) {
    __shared__ float red[33];
    float s = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) s += partial[i];
    const float total = block_sum(s, red);
    if (threadIdx.x == 0) {
        const float norm = sqrtf(total);
        const bool finite = isfinite(norm);
        float coef = 1.f;
        if (clip && finite) coef = fminf(1.f, hyper[7] / (norm + 1e-6f));
        stats[0] = norm;
        stats[1] = coef;
        stats[2] = finite ? 1.f : 0.f;
        stats[3] = total;
    }
}

__global__ void __launch_bounds__(512)
adamw_kernel(__nv_bfloat16* __restrict__ p, const __nv_bfloat16* __restrict__ g, __nv_bfloat16* __restrict__ m,
             __nv_bfloat16* __restrict__ v, int64_t nvec, const float* __restrict__ hyper,
             const float* __restrict__ stats) {
    const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
    const float bc1 = hyper[5], bc2_sqrt = hyper[6];
    const float coef = stats ? stats[1] : 1.f;
    if (stats && stats[2] == 0.f) return;  // non-finite gradient norm: leave the state untouched
    const float step_size = lr / bc1;
    const float decay = 1.f - lr * wd;
    const float vcut = adam_vcut(eps, bc2_sqrt);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float pf[8], gf[8], mf[8], vf[8];
        unpack8(ld8_stream(p + i * 8), pf);
        unpack8(ld8_stream(g + i * 8), gf);
        unpack8(ld8_stream(m + i * 8), mf);
        unpack8(ld8_stream(v + i * 8), vf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gr = gf[j] * coef;
            mf[j] = mf[j] + (1.f - b1) * (gr - mf[j]);           // lerp(m, g, 1-b1)
            vf[j] = b2 * vf[j] + (1.f - b2) * gr * gr;
            const float denom = adam_denom(vf[j], bc2_sqrt, eps, vcut);
            pf[j] = pf[j] * decay - step_size * adam_quot(mf[j], denom);
        }
        st8_stream(p + i * 8, pack8(pf));
        st8_stream(m + i * 8, pack8(mf));
        st8_stream(v + i * 8, pack8(vf));
    }
}

}  // namespace ftl

using namespace ftl;

FTL_API int ftl_grad_norm_partials(void) { return kNumSMs * 8; }

// scratch: fp32 [ftl_grad_norm_partials()].  stats: fp32 [4].
FTL_API int ftl_grad_norm(const void* g, int64_t n, const float* hyper, float* scratch, float* stats, int clip,
                          cudaStream_t stream) {
    if (n % 8 || n <= 0) return (int)cudaErrorInvalidValue;
    const int parts = ftl_grad_norm_partials();
    grad_sumsq_kernel<<<parts, kNormSqThreads, 0, stream>>>((const __nv_bfloat16*)g, n / 8, scratch);
    FTL_LAUNCH_CHECK();
    grad_norm_finalize_kernel<<<1, 1024, 0, stream>>>(scratch, parts, hyper, stats, clip);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_adamw_step(void* p, const void* g, void* m, void* v, int64_t n, const float* hyper, const float* stats,
                           cudaStream_t stream) {
    if (n % 8 || n <= 0) return (int)cudaErrorInvalidValue;
    const int64_t nvec = n / 8;
    int64_t blocks = (nvec + 511) / 512;
    const int64_t cap = (int64_t)kNumSMs * 16;
    if (blocks > cap) blocks = cap;
    adamw_kernel<<<(unsigned)blocks, 512, 0, stream>>>((__nv_bfloat16*)p, (const __nv_bfloat16*)g, (__nv_bfloat16*)m,
                                                       (__nv_bfloat16*)v, nvec, hyper, stats);
    FTL_LAUNCH_CHECK();
    return 0;
}
