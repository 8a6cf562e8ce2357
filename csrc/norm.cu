// RMSNorm forward / backward, fused with the residual add (replaces K2 + K8).
//
// Arithmetic follows /root/reference/model.py:43-48: statistics in fp32, the normalised value is
// rounded to bf16 *before* the multiply by the bf16 weight.  The residual add of the block
// (model.py:310-311, `h = x + attention(...)`) is a bf16 add and is done in the same pass.
//
// One CTA per row in the forward (row = 8 KB at dim 4096: 512 threads x one 16-byte vector held in
// registers, one block reduction).  The backward is persistent: each CTA walks rows with a
// grid stride and keeps its slice of d(weight) in registers, then writes one fp32 partial row into
// scratch; a second kernel folds the partials in fixed order (deterministic, no atomics).
#include "common/util.cuh"

namespace ftl {

constexpr int kNormMaxVec = 4;  // vectors of 8 bf16 cached per thread

template <bool kAdd>
__global__ void __launch_bounds__(1024)
rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ delta,
                   __nv_bfloat16* __restrict__ h_out, const __nv_bfloat16* __restrict__ w,
                   __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out, int dim, float eps) {
    __shared__ float red[33];
    const int64_t row = blockIdx.x;
    const int nvec = dim >> 3;
    const __nv_bfloat16* xr = x + row * dim;
    bf16x8 cache[kNormMaxVec];
    float ss = 0.f;
#pragma unroll
    for (int k = 0; k < kNormMaxVec; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
        if (i < nvec) {
            bf16x8 a = ld8(xr + i * 8);
            if (kAdd) {
                bf16x8 b = ld8(delta + row * dim + i * 8);
#pragma unroll
                for (int j = 0; j < 4; ++j) a.v[j] = __hadd2(a.v[j], b.v[j]);  // bf16 add, round-to-nearest
                st8(h_out + row * dim + i * 8, a);
            }
            cache[k] = a;
            float f[8];
            unpack8(a, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) ss += f[j] * f[j];
        }
    }
    const float total = block_sum(ss, red);
    const float rstd = rsqrtf(total / (float)dim + eps);
    if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
#pragma unroll
    for (int k = 0; k < kNormMaxVec; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
        if (i < nvec) {
            float f[8], g[8];
            unpack8(cache[k], f);
            unpack8(ld8(w + i * 8), g);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = round_bf16(f[j] * rstd) * g[j];
            st8(y + row * dim + i * 8, pack8(f));
        }
    }
}

// dx = rstd * (dy*w - xhat * mean(dy*w*xhat)) [+ dres];  dw_partial[cta] = sum_rows dy * bf16(xhat)
template <bool kAddRes>
__global__ void __launch_bounds__(256)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                   const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd_in,
                   const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                   float* __restrict__ dw_partial, int rows, int dim) {
    __shared__ float red[33];
    const int nvec = dim >> 3;
    float dw_acc[kNormMaxVec][8];
    float wreg[kNormMaxVec][8];
#pragma unroll
    for (int k = 0; k < kNormMaxVec; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
#pragma unroll
        for (int j = 0; j < 8; ++j) dw_acc[k][j] = 0.f;
        if (i < nvec) unpack8(ld8(w + i * 8), wreg[k]);
    }
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const float rstd = rstd_in[row];
        float xh[kNormMaxVec][8], dyw[kNormMaxVec][8];
        float dot = 0.f;
#pragma unroll
        for (int k = 0; k < kNormMaxVec; ++k) {
            const int i = threadIdx.x + k * blockDim.x;
            if (i < nvec) {
                float d[8];
                unpack8(ld8(x + row * dim + i * 8), xh[k]);
                unpack8(ld8(dy + row * dim + i * 8), d);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xh[k][j] *= rstd;
                    dw_acc[k][j] += d[j] * round_bf16(xh[k][j]);
                    dyw[k][j] = d[j] * wreg[k][j];
                    dot += dyw[k][j] * xh[k][j];
                }
            }
        }
        const float mean = block_sum(dot, red) / (float)dim;
#pragma unroll
        for (int k = 0; k < kNormMaxVec; ++k) {
            const int i = threadIdx.x + k * blockDim.x;
            if (i < nvec) {
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = rstd * (dyw[k][j] - xh[k][j] * mean);
                if (kAddRes) {
                    float r[8];
                    unpack8(ld8(dres + row * dim + i * 8), r);
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] += r[j];
                }
                st8(dx + row * dim + i * 8, pack8(o));
            }
        }
    }
    float* out = dw_partial + (int64_t)blockIdx.x * dim;
#pragma unroll
    for (int k = 0; k < kNormMaxVec; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
        if (i < nvec) {
            *reinterpret_cast<float4*>(out + i * 8) = make_float4(dw_acc[k][0], dw_acc[k][1], dw_acc[k][2], dw_acc[k][3]);
            *reinterpret_cast<float4*>(out + i * 8 + 4) = make_float4(dw_acc[k][4], dw_acc[k][5], dw_acc[k][6], dw_acc[k][7]);
        }
    }
}

__global__ void rmsnorm_dw_finalize_kernel(const float* __restrict__ partial, int nparts, int dim,
                                           __nv_bfloat16* __restrict__ dw) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= dim) return;
    float s = 0.f;
    for (int p = 0; p < nparts; ++p) s += partial[(int64_t)p * dim + j];  // fixed order
    dw[j] = __float2bfloat16_rn(s);
}

static inline int norm_threads(int dim) {
    int nvec = dim / 8;
    int t = (nvec + kNormMaxVec - 1) / kNormMaxVec;
    if (nvec >= 256 && t < 256) t = 256;  // prefer one or two vectors per thread at dim 2048-4096
    t = (t + 31) / 32 * 32;
    if (t < 32) t = 32;
    return t;
}

}  // namespace ftl

using namespace ftl;

FTL_API int ftl_rmsnorm_partials(int rows) { return rows < 2 * kNumSMs ? rows : 2 * kNumSMs; }

FTL_API int ftl_rmsnorm_fwd(const void* x, const void* delta, void* h_out, const void* w, void* y, float* rstd,
                            int rows, int dim, float eps, cudaStream_t stream) {
    if (dim % 8 || dim > 8 * kNormMaxVec * 1024 || rows <= 0) return (int)cudaErrorInvalidValue;
    const int threads = norm_threads(dim);
    auto X = (const __nv_bfloat16*)x; auto D = (const __nv_bfloat16*)delta; auto H = (__nv_bfloat16*)h_out;
    auto W = (const __nv_bfloat16*)w; auto Y = (__nv_bfloat16*)y;
    if (delta) rmsnorm_fwd_kernel<true><<<rows, threads, 0, stream>>>(X, D, H, W, Y, rstd, dim, eps);
    else rmsnorm_fwd_kernel<false><<<rows, threads, 0, stream>>>(X, nullptr, nullptr, W, Y, rstd, dim, eps);
    FTL_LAUNCH_CHECK();
    return 0;
}

// scratch: fp32 [ftl_rmsnorm_partials(rows), dim]
FTL_API int ftl_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                            void* dx, void* dw, float* scratch, int rows, int dim, cudaStream_t stream) {
    if (dim % 8 || dim > 8 * kNormMaxVec * 256 || rows <= 0) return (int)cudaErrorInvalidValue;
    int threads = norm_threads(dim);  // the backward keeps 4 fp32 row slices in registers: at most 256 threads
    if (threads > 256) threads = 256;
    const int parts = ftl_rmsnorm_partials(rows);
    auto DY = (const __nv_bfloat16*)dy; auto X = (const __nv_bfloat16*)x; auto W = (const __nv_bfloat16*)w;
    auto DR = (const __nv_bfloat16*)dres; auto DX = (__nv_bfloat16*)dx;
    if (dres) rmsnorm_bwd_kernel<true><<<parts, threads, 0, stream>>>(DY, X, W, rstd, DR, DX, scratch, rows, dim);
    else rmsnorm_bwd_kernel<false><<<parts, threads, 0, stream>>>(DY, X, W, rstd, nullptr, DX, scratch, rows, dim);
    FTL_LAUNCH_CHECK();
    rmsnorm_dw_finalize_kernel<<<(dim + 255) / 256, 256, 0, stream>>>(scratch, parts, dim, (__nv_bfloat16*)dw);
    FTL_LAUNCH_CHECK();
    return 0;
}
