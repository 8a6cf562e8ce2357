// RoPE (N7), SwiGLU elementwise (K10), embedding gather + deterministic backward (N9).
#include "common/util.cuh"

namespace ftl {

// In-place rotary embedding on the q and k column blocks of a fused [tokens, row_stride] qkv buffer.
// Adjacent pairs (x[2i], x[2i+1]) are rotated by angle[pos, i] in fp32 and rounded once to bf16 —
// the interleaved convention and rounding of /root/reference/model.py:121-126.  `sign` = -1 rotates
// backwards, which is the exact transpose and therefore the backward pass.
__global__ void rope_kernel(__nv_bfloat16* __restrict__ qkv, const float* __restrict__ cos_t,
                            const float* __restrict__ sin_t, int64_t tokens, int seq_len, int row_stride,
                            int rot_cols, int head_dim, float sign) {
    const int vec_per_row = rot_cols >> 3;
    const int64_t total = tokens * vec_per_row;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = idx / vec_per_row;
        const int c = (int)(idx - t * vec_per_row) << 3;   // column inside the rotated block
        const int pos = (int)(t % seq_len);
        const int pair0 = (c % head_dim) >> 1;              // first of 4 pairs handled here
        __nv_bfloat16* p = qkv + t * row_stride + c;
        float f[8];
        unpack8(ld8(p), f);
        const float4 cs = *reinterpret_cast<const float4*>(cos_t + (int64_t)pos * (head_dim >> 1) + pair0);
        const float4 sn = *reinterpret_cast<const float4*>(sin_t + (int64_t)pos * (head_dim >> 1) + pair0);
        const float cc[4] = {cs.x, cs.y, cs.z, cs.w};
        const float ss[4] = {sn.x * sign, sn.y * sign, sn.z * sign, sn.w * sign};
        float o[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            o[2 * j] = f[2 * j] * cc[j] - f[2 * j + 1] * ss[j];
            o[2 * j + 1] = f[2 * j] * ss[j] + f[2 * j + 1] * cc[j];
        }
        st8(p, pack8(o));
    }
}

// act = bf16(bf16(silu(gate)) * up): the two roundings of F.silu(w1 x) * w3 x in bf16 (model.py:254)
__global__ void swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gate, const __nv_bfloat16* __restrict__ up,
                                  __nv_bfloat16* __restrict__ act, int64_t nvec) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float g[8], u[8];
        unpack8(ld8(gate + i * 8), g);
        unpack8(ld8(up + i * 8), u);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = round_bf16(g[j] / (1.f + __expf(-g[j]))) * u[j];
        st8(act + i * 8, pack8(g));
    }
}

// dgate = dact * up * silu'(gate), dup = dact * silu(gate); may run in place over gate/up.
__global__ void swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ dact, const __nv_bfloat16* gate,
                                  const __nv_bfloat16* up, __nv_bfloat16* dgate, __nv_bfloat16* dup, int64_t nvec) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float d[8], g[8], u[8], og[8], ou[8];
        unpack8(ld8(dact + i * 8), d);
        unpack8(ld8(gate + i * 8), g);
        unpack8(ld8(up + i * 8), u);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float sig = 1.f / (1.f + __expf(-g[j]));
            const float silu = g[j] * sig;
            og[j] = d[j] * u[j] * (sig * (1.f + g[j] * (1.f - sig)));
            ou[j] = d[j] * silu;
        }
        st8(dgate + i * 8, pack8(og));
        st8(dup + i * 8, pack8(ou));
    }
}

__global__ void embedding_fwd_kernel(const int64_t* __restrict__ tokens, const __nv_bfloat16* __restrict__ table,
                                     __nv_bfloat16* __restrict__ out, int dim, int vocab) {
    const int64_t t = blockIdx.x;
    int64_t id = tokens[t];
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    const __nv_bfloat16* src = table + id * dim;
    for (int i = threadIdx.x; i < (dim >> 3); i += blockDim.x) st8(out + t * dim + i * 8, ld8(src + i * 8));
}

// Deterministic scatter-add: `sorted`/`perm` are the token ids sorted ascending (stable) and the
// positions they came from.  The CTA at the head of a run of equal ids sums that run in position
// order in fp32 and writes the row once.  The table gradient must be zero beforehand.
__global__ void embedding_bwd_kernel(const int64_t* __restrict__ sorted, const int64_t* __restrict__ perm,
                                     const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dtable,
                                     int64_t tokens, int dim, int vocab) {
    const int64_t b = blockIdx.x;
    const int64_t id = sorted[b];
    if (b > 0 && sorted[b - 1] == id) return;
    if (id < 0 || id >= vocab) return;
    for (int i = threadIdx.x; i < (dim >> 3); i += blockDim.x) {
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int64_t j = b; j < tokens && sorted[j] == id; ++j) {
            float f[8];
            unpack8(ld8(dout + perm[j] * dim + i * 8), f);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += f[k];
        }
        st8(dtable + id * dim + i * 8, pack8(acc));
    }
}

static inline int grid_for(int64_t work_items, int threads) {
    int64_t blocks = (work_items + threads - 1) / threads;
    const int64_t cap = (int64_t)kNumSMs * 16;
    return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

}  // namespace ftl

using namespace ftl;

FTL_API int ftl_rope_inplace(void* qkv, const float* cos_t, const float* sin_t, int64_t tokens, int seq_len,
                             int row_stride, int rot_cols, int head_dim, int inverse, cudaStream_t stream) {
    if (rot_cols % 8 || head_dim % 8 || row_stride % 8 || tokens <= 0) return (int)cudaErrorInvalidValue;
    const int64_t total = tokens * (rot_cols / 8);
    rope_kernel<<<grid_for(total, 256), 256, 0, stream>>>((__nv_bfloat16*)qkv, cos_t, sin_t, tokens, seq_len,
                                                          row_stride, rot_cols, head_dim, inverse ? -1.f : 1.f);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_swiglu_fwd(const void* gate, const void* up, void* act, int64_t n, cudaStream_t stream) {
    if (n % 8) return (int)cudaErrorInvalidValue;
    swiglu_fwd_kernel<<<grid_for(n / 8, 256), 256, 0, stream>>>((const __nv_bfloat16*)gate, (const __nv_bfloat16*)up,
                                                                (__nv_bfloat16*)act, n / 8);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_swiglu_bwd(const void* dact, const void* gate, const void* up, void* dgate, void* dup, int64_t n,
                           cudaStream_t stream) {
    if (n % 8) return (int)cudaErrorInvalidValue;
    swiglu_bwd_kernel<<<grid_for(n / 8, 256), 256, 0, stream>>>(
        (const __nv_bfloat16*)dact, (const __nv_bfloat16*)gate, (const __nv_bfloat16*)up, (__nv_bfloat16*)dgate,
        (__nv_bfloat16*)dup, n / 8);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_embedding_fwd(const int64_t* tokens, const void* table, void* out, int64_t ntok, int dim, int vocab,
                              cudaStream_t stream) {
    if (dim % 8 || ntok <= 0) return (int)cudaErrorInvalidValue;
    embedding_fwd_kernel<<<(unsigned)ntok, 128, 0, stream>>>(tokens, (const __nv_bfloat16*)table, (__nv_bfloat16*)out,
                                                             dim, vocab);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_embedding_bwd(const int64_t* sorted_tokens, const int64_t* perm, const void* dout, void* dtable,
                              int64_t ntok, int dim, int vocab, cudaStream_t stream) {
    if (dim % 8 || ntok <= 0) return (int)cudaErrorInvalidValue;
    embedding_bwd_kernel<<<(unsigned)ntok, 128, 0, stream>>>(sorted_tokens, perm, (const __nv_bfloat16*)dout,
                                                             (__nv_bfloat16*)dtable, ntok, dim, vocab);
    FTL_LAUNCH_CHECK();
    return 0;
}
