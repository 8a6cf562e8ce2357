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

// Embedding backward without a sort (replaces torch.sort + embedding_bwd_kernel + the index_fill that cleared the
// previous step's rows — the last library kernels of the step).  CTA t handles token position t:
//   * it is the OWNER of row id = tokens[t] iff no earlier position holds the same id; the owner adds the dout rows of
//     every position with that id in ascending position order (fp32, deterministic) and writes the row once;
//   * with `prev` (the ids of the previous step, whose rows are the only non-zero rows of the table gradient), CTAs
//     [T, 2T) clear the rows that the previous step wrote and this step does not write.
// The scan over positions is a block-cooperative ballot per 128 positions: T^2 / 128 compare rounds in total, a few
// microseconds at T = 2048.  sumsq (optional, [T]): sum of squares of the bf16 row the owner stored, 0 elsewhere.
constexpr int kEmbThreads = 128, kEmbVec = 4;

__global__ void __launch_bounds__(kEmbThreads)
embedding_bwd_direct_kernel(const int64_t* __restrict__ tokens, const int64_t* __restrict__ prev,
                            const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dtable, int T, int dim,
                            int vocab, float* __restrict__ sumsq) {
    __shared__ unsigned masks[kEmbThreads / 32];
    __shared__ float red[33];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nvec = dim >> 3;
    if ((int)blockIdx.x >= T) {
        // ---- clear a row that only the previous step touched
        const int t = blockIdx.x - T;
        const int64_t id = prev[t];
        if (id < 0 || id >= vocab) return;
        bool hit = false;
        for (int j = tid; j < t; j += kEmbThreads) hit |= prev[j] == id;        // an earlier position clears it
        for (int j = tid; j < T; j += kEmbThreads) hit |= tokens[j] == id;      // this step's owner overwrites it
        if (__syncthreads_or(hit)) return;
        bf16x8 z;
#pragma unroll
        for (int k = 0; k < 4; ++k) z.v[k] = __floats2bfloat162_rn(0.f, 0.f);
        for (int i = tid; i < nvec; i += kEmbThreads) st8(dtable + id * dim + i * 8, z);
        return;
    }
    const int t = blockIdx.x;
    const int64_t id = tokens[t];
    bool dup = id < 0 || id >= vocab;
    for (int j = tid; j < t; j += kEmbThreads) dup |= tokens[j] == id;
    if (__syncthreads_or(dup)) {
        if (sumsq && tid == 0) sumsq[t] = 0.f;
        return;
    }
    float ss = 0.f;
    for (int i0 = 0; i0 < nvec; i0 += kEmbVec * kEmbThreads) {
        float acc[kEmbVec][8];
#pragma unroll
        for (int k = 0; k < kEmbVec; ++k)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
        for (int c = (t / kEmbThreads) * kEmbThreads; c < T; c += kEmbThreads) {
            const int j = c + tid;
            const unsigned m = __ballot_sync(0xffffffffu, j >= t && j < T && tokens[j] == id);
            __syncthreads();                       // previous round's masks have been consumed
            if (lane == 0) masks[wid] = m;
            __syncthreads();
#pragma unroll
            for (int w = 0; w < kEmbThreads / 32; ++w) {
                unsigned mm = masks[w];
                while (mm) {
                    const int pos = c + w * 32 + __ffs(mm) - 1;
                    mm &= mm - 1;
#pragma unroll
                    for (int k = 0; k < kEmbVec; ++k) {
                        const int i = i0 + tid + k * kEmbThreads;
                        if (i < nvec) {
                            float f[8];
                            unpack8(ld8(dout + (int64_t)pos * dim + i * 8), f);
#pragma unroll
                            for (int q = 0; q < 8; ++q) acc[k][q] += f[q];
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < kEmbVec; ++k) {
            const int i = i0 + tid + k * kEmbThreads;
            if (i < nvec) {
                const bf16x8 o = pack8(acc[k]);
                st8(dtable + id * dim + i * 8, o);
                float f[8];
                unpack8(o, f);
#pragma unroll
                for (int q = 0; q < 8; ++q) ss += f[q] * f[q];
            }
        }
    }
    if (sumsq) {
        const float total = block_sum(ss, red);
        if (tid == 0) sumsq[t] = total;
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

// Sort-free deterministic embedding backward (see embedding_bwd_direct_kernel).  `prev_tokens` may be null (then the
// caller guarantees that every row this step does not write is already zero, e.g. after a memset).  sumsq: fp32 [ntok] or null.
FTL_API int ftl_embedding_bwd_direct(const int64_t* tokens, const int64_t* prev_tokens, const void* dout, void* dtable,
                                     int64_t ntok, int dim, int vocab, float* sumsq, cudaStream_t stream) {
    if (dim % 8 || ntok <= 0 || ntok > (1 << 24)) return (int)cudaErrorInvalidValue;
    const unsigned grid = (unsigned)(prev_tokens ? 2 * ntok : ntok);
    embedding_bwd_direct_kernel<<<grid, kEmbThreads, 0, stream>>>(tokens, prev_tokens, (const __nv_bfloat16*)dout,
                                                                  (__nv_bfloat16*)dtable, (int)ntok, dim, vocab, sumsq);
    FTL_LAUNCH_CHECK();
    return 0;
}
