// Multi-GPU data plane over NVLink 5 / NVSwitch peer memory (the reference has no
// communication code at all — SURVEY §2.3).  One process per GPU; every rank maps every peer's parameter and
// gradient buffers (CUDA IPC handles exchanged once over the torch.distributed control plane) and the kernels
// below read and write peer memory directly:
//
//   dp_reduce_scatter_kernel   rank r sums shard r of every rank's gradient buffer with 16-byte peer loads in
//                              fixed rank order (deterministic), writes the bf16 sum into its own buffer and
//                              produces sum-of-squares partials for the global gradient norm  (RS ⊕ norm²)
//   publish / finalize         every rank stores {shard sumsq, local loss} into a slot on every peer; after the
//                              barrier each rank folds the W slots in the same order -> identical norm and clip
//                              coefficient everywhere, no host round trip
//   dp_adamw_allgather_kernel  the sharded AdamW update of csrc/optim.cu fused with the parameter all-gather:
//                              each updated 16-byte vector is stored to the local copy and to every peer's
//                              copy in the same pass                                   (clip ⊕ AdamW ⊕ AG)
//   barrier_kernel             epoch-counter barrier on flags in peer memory (st.release.sys / ld.acquire.sys)
//                              with an abort word and a wall-clock budget: a failed or interrupted peer makes
//                              the others fall through with a status instead of spinning forever
#include "common/util.cuh"

namespace ftl {

constexpr int kMaxRanks = 8;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t gtimer() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ int4 ld_peer16(const void* p) {
    int4 r;
    asm volatile("ld.global.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// status word: 0 ok, 1 barrier timed out, 2 a peer raised the abort word, 3 the global gradient norm was not finite.
// Sticky until the host clears it: once it is non-zero NO later update is applied, so the device state stays that of the
// last good step however far ahead the host has already enqueued.
// flags layout per rank (uint32): [0..W) arrival epochs written by peers, [kMaxRanks] abort word.
__global__ void barrier_kernel(uint32_t* const* __restrict__ peer_flags, int rank, int world, uint32_t epoch,
                               uint32_t* __restrict__ status, uint64_t timeout_ns) {
    const int t = threadIdx.x;
    if (t >= world) return;
    __threadfence_system();                       // order this rank's earlier writes (incl. peer stores) first
    st_release_sys(peer_flags[t] + rank, epoch);  // tell rank t that `rank` has arrived
    const uint32_t* mine = peer_flags[rank];
    const uint64_t t0 = gtimer();
    uint32_t polls = 0;
    while ((int32_t)(ld_acquire_sys(mine + t) - epoch) < 0) {
        if ((++polls & 0x3ff) == 0) {
            if (ld_acquire_sys(mine + kMaxRanks) != 0) { atomicMax(status, 2u); return; }
            if (gtimer() - t0 > timeout_ns) { atomicMax(status, 1u); return; }
        }
    }
}

__global__ void raise_abort_kernel(uint32_t* const* __restrict__ peer_flags, int world) {
    const int t = threadIdx.x;
    if (t < world) st_release_sys(peer_flags[t] + kMaxRanks, 1u);
}

template <int W>
__global__ void __launch_bounds__(512)
dp_reduce_scatter_kernel(__nv_bfloat16* const* __restrict__ peer_g, int rank, int64_t lo_vec, int64_t nvec,
                         float* __restrict__ partial, const uint32_t* __restrict__ status,
                         __nv_bfloat16* __restrict__ dst_compact) {
    __shared__ float red[33];
    float acc = 0.f;
    if (*status == 0) {
        const __nv_bfloat16* src[W];
#pragma unroll
        for (int r = 0; r < W; ++r) src[r] = peer_g[r] + lo_vec * 8;
        // in place (this rank's own range of its gradient buffer) or into a compact shard buffer (fully-sharded mode
        // with gradient slots that the backward reuses)
        __nv_bfloat16* dst = dst_compact ? dst_compact : peer_g[rank] + lo_vec * 8;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
            int4 raw[W];
#pragma unroll
            for (int r = 0; r < W; ++r) raw[r] = ld_peer16(src[r] + i * 8);   // W independent loads in flight
            float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int r = 0; r < W; ++r) {                                       // fixed order: rank 0, 1, ...
                float f[8];
                unpack8(*reinterpret_cast<bf16x8*>(&raw[r]), f);
#pragma unroll
                for (int j = 0; j < 8; ++j) s[j] += f[j];
            }
            const bf16x8 o = pack8(s);
            if (W > 1 || dst_compact) st8(dst + i * 8, o);   // a single rank already holds the sum in place
            float f[8];
            unpack8(o, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
        }
    }
    const float total = block_sum(acc, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}

// slots layout per rank (fp32): [2*r] = sumsq of rank r's shard, [2*r+1] = rank r's local loss
__global__ void publish_kernel(const float* __restrict__ partial, int nparts, const float* __restrict__ loss_local,
                               float* const* __restrict__ peer_slots, int rank, int world) {
    __shared__ float red[33];
    float s = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) s += partial[i];
    const float total = block_sum(s, red);
    if (threadIdx.x < world) {
        float* dst = peer_slots[threadIdx.x] + 2 * rank;
        dst[0] = total;
        dst[1] = loss_local ? loss_local[0] : 0.f;
        __threadfence_system();
    }
}

__global__ void dp_finalize_kernel(const float* __restrict__ slots, int world, const float* __restrict__ hyper,
                                   float* __restrict__ stats, float* __restrict__ loss_out, int clip,
                                   uint32_t* __restrict__ status, float* __restrict__ result) {
    if (threadIdx.x != 0) return;
    float total = 0.f, loss = 0.f;
    for (int r = 0; r < world; ++r) {   // same order on every rank -> bitwise-identical norm everywhere
        total += slots[2 * r];
        loss += slots[2 * r + 1];
    }
    const float norm = sqrtf(total);
    const bool finite = isfinite(norm);
    float coef = 1.f;
    if (clip && finite) coef = fminf(1.f, hyper[7] / (norm + 1e-6f));
    stats[0] = norm; stats[1] = coef; stats[2] = finite ? 1.f : 0.f; stats[3] = total;
    if (loss_out) loss_out[0] = loss;
    // every rank folds the same slots in the same order, so all of them gate the same step (reference: the step that
    // meets a non-finite norm raises before optimizer.step(), utils.py:61, train.py:104-107)
    if (!finite && status) atomicMax(status, 3u);
    // the step's whole report in one place, so that the host needs ONE device-to-host copy and no packing kernels
    if (result) {
        result[0] = loss; result[1] = norm; result[2] = coef; result[3] = finite ? 1.f : 0.f;
        result[4] = status ? (float)*status : 0.f;
    }
}

template <int W>
__global__ void __launch_bounds__(512)
dp_adamw_allgather_kernel(__nv_bfloat16* const* __restrict__ peer_p, const __nv_bfloat16* __restrict__ g,
                          __nv_bfloat16* __restrict__ m, __nv_bfloat16* __restrict__ v, int rank, int64_t lo_vec,
                          int64_t nvec, const float* __restrict__ hyper, const float* __restrict__ stats,
                          const uint32_t* __restrict__ status, unsigned long long* __restrict__ commits) {
    if (*status != 0 || stats[2] == 0.f) return;   // aborted step or non-finite norm: leave every copy untouched
    if (commits && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(commits, 1ull);   // this range's update is applied
    const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
    const float bc1 = hyper[5], bc2_sqrt = hyper[6], coef = stats[1];
    const float step_size = lr / bc1, decay = 1.f - lr * wd;
    const float vcut = adam_vcut(eps, bc2_sqrt);
    __nv_bfloat16* dst[W];
#pragma unroll
    for (int r = 0; r < W; ++r) dst[r] = peer_p[r] + lo_vec * 8;
    const __nv_bfloat16* p = peer_p[rank] + lo_vec * 8;
    g += lo_vec * 8; m += lo_vec * 8; v += lo_vec * 8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float pf[8], gf[8], mf[8], vf[8];
        unpack8(ld8_stream(p + i * 8), pf);
        unpack8(ld8_stream(g + i * 8), gf);
        unpack8(ld8_stream(m + i * 8), mf);
        unpack8(ld8_stream(v + i * 8), vf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gr = gf[j] * coef;
            mf[j] = mf[j] + (1.f - b1) * (gr - mf[j]);
            vf[j] = b2 * vf[j] + (1.f - b2) * gr * gr;
            const float denom = adam_denom(vf[j], bc2_sqrt, eps, vcut);
            pf[j] = pf[j] * decay - step_size * adam_quot(mf[j], denom);
        }
        const bf16x8 np = pack8(pf);
        st8_stream(m + i * 8, pack8(mf));
        st8_stream(v + i * 8, pack8(vf));
#pragma unroll
        for (int r = 0; r < W; ++r) st8_stream(dst[r] + i * 8, np);   // local copy + W-1 peer copies (all-gather)
    }
}

// Plain parameter all-gather (N13): dst[r * n : (r+1) * n] = owner r's shard (n elements at `off` in its compact
// buffer), for r = 0..W-1, through 16-byte peer loads.  Used for the operands whose gather is not fused into a GEMM.
template <int W>
__global__ void __launch_bounds__(512)
fsdp_gather_kernel(const __nv_bfloat16* const* __restrict__ peer_shards, int64_t off, __nv_bfloat16* __restrict__ dst,
                   int64_t nvec, int rank, const uint32_t* __restrict__ status) {
    if (*status != 0) return;
    // four independent peer loads in flight per thread (NVLink round trips are ~3 us: bytes in flight are what set the rate)
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, total = nvec * W;
    for (int64_t i0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i0 < total; i0 += 4 * stride) {
        int4 v[4];
        int64_t d[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t i = i0 + u * stride;
            d[u] = -1;
            if (i < total) {
                const int rr = (int)(i / nvec);
                const int r = (rr + rank) % W;                     // start with the own shard, peers rotate
                const int64_t k = i - (int64_t)rr * nvec;
                v[u] = ld_peer16(peer_shards[r] + off + k * 8);
                d[u] = ((int64_t)r * nvec + k) * 8;
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (d[u] >= 0) *reinterpret_cast<int4*>(dst + d[u]) = v[u];
    }
}

// AdamW on one compact shard (FSDP): p, m, v are shard-local, g points at the reduced gradient range.
__global__ void __launch_bounds__(512)
adamw_shard_kernel(__nv_bfloat16* __restrict__ p, const __nv_bfloat16* __restrict__ g, __nv_bfloat16* __restrict__ m,
                   __nv_bfloat16* __restrict__ v, int64_t nvec, const float* __restrict__ hyper,
                   const float* __restrict__ stats, const uint32_t* __restrict__ status,
                   unsigned long long* __restrict__ commits) {
    if (*status != 0 || stats[2] == 0.f) return;
    if (commits && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(commits, 1ull);
    const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
    const float bc1 = hyper[5], bc2_sqrt = hyper[6], coef = stats[1];
    const float step_size = lr / bc1, decay = 1.f - lr * wd;
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
            mf[j] = mf[j] + (1.f - b1) * (gr - mf[j]);
            vf[j] = b2 * vf[j] + (1.f - b2) * gr * gr;
            const float denom = adam_denom(vf[j], bc2_sqrt, eps, vcut);
            pf[j] = pf[j] * decay - step_size * adam_quot(mf[j], denom);
        }
        st8_stream(p + i * 8, pack8(pf));
        st8_stream(m + i * 8, pack8(mf));
        st8_stream(v + i * 8, pack8(vf));
    }
}

// ---- embedding bucket: sparse reduce-scatter --------------------------------------------------------------------------
// The table gradient is 1 GB per rank of which only the rows of the step's tokens (<= T per rank) are non-zero, and it is
// the LAST gradient of the backward: its dense reduce-scatter (every rank pulling W-1 shards of zeros, 1.5 ms at W = 8
// after a 0.35 ms memset) sat between the end of one step and the first kernel of the next.  Here every rank first
// publishes its token ids to all peers (emb_tokens_push_kernel, at the start of the backward); after the barrier the
// owner of a row range reduces only the rows that occur in ANY rank's list: the CTA of the first occurrence of an id (in
// rank-major, position-minor order) adds the rows of the ranks that hold the id, in rank order — bit-identical to the
// dense fixed-order sum, since all other addends are +0 — and writes the bf16 total and a sum-of-squares partial.
// CTAs [W*T, 2*W*T) clear the rows the PREVIOUS step's union wrote in this rank's range and this step's does not.
// toks: [W][cap] int64 (this rank's copy of everybody's ids); [lo, hi): this rank's element range of the flat gradient.
__global__ void emb_tokens_push_kernel(const int64_t* __restrict__ tokens, int64_t* const* __restrict__ peer_toks,
                                       int64_t slot_off, int rank, int T, int cap) {
    int64_t* dst = peer_toks[blockIdx.x] + slot_off + (int64_t)rank * cap;
    for (int i = threadIdx.x; i < T; i += blockDim.x) dst[i] = tokens[i];
    __threadfence_system();
}

template <int W>
__global__ void __launch_bounds__(128)
emb_sparse_reduce_kernel(__nv_bfloat16* const* __restrict__ peer_g, int rank, const int64_t* __restrict__ toks,
                         const int64_t* __restrict__ toks_prev, int T, int cap, int64_t emb_off, int dim, int vocab,
                         int64_t lo, int64_t hi, float* __restrict__ partial, const uint32_t* __restrict__ status) {
    __shared__ float red[33];
    const int tid = threadIdx.x;
    const int WT = W * T;
    const bool clearing = (int)blockIdx.x >= WT;
    const int idx = clearing ? blockIdx.x - WT : blockIdx.x;
    const int q = idx / T, t = idx - q * T;
    if (*status != 0) {
        if (!clearing && tid == 0) partial[idx] = 0.f;
        return;
    }
    const int64_t* list = clearing ? toks_prev : toks;
    const int64_t id = list[(int64_t)q * cap + t];
    // this rank's part of the row
    const int64_t r0 = emb_off + id * dim;
    const int64_t e0 = r0 > lo ? r0 : lo, e1 = (r0 + dim) < hi ? (r0 + dim) : hi;
    bool skip = id < 0 || id >= vocab || e0 >= e1;
    // first occurrence in (rank, position) order?
    if (!skip) {
        for (int qq = 0; qq <= q; ++qq) {
            const int n = qq < q ? T : t;
            for (int j = tid; j < n; j += blockDim.x) skip |= list[(int64_t)qq * cap + j] == id;
        }
    }
    if (clearing && !skip)   // rows of the current union are (over)written by their owner CTA below
        for (int j = tid; j < WT; j += blockDim.x) skip |= toks[(int64_t)(j / T) * cap + (j % T)] == id;
    if (__syncthreads_or(skip)) {
        if (!clearing && tid == 0) partial[idx] = 0.f;
        return;
    }
    __nv_bfloat16* mine = peer_g[rank];
    const int64_t v0 = e0 >> 3, v1 = e1 >> 3;          // ranges are multiples of 8 elements
    if (clearing) {
        bf16x8 z;
#pragma unroll
        for (int k = 0; k < 4; ++k) z.v[k] = __floats2bfloat162_rn(0.f, 0.f);
        for (int64_t i = v0 + tid; i < v1; i += blockDim.x) st8(mine + i * 8, z);
        return;
    }
    bool has[W];
#pragma unroll
    for (int qq = 0; qq < W; ++qq) {
        bool h = false;
        for (int j = tid; j < T; j += blockDim.x) h |= toks[(int64_t)qq * cap + j] == id;
        has[qq] = __syncthreads_or(h);
    }
    float ss = 0.f;
    for (int64_t i = v0 + tid; i < v1; i += blockDim.x) {
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int qq = 0; qq < W; ++qq) {               // fixed order: rank 0, 1, ...
            if (has[qq]) {
                const int4 raw = ld_peer16(peer_g[qq] + i * 8);
                float f[8];
                unpack8(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
                for (int j = 0; j < 8; ++j) s[j] += f[j];
            }
        }
        const bf16x8 o = pack8(s);
        st8(mine + i * 8, o);
        float f[8];
        unpack8(o, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += f[j] * f[j];
    }
    const float total = block_sum(ss, red);
    if (tid == 0) partial[idx] = total;
}

static inline int comm_grid(int64_t nvec, int max_ctas) {
    int64_t blocks = (nvec + 511) / 512;
    int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)kNumSMs * 4;
    return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

}  // namespace ftl

using namespace ftl;

// ---- symmetric buffers: cudaMalloc + CUDA IPC ------------------------------------------------------------
FTL_API int ftl_symm_alloc(void** ptr, int64_t bytes) {   // allocates (set-up time only)
    cudaError_t e = cudaMalloc(ptr, (size_t)bytes);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaMemset(*ptr, 0, (size_t)bytes);
}
FTL_API int ftl_symm_free(void* ptr) { return (int)cudaFree(ptr); }
FTL_API int ftl_ipc_export(void* ptr, void* handle64) {
    return (int)cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle64), ptr);
}
FTL_API int ftl_ipc_import(const void* handle64, void** ptr) {
    cudaIpcMemHandle_t h = *reinterpret_cast<const cudaIpcMemHandle_t*>(handle64);
    return (int)cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}
FTL_API int ftl_ipc_close(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }
FTL_API int ftl_comm_max_ranks(void) { return kMaxRanks; }
FTL_API int ftl_comm_flag_words(void) { return kMaxRanks + 8; }

// ---- kernels ---------------------------------------------------------------------------------------------
FTL_API int ftl_barrier(void* const* peer_flags, int rank, int world, uint32_t epoch, void* status, int64_t timeout_ms,
                        cudaStream_t stream) {
    if (world < 1 || world > kMaxRanks) return (int)cudaErrorInvalidValue;
    barrier_kernel<<<1, 32, 0, stream>>>((uint32_t* const*)peer_flags, rank, world, epoch, (uint32_t*)status,
                                         (uint64_t)timeout_ms * 1000000ull);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_raise_abort(void* const* peer_flags, int world, cudaStream_t stream) {
    raise_abort_kernel<<<1, 32, 0, stream>>>((uint32_t* const*)peer_flags, world);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_dp_partials(int max_ctas) { return max_ctas > 0 ? max_ctas : kNumSMs * 4; }

// Reduce shard [lo, lo+n) (elements, multiples of 8) of every rank's gradient buffer into this rank's buffer.
FTL_API int ftl_dp_reduce_scatter(void* const* peer_g, int rank, int world, int64_t lo, int64_t n, float* partial,
                                  const void* status, int max_ctas, cudaStream_t stream) {
    if (lo % 8 || n % 8 || n <= 0) return (int)cudaErrorInvalidValue;
    const int grid = ftl_dp_partials(max_ctas);
    auto P = (__nv_bfloat16* const*)peer_g;
    auto S = (const uint32_t*)status;
#define RS_CASE(W) case W: dp_reduce_scatter_kernel<W><<<grid, 512, 0, stream>>>(P, rank, lo / 8, n / 8, partial, S, nullptr); break;
    switch (world) { RS_CASE(1) RS_CASE(2) RS_CASE(3) RS_CASE(4) RS_CASE(5) RS_CASE(6) RS_CASE(7) RS_CASE(8)
        default: return (int)cudaErrorInvalidValue; }
#undef RS_CASE
    FTL_LAUNCH_CHECK();
    return 0;
}

// Same reduction, but the bf16 sum of [lo, lo+n) goes to dst[0, n) (a compact shard buffer) instead of in place.
FTL_API int ftl_dp_reduce_scatter_to(void* const* peer_g, int rank, int world, int64_t lo, int64_t n, void* dst,
                                     float* partial, const void* status, int max_ctas, cudaStream_t stream) {
    if (lo % 8 || n % 8 || n <= 0 || !dst || (reinterpret_cast<uintptr_t>(dst) & 15)) return (int)cudaErrorInvalidValue;
    const int grid = ftl_dp_partials(max_ctas);
    auto P = (__nv_bfloat16* const*)peer_g;
    auto S = (const uint32_t*)status;
#define RS_CASE(W) case W: dp_reduce_scatter_kernel<W><<<grid, 512, 0, stream>>>(P, rank, lo / 8, n / 8, partial, S, \
        (__nv_bfloat16*)dst); break;
    switch (world) { RS_CASE(1) RS_CASE(2) RS_CASE(3) RS_CASE(4) RS_CASE(5) RS_CASE(6) RS_CASE(7) RS_CASE(8)
        default: return (int)cudaErrorInvalidValue; }
#undef RS_CASE
    FTL_LAUNCH_CHECK();
    return 0;
}

// Publish this rank's token ids of the step into slot `slot_off` (elements) of every peer's id buffer ([.][W][cap] int64).
FTL_API int ftl_emb_tokens_push(const int64_t* tokens, void* const* peer_toks, int64_t slot_off, int rank, int world, int T,
                                int cap, cudaStream_t stream) {
    if (world < 1 || world > kMaxRanks || T <= 0 || T > cap) return (int)cudaErrorInvalidValue;
    emb_tokens_push_kernel<<<world, 256, 0, stream>>>(tokens, (int64_t* const*)peer_toks, slot_off, rank, T, cap);
    FTL_LAUNCH_CHECK();
    return 0;
}

// Sparse reduce-scatter of the embedding-table gradient over this rank's element range [lo, hi) of the flat gradient
// (see emb_sparse_reduce_kernel).  toks / toks_prev: this rank's [W][cap] id lists of this / the previous step (prev may
// be NULL on the first step, when the range is still all zero).  partial: fp32 [world * T].
FTL_API int ftl_emb_sparse_reduce(void* const* peer_g, int rank, int world, const int64_t* toks, const int64_t* toks_prev,
                                  int T, int cap, int64_t emb_off, int dim, int vocab, int64_t lo, int64_t hi,
                                  float* partial, const void* status, cudaStream_t stream) {
    if (lo % 8 || hi % 8 || emb_off % 8 || dim % 8 || T <= 0 || T > cap || lo >= hi) return (int)cudaErrorInvalidValue;
    const unsigned grid = (unsigned)((toks_prev ? 2 : 1) * world * T);
    auto P = (__nv_bfloat16* const*)peer_g;
    auto S = (const uint32_t*)status;
#define ESR_CASE(W) case W: emb_sparse_reduce_kernel<W><<<grid, 128, 0, stream>>>(P, rank, toks, toks_prev, T, cap, emb_off, \
        dim, vocab, lo, hi, partial, S); break;
    switch (world) { ESR_CASE(1) ESR_CASE(2) ESR_CASE(3) ESR_CASE(4) ESR_CASE(5) ESR_CASE(6) ESR_CASE(7) ESR_CASE(8)
        default: return (int)cudaErrorInvalidValue; }
#undef ESR_CASE
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_dp_publish(const float* partial, int nparts, const float* loss_local, void* const* peer_slots, int rank,
                           int world, cudaStream_t stream) {
    publish_kernel<<<1, 1024, 0, stream>>>(partial, nparts, loss_local, (float* const*)peer_slots, rank, world);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_dp_finalize(const float* slots, int world, const float* hyper, float* stats, float* loss_out, int clip,
                            void* status, float* result, cudaStream_t stream) {
    dp_finalize_kernel<<<1, 32, 0, stream>>>(slots, world, hyper, stats, loss_out, clip, (uint32_t*)status, result);
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_dp_adamw_allgather(void* const* peer_p, const void* g, void* m, void* v, int rank, int world, int64_t lo,
                                   int64_t n, const float* hyper, const float* stats, const void* status, void* commits,
                                   int max_ctas, cudaStream_t stream) {
    if (lo % 8 || n % 8 || n <= 0) return (int)cudaErrorInvalidValue;
    const int grid = comm_grid(n / 8, max_ctas);
    auto P = (__nv_bfloat16* const*)peer_p;
    auto S = (const uint32_t*)status;
#define AG_CASE(W) case W: dp_adamw_allgather_kernel<W><<<grid, 512, 0, stream>>>(P, (const __nv_bfloat16*)g, \
        (__nv_bfloat16*)m, (__nv_bfloat16*)v, rank, lo / 8, n / 8, hyper, stats, S, (unsigned long long*)commits); break;
    switch (world) { AG_CASE(1) AG_CASE(2) AG_CASE(3) AG_CASE(4) AG_CASE(5) AG_CASE(6) AG_CASE(7) AG_CASE(8)
        default: return (int)cudaErrorInvalidValue; }
#undef AG_CASE
    FTL_LAUNCH_CHECK();
    return 0;
}


FTL_API int ftl_fsdp_gather(const void* const* peer_shards, int64_t off, void* dst, int64_t n_per_rank, int rank, int world,
                            const void* status, int max_ctas, cudaStream_t stream) {
    if (off % 8 || n_per_rank % 8 || n_per_rank <= 0) return (int)cudaErrorInvalidValue;
    const int grid = comm_grid(n_per_rank / 8 * world, max_ctas);
    auto P = (const __nv_bfloat16* const*)peer_shards;
    auto S = (const uint32_t*)status;
#define FG_CASE(W) case W: fsdp_gather_kernel<W><<<grid, 512, 0, stream>>>(P, off, (__nv_bfloat16*)dst, n_per_rank / 8, rank, S); break;
    switch (world) { FG_CASE(1) FG_CASE(2) FG_CASE(3) FG_CASE(4) FG_CASE(5) FG_CASE(6) FG_CASE(7) FG_CASE(8)
        default: return (int)cudaErrorInvalidValue; }
#undef FG_CASE
    FTL_LAUNCH_CHECK();
    return 0;
}

FTL_API int ftl_adamw_shard(void* p, const void* g, void* m, void* v, int64_t n, const float* hyper, const float* stats,
                            const void* status, void* commits, int max_ctas, cudaStream_t stream) {
    if (n % 8 || n <= 0) return (int)cudaErrorInvalidValue;
    adamw_shard_kernel<<<comm_grid(n / 8, max_ctas), 512, 0, stream>>>(
        (__nv_bfloat16*)p, (const __nv_bfloat16*)g, (__nv_bfloat16*)m, (__nv_bfloat16*)v, n / 8, hyper, stats,
        (const uint32_t*)status, (unsigned long long*)commits);
    FTL_LAUNCH_CHECK();
    return 0;
}
