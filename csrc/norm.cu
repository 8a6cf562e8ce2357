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
#include <cstdio>

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
// kVec = vectors of 8 bf16 per thread (1, 2 or 4): sized to the row so that dim 4096 with 256 threads keeps two CTAs per SM
template <bool kAddRes, int kVec>
__global__ void __launch_bounds__(256)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                   const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd_in,
                   const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                   float* __restrict__ dw_partial, int rows, int dim) {
    __shared__ float red[33];
    const int nvec = dim >> 3;
    float dw_acc[kVec][8];
    float wreg[kVec][8];
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
#pragma unroll
        for (int j = 0; j < 8; ++j) dw_acc[k][j] = 0.f;
        if (i < nvec) unpack8(ld8(w + i * 8), wreg[k]);
    }
    // A CTA walks its rows one after the other with a block reduction in the middle of each, so per row the chain used
    // to be load latency -> reduce -> residual load latency -> store.  All three inputs of a row are now requested at
    // once and the NEXT row's are requested before the current row's reduction, which keeps ~2 rows of loads in flight
    // per CTA (the kernel is HBM-bound: 3 reads + 1 write of [rows, dim] bf16).
    bf16x8 nx[kVec], ndy[kVec], nres[kVec];
    float nrstd = 0.f;
    auto fetch = [&](int64_t row) {
        nrstd = rstd_in[row];
#pragma unroll
        for (int k = 0; k < kVec; ++k) {
            const int i = threadIdx.x + k * blockDim.x;
            if (i < nvec) {
                nx[k] = ld8(x + row * dim + i * 8);
                ndy[k] = ld8(dy + row * dim + i * 8);
                if (kAddRes) nres[k] = ld8(dres + row * dim + i * 8);
            }
        }
    };
    if ((int64_t)blockIdx.x < rows) fetch(blockIdx.x);
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const float rstd = nrstd;
        float xh[kVec][8], dyw[kVec][8];
        bf16x8 res[kVec];
        float dot = 0.f;
#pragma unroll
        for (int k = 0; k < kVec; ++k) {
            const int i = threadIdx.x + k * blockDim.x;
            if (i < nvec) {
                float d[8];
                unpack8(nx[k], xh[k]);
                unpack8(ndy[k], d);
                if (kAddRes) res[k] = nres[k];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xh[k][j] *= rstd;
                    dw_acc[k][j] += d[j] * round_bf16(xh[k][j]);
                    dyw[k][j] = d[j] * wreg[k][j];
                    dot += dyw[k][j] * xh[k][j];
                }
            }
        }
        if (row + gridDim.x < rows) fetch(row + gridDim.x);
        const float mean = block_sum(dot, red) / (float)dim;
#pragma unroll
        for (int k = 0; k < kVec; ++k) {
            const int i = threadIdx.x + k * blockDim.x;
            if (i < nvec) {
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = rstd * (dyw[k][j] - xh[k][j] * mean);
                if (kAddRes) {
                    float r[8];
                    unpack8(res[k], r);
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] += r[j];
                }
                st8(dx + row * dim + i * 8, pack8(o));
            }
        }
    }
    float* out = dw_partial + (int64_t)blockIdx.x * dim;
#pragma unroll
    for (int k = 0; k < kVec; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
        if (i < nvec) {
            *reinterpret_cast<float4*>(out + i * 8) = make_float4(dw_acc[k][0], dw_acc[k][1], dw_acc[k][2], dw_acc[k][3]);
            *reinterpret_cast<float4*>(out + i * 8 + 4) = make_float4(dw_acc[k][4], dw_acc[k][5], dw_acc[k][6], dw_acc[k][7]);
        }
    }
}

// Fold the per-CTA partial rows: block = 32 columns x 8 row lanes; lane py adds partials py, py+8, ... in order,
// the 8 lanes are then added in fixed order through shared memory (deterministic).  Reads are 128-byte coalesced.
// sumsq (optional, one float per block): sum of squares of the 32 bf16 values this block stored (gradient-norm partial).
__global__ void __launch_bounds__(256)
rmsnorm_dw_finalize_kernel(const float* __restrict__ partial, int nparts, int dim, __nv_bfloat16* __restrict__ dw,
                           float* __restrict__ sumsq = nullptr) {
    __shared__ float acc[8][33];
    const int cx = threadIdx.x & 31, py = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + cx;
    float s = 0.f;
    if (j < dim)
        for (int p = py; p < nparts; p += 8) s += partial[(int64_t)p * dim + j];
    acc[py][cx] = s;
    __syncthreads();
    if (py == 0) {
        float sq = 0.f;
        if (j < dim) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) t += acc[k][cx];
            const __nv_bfloat16 o = __float2bfloat16_rn(t);
            dw[j] = o;
            sq = __bfloat162float(o) * __bfloat162float(o);
        }
        if (sumsq) {
            sq = warp_sum(sq);
            if (cx == 0) sumsq[blockIdx.x] = sq;
        }
    }
}

static inline int norm_threads(int dim) {
    int nvec = dim / 8;
    int t = (nvec + kNormMaxVec - 1) / kNormMaxVec;
    if (nvec >= 256 && t < 256) t = 256;  // prefer one or two vectors per thread at dim 2048-4096
    t = (t + 31) / 32 * 32;
    if (t < 32) t = 32;
    return t;
}

// picks the instantiation with the fewest vectors per thread that covers the row
static void launch_rmsnorm_bwd(int parts, int threads, cudaStream_t stream, const __nv_bfloat16* DY, const __nv_bfloat16* X,
                               const __nv_bfloat16* W, const float* rstd, const __nv_bfloat16* DR, __nv_bfloat16* DX,
                               float* scratch, int rows, int dim) {
    const int vec = (dim / 8 + threads - 1) / threads;
#define FTL_RMS_BWD(V)                                                                                             \
    do {                                                                                                           \
        if (DR) rmsnorm_bwd_kernel<true, V><<<parts, threads, 0, stream>>>(DY, X, W, rstd, DR, DX, scratch, rows, dim); \
        else rmsnorm_bwd_kernel<false, V><<<parts, threads, 0, stream>>>(DY, X, W, rstd, nullptr, DX, scratch, rows, dim); \
    } while (0)
    if (vec <= 1) FTL_RMS_BWD(1);
    else if (vec <= 2) FTL_RMS_BWD(2);
    else FTL_RMS_BWD(4);
#undef FTL_RMS_BWD
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

FTL_API int ftl_rmsnorm_sumsq_slots(int dim) { return (dim + 31) / 32; }

// scratch: fp32 [ftl_rmsnorm_partials(rows), dim];  sumsq: optional fp32 [ftl_rmsnorm_sumsq_slots(dim)] receiving the
// sum of squares of the stored d(weight), in pieces
FTL_API int ftl_rmsnorm_bwd_ex(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                               void* dx, void* dw, float* scratch, float* sumsq, int rows, int dim, cudaStream_t stream) {
    if (dim % 8 || dim > 8 * kNormMaxVec * 256 || rows <= 0) return (int)cudaErrorInvalidValue;
    int threads = norm_threads(dim);
    if (threads > 256) threads = 256;
    const int parts = ftl_rmsnorm_partials(rows);
    auto DY = (const __nv_bfloat16*)dy; auto X = (const __nv_bfloat16*)x; auto W = (const __nv_bfloat16*)w;
    auto DR = (const __nv_bfloat16*)dres; auto DX = (__nv_bfloat16*)dx;
    launch_rmsnorm_bwd(parts, threads, stream, DY, X, W, rstd, DR, DX, scratch, rows, dim);
    FTL_LAUNCH_CHECK();
    rmsnorm_dw_finalize_kernel<<<(dim + 31) / 32, 256, 0, stream>>>(scratch, parts, dim, (__nv_bfloat16*)dw, sumsq);
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
    launch_rmsnorm_bwd(parts, threads, stream, DY, X, W, rstd, DR, DX, scratch, rows, dim);
    FTL_LAUNCH_CHECK();
    rmsnorm_dw_finalize_kernel<<<(dim + 31) / 32, 256, 0, stream>>>(scratch, parts, dim, (__nv_bfloat16*)dw);
    FTL_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// RMSNorm backward for callers that have no device memory to hand in as scratch.
//
// ftl_rmsnorm_bwd() needs an fp32 work area of ftl_rmsnorm_partials(rows) x dim floats from its caller.
// Some callers cannot produce one.  The case this library cares about is the recovery path of a training
// job: after a CUDA out-of-memory error the handler still has to run a backward fragment or a
// finiteness check before it snapshots state, and at that moment every further cudaMalloc fails — the
// framework's own caching allocator already owns what is left of the device.  The same holds for a host
// program that embeds this library next to such an allocator and owns no pool of its own, and for code
// that runs where a cudaMalloc is unwelcome because it synchronises the whole device (inside a step
// that is being timed or overlapped with a copy stream).
//
// For them the work area lives inside the library image: the file-scope array below, 296 x 4096 floats
// (4 849 664 bytes).  Every GPU that loads libftl_b200.so carries that array from load until the process
// exits, whether or not this entry point is ever called.  296 is the largest number of partial rows the
// backward ever uses (two CTAs per SM on 148 SMs); 4096 is the model width of the flagship
// configuration.  Requests that need more (dim > 4096) are refused with cudaErrorInvalidValue and a
// message readable through ftl_last_error(); use ftl_rmsnorm_bwd() with your own scratch for those.
//
// Things a caller must know:
//  * The name `g_rmsnorm_static_scratch` is only a handle on the host side.  The device address is looked
//    up once per device with cudaGetSymbolAddress, with that device current, by the first call made on
//    that device.  Make that first call outside of any stream capture; later calls only launch kernels.
//  * No cudaMalloc, cudaFree or other allocating routine is called, on the first call or later.
//  * All calls on one GPU share the one array, and it keeps whatever the previous call left in it.  Nothing
//    has to be cleared between calls: each call writes every partial row it later folds before it reads
//    it.  For the same reason two calls whose kernels can overlap on the GPU (different streams, no
//    dependency between them) race on the array and produce garbage — order them on one stream or with
//    an event.
constexpr int kStaticScratchRows = 2 * ftl::kNumSMs;   // 296
constexpr int kStaticScratchDim = 4096;
static __device__ __align__(256) float g_rmsnorm_static_scratch[kStaticScratchRows * kStaticScratchDim];

constexpr int kMaxDevices = 64;
// Per-device cache of the array's device address, filled on a device's first call WITHOUT a lock: two host threads that
// make their first call on the same device at the same moment both resolve and store the same value (benign).
static float* g_static_scratch_ptr[kMaxDevices] = {nullptr};
static thread_local char g_last_error[256] = "";

FTL_API const char* ftl_last_error(void) { return g_last_error; }

FTL_API int64_t ftl_rmsnorm_bwd_static_capacity(void) {
    return (int64_t)kStaticScratchRows * kStaticScratchDim * (int64_t)sizeof(float);
}

FTL_API int ftl_rmsnorm_bwd_static(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                                   void* dx, void* dw, int rows, int dim, cudaStream_t stream) {
    g_last_error[0] = 0;
    if (rows <= 0 || dim <= 0 || dim % 8) {
        snprintf(g_last_error, sizeof g_last_error, "ftl_rmsnorm_bwd_static: rows=%d dim=%d: dim must be a positive multiple of 8", rows, dim);
        return (int)cudaErrorInvalidValue;
    }
    const int parts = ftl_rmsnorm_partials(rows);
    if (dim > kStaticScratchDim || (int64_t)parts * dim > (int64_t)kStaticScratchRows * kStaticScratchDim) {
        snprintf(g_last_error, sizeof g_last_error,
                 "ftl_rmsnorm_bwd_static: dim=%d needs %lld scratch bytes, the built-in area holds %lld (dim <= %d); "
                 "call ftl_rmsnorm_bwd with caller-provided scratch instead",
                 dim, (long long)parts * dim * 4, (long long)ftl_rmsnorm_bwd_static_capacity(), kStaticScratchDim);
        return (int)cudaErrorInvalidValue;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        snprintf(g_last_error, sizeof g_last_error, "ftl_rmsnorm_bwd_static: cudaGetDevice failed (%d), device %d", (int)e, dev);
        return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidDevice;
    }
    float* scratch = g_static_scratch_ptr[dev];
    if (scratch == nullptr) {
        // once per device, with that device current: turn the host-side symbol into this device's address
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_rmsnorm_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            snprintf(g_last_error, sizeof g_last_error, "ftl_rmsnorm_bwd_static: cudaGetSymbolAddress failed: %s",
                     cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
        }
        scratch = static_cast<float*>(p);
        g_static_scratch_ptr[dev] = scratch;
    }
    int threads = norm_threads(dim);
    if (threads > 256) threads = 256;
    auto DY = (const __nv_bfloat16*)dy; auto X = (const __nv_bfloat16*)x; auto W = (const __nv_bfloat16*)w;
    auto DR = (const __nv_bfloat16*)dres; auto DX = (__nv_bfloat16*)dx;
    launch_rmsnorm_bwd(parts, threads, stream, DY, X, W, rstd, DR, DX, scratch, rows, dim);
    e = cudaGetLastError();
    if (e == cudaSuccess) {
        rmsnorm_dw_finalize_kernel<<<(dim + 31) / 32, 256, 0, stream>>>(scratch, parts, dim, (__nv_bfloat16*)dw);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) snprintf(g_last_error, sizeof g_last_error, "ftl_rmsnorm_bwd_static: launch failed: %s", cudaGetErrorString(e));
    return (int)e;
}
