// Shared device helpers for the sm_100a kernels: 128-bit bf16 vectors, warp/block reductions.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define FTL_API extern "C" __attribute__((visibility("default")))

#define FTL_LAUNCH_CHECK()                         \
    do {                                           \
        cudaError_t _e = cudaGetLastError();       \
        if (_e != cudaSuccess) return (int)_e;     \
    } while (0)

namespace ftl {

constexpr int kNumSMs = 148;  // B200

struct alignas(16) bf16x8 {
    __nv_bfloat162 v[4];
};

__device__ __forceinline__ bf16x8 ld8(const __nv_bfloat16* p) {
    return *reinterpret_cast<const bf16x8*>(p);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const bf16x8& x) {
    *reinterpret_cast<bf16x8*>(p) = x;
}
// 32-byte stores (sm_100: 256-bit global accesses): a thread that owns a whole row segment writes full 32-byte sectors,
// where two 16-byte stores from different instructions each reach L2 as a partial-sector write (read-modify-write when
// the line is not resident).  The attention epilogues, one row per thread, spent ~9 us per 32 KB tile on those.
__device__ __forceinline__ void st_row32(__nv_bfloat16* p, const bf16x8& x, const bf16x8& y) {
    const int4& a = *reinterpret_cast<const int4*>(&x);
    const int4& b = *reinterpret_cast<const int4*>(&y);
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w) : "memory");
}
// 32-byte load of a row segment owned by one thread: half as many per-line transactions as two 16-byte loads
__device__ __forceinline__ void ld_row32(const void* p, uint32_t* d) {
    asm volatile("ld.global.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]) : "l"(p));
}
__device__ __forceinline__ void st8f(float* p, const uint32_t* v) {      // 8 fp32 values given as raw bits
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
// streaming variants: data touched once (optimizer state, gradients) should not displace L2 residents
__device__ __forceinline__ bf16x8 ld8_stream(const __nv_bfloat16* p) {
    int4 r;
    asm volatile("ld.global.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return *reinterpret_cast<bf16x8*>(&r);
}
__device__ __forceinline__ void st8_stream(__nv_bfloat16* p, const bf16x8& x) {
    const int4& r = *reinterpret_cast<const int4*>(&x);
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w) : "memory");
}

__device__ __forceinline__ void unpack8(const bf16x8& x, float (&f)[8]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(x.v[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ bf16x8 pack8(const float (&f)[8]) {
    bf16x8 x;
#pragma unroll
    for (int i = 0; i < 4; ++i) x.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return x;
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum, result broadcast to every thread.  `red` is >= 33 floats of shared memory.
// Fixed reduction order: deterministic for a fixed block size.
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect `red` from a previous use
    if (lane == 0) red[wid] = v;
    __syncthreads();
    if (wid == 0) {
        float t = lane < nw ? red[lane] : 0.f;
        t = warp_sum(t);
        if (lane == 0) red[32] = t;
    }
    __syncthreads();
    return red[32];
}
__device__ __forceinline__ float block_max(float v, float* red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    if (wid == 0) {
        float t = lane < nw ? red[lane] : -INFINITY;
        t = warp_max(t);
        if (lane == 0) red[32] = t;
    }
    __syncthreads();
    return red[32];
}

// AdamW's denominator sqrt(v) / bc2_sqrt + eps.  sqrtf() and the IEEE division send zero / very small arguments to
// fix-up subroutines, which made the update of mostly-untouched embedding rows (v == 0) compute-bound: 3.2 ms for a
// bucket that streams in 1.4 ms.  Below vcut = (eps * bc2_sqrt * 2^-26)^2 the root term is smaller than half an ulp of
// eps and vanishes in the rounding of "+ eps", so the result is eps exactly: those elements are given a harmless
// argument (1.0) and eps is selected afterwards.  Bit-identical to the plain expression; eps == 0 turns it off.
__device__ __forceinline__ float adam_vcut(float eps, float bc2_sqrt) {
    const float t = eps * bc2_sqrt * 0x1p-26f;
    return t * t;
}
__device__ __forceinline__ float adam_denom(float v, float bc2_sqrt, float eps, float vcut) {
    const bool dead = v < vcut;
    const float d = sqrtf(dead ? 1.f : v) / bc2_sqrt + eps;
    return dead ? eps : d;
}
// m / denom, bit-identical to the plain quotient, with a zero numerator kept off the IEEE division's slow path (its
// range check sends zero operands to a fix-up subroutine; first moments of untouched embedding rows are exactly zero).
__device__ __forceinline__ float adam_quot(float m, float denom) {
    float num = m == 0.f ? 1.f : m;
    asm("" : "+f"(num));                      // opaque: otherwise the compiler folds the select back into m / denom
    const float q = num / denom;
    return m == 0.f ? copysignf(0.f, m) : q;
}

}  // namespace ftl
