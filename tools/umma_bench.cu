// Cycles per tcgen05.mma (M = 128, K = 16, bf16) as a function of N, operand sources and concurrent traffic.
// One CTA per SM; thread 0 of warp 0 issues REP groups of 8 MMAs, commits, waits; optional background warps read
// TMEM (tcgen05.ld) or write shared memory while the MMAs run.  Operand contents are irrelevant (smem is left as is).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I csrc tools/umma_bench.cu -o build/umma_bench -lcuda
//   build/umma_bench            (on the GPU box)
//
// The attention kernels' tile shapes were chosen from this table (profiles/umma_shapes.md).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>

#include "common/ptx.cuh"

using namespace ftl;

struct Cfg {
    int n;          // UMMA N
    int ts;         // A from TMEM
    int b_mn;       // B MN-major
    int ldtm_warps; // background warps doing tcgen05.ld
    int sts_warps;  // background warps doing st.shared
    int acc_same;   // unused (kept for the table)
};

// descriptors are stepped by compile-time constants, as in the attention kernels, so that the issue loop is a run of
// back-to-back UTCHMMA instructions and the measurement shows the tensor pipe, not address arithmetic
template <int N, bool TS, bool B_MN>
__device__ __forceinline__ void issue8(uint32_t d, uint32_t tmem_a, uint64_t da, uint64_t db, uint32_t idesc) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint64_t dbk = B_MN ? db + (uint64_t)((k * 16 * 128) >> 4) : db + (uint64_t)(((k >> 2) * N * 128 + (k & 3) * 32) >> 4);
        if (TS) ptx::umma_bf16_ts(d, tmem_a + k * 8, dbk, idesc, 1u);
        else ptx::umma_bf16(d, da + (uint64_t)(((k >> 2) * 128 * 128 + (k & 3) * 32) >> 4), dbk, idesc, 1u);
    }
}

template <int N, bool TS, bool B_MN>
__global__ void __launch_bounds__(320, 1) umma_bench_kernel(Cfg c, int rep, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                 // 32 KB
    uint8_t* sB = smem + 32768;         // 64 KB (N up to 256)
    uint8_t* sX = smem + 98304;         // 32 KB scratch for the st.shared warps
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    __shared__ volatile int stop;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); stop = 0; }
    if (warp == 1) ptx::tmem_alloc<512>(&tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (warp == 0) {
        if (ptx::elect_one()) {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, N, false, B_MN);
            const uint64_t da = ptx::umma_smem_desc(ptx::smem_u32(sA), 0, 1024);
            const uint64_t db = B_MN ? ptx::umma_smem_desc(ptx::smem_u32(sB), 128 * 128, 1024) : ptx::umma_smem_desc(ptx::smem_u32(sB), 0, 1024);
            const long long t0 = clock64();
            for (int r = 0; r < rep; ++r) issue8<N, TS, B_MN>(tmem + 256, tmem, da, db, idesc);
            const long long t1 = clock64();                 // issue time
            ptx::umma_commit(&bar);
            ptx::mbar_wait(&bar, 0);
            const long long t2 = clock64();                 // completion
            stop = 1;
            if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
        }
    } else if (warp >= 2 && warp < 2 + c.ldtm_warps) {
        uint32_t v[32];
        uint32_t acc = 0;
        const uint32_t addr = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        long long n = 0;
        while (!stop) {
            ptx::tmem_ld32(addr + 128, v);
            ptx::tmem_ld32(addr + 160, v);
            ptx::tmem_ld32(addr + 192, v);
            ptx::tmem_ld32(addr + 224, v);
            ptx::tmem_ld_wait();
            acc += v[0];
            ++n;
        }
        if (acc == 0x12345 && lane == 0) out[7] = acc;
        if (blockIdx.x == 0 && warp == 2 && lane == 0) out[2] = n;
    } else if (warp >= 6 && warp < 6 + c.sts_warps) {
        long long n = 0;
        const uint32_t base = ptx::smem_u32(sX) + threadIdx.x * 16;
        while (!stop) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(base + ((i * 4096) & 32767)), "r"((int)n) : "memory");
            ++n;
        }
        if (blockIdx.x == 0 && warp == 6 && lane == 0) out[3] = n;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem);
}

int main() {
    long long* out;
    cudaMalloc(&out, 64);
    const int smem = 98304 + 32768 + 1024;
    const int rep = 512;
    const Cfg cfgs[] = {
        {64, 0, 0, 0, 0, 1},  {128, 0, 0, 0, 0, 1}, {256, 0, 0, 0, 0, 1},
        {128, 0, 1, 0, 0, 1}, {64, 0, 1, 0, 0, 1},  {128, 1, 1, 0, 0, 1}, {64, 1, 1, 0, 0, 1},  {128, 1, 0, 0, 0, 1},
        {128, 0, 0, 4, 0, 1}, {128, 1, 1, 4, 0, 1}, {128, 0, 0, 0, 4, 1}, {128, 1, 1, 0, 4, 1}, {128, 0, 0, 4, 4, 1},
        {256, 0, 0, 4, 0, 1}, {64, 0, 0, 4, 0, 1},
    };
    printf("| N | A | B | ldtm warps | sts warps | issue cyc/MMA | total cyc/MMA | bg ld iters | bg st iters |\n|---|---|---|---|---|---|---|---|---|\n");
    for (int grid : {1, 148}) {
        printf("grid %d\n", grid);
        for (const Cfg& c : cfgs) {
            cudaMemset(out, 0, 64);
#define LAUNCH(NN, TT, BB) { cudaFuncSetAttribute(umma_bench_kernel<NN, TT, BB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); \
                             umma_bench_kernel<NN, TT, BB><<<grid, 320, smem>>>(c, rep, out); }
            if (c.n == 64 && !c.ts && !c.b_mn) LAUNCH(64, false, false)
            else if (c.n == 128 && !c.ts && !c.b_mn) LAUNCH(128, false, false)
            else if (c.n == 256 && !c.ts && !c.b_mn) LAUNCH(256, false, false)
            else if (c.n == 128 && !c.ts && c.b_mn) LAUNCH(128, false, true)
            else if (c.n == 64 && !c.ts && c.b_mn) LAUNCH(64, false, true)
            else if (c.n == 128 && c.ts && c.b_mn) LAUNCH(128, true, true)
            else if (c.n == 64 && c.ts && c.b_mn) LAUNCH(64, true, true)
            else if (c.n == 128 && c.ts && !c.b_mn) LAUNCH(128, true, false)
            else continue;
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            long long h[8];
            cudaMemcpy(h, out, 64, cudaMemcpyDeviceToHost);
            printf("| %d | %s | %s | %d | %d | %.1f | %.1f | %lld | %lld |\n", c.n, c.ts ? "tmem" : "smem", c.b_mn ? "MN" : "K", c.ldtm_warps,
                   c.sts_warps, h[0] / (rep * 8.0), h[1] / (rep * 8.0), h[2], h[3]);
        }
    }
    return 0;
}
