// bf16 GEMM on the 5th-generation tensor cores: TMA -> 128B-swizzled shared memory ->
// tcgen05.mma (one elected thread) -> fp32 accumulators in TMEM -> tcgen05.ld -> fused epilogue -> global.
//
//   C[M, N] = sum_r A(m, r) * B(n, r)
//
// A and B are each either "K-major" (reduction index contiguous: x @ W^T, the forward Linear of
// /root/reference/model.py:195, 215, 254, 379) or "MN-major" (output index contiguous: the operand is the
// transpose of a row-major matrix, which is what autograd's dgrad dY @ W and wgrad dY^T @ X need), so the
// three GEMM flavours of a training step run through one kernel without transposing anything in memory.
// The reduction may be split over two (A, B) pairs (kb < nk1 reads pair 1, the rest pair 2), which turns
// d(hn) = dgate @ W1 + dup @ W3 into a single GEMM over K = 2 x 14336 with no accumulate pass.
//
// Fused epilogues:
//   EPI_PLAIN   bf16 store.
//   EPI_SWIGLU  B tile = 128 rows of w1 | 128 rows of w3 (two TMA boxes, one N=256 MMA); the epilogue
//               writes gate, up (bf16, saved for backward) and act = bf16(bf16(silu(gate)) * up)  (K9+K10).
//   EPI_ROPE    rotary embedding on the q|k columns of the fused QKV projection (K3+K4): the accumulator
//               is rounded to bf16 (the Linear's output dtype), adjacent pairs are rotated in fp32 with
//               the interleaved convention of model.py:121-126, and rounded once more.
//
//   EPI_SWIGLU_BWD  the w2 dgrad GEMM (dact = dx @ w2) whose epilogue reads the saved gate / up tiles and writes
//               dgate = dact * up * silu'(gate) and dup = dact * silu(gate): dact never exists in memory.
//   sumsq       any plain GEMM can also return per-(CTA, epilogue warp) sums of squares of the bf16 values it stored:
//               the weight-gradient GEMMs produce the global gradient norm's partials as they write the gradient.
//
// Two kernels share the epilogues.  Default (gemm_tc2_kernel): persistent CTA pairs, tcgen05 cta_group::2 — the two
// SMs of a TPC compute one 256 x 256 tile, each CTA stages its 128 rows of A and half of the B tile (6-stage, 32 KB
// per stage TMA ring), the leader issues M=256 UMMAs, two accumulator stages of 256 TMEM columns let the epilogue of
// tile i overlap the MMAs of tile i+1; warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2-5 = epilogue
// (and, in the FSDP variants, the NVLink gather of the weight operand).  gemm_tc_kernel (FTL_GEMM_CLUSTER=1/2/4) is
// the earlier single-CTA-MMA kernel with a 4-stage ring and optional cluster multicast of B; kept as a baseline.
// Tiles are walked M-fastest, so CTAs running at the same time share a B (weight) tile through L2.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cudaTypedefs.h>

#include <stdlib.h>

#include "common/ptx.cuh"
#include "common/tmap_cache.cuh"
#include "common/util.cuh"

namespace ftl {

constexpr int BM = 128, BN = 256, BK = 64, kStages = 4, kAccStages = 2;
constexpr int kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kGemmThreads = 192;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;

enum { EPI_PLAIN = 0, EPI_SWIGLU = 1, EPI_ROPE = 2, EPI_SWIGLU_BWD = 3, EPI_ACC32 = 4 };

struct GemmArgs {
    __nv_bfloat16* C;       // PLAIN/ROPE: [M, N]; SWIGLU: act [M, N/2]
    __nv_bfloat16* C_gate;  // SWIGLU: raw gate / up outputs (may be null)
    __nv_bfloat16* C_up;
    const float* cos_t;     // ROPE: fp32 [seq_len, head_dim/2]
    const float* sin_t;
    __nv_bfloat16* C2;      // SWIGLU_BWD: dup (C = dgate)
    const __nv_bfloat16* in_gate;   // SWIGLU_BWD: saved pre-activations [M, N] (ldc)
    const __nv_bfloat16* in_up;
    float* sumsq;           // optional [gridDim.x * 4]: sum of squares of the bf16 outputs per (CTA, epilogue warp)
    float* C32;             // optional fp32 [M, N] (ldc) accumulator across calls (chunked reductions): see acc_mode
    int acc_mode;           // EPI_ACC32: bit 0 = add to the old C32 (else overwrite it), bit 1 = also write C = bf16(C32)
    int64_t ldc;
    int M, N, K;            // N counts accumulator columns (SWIGLU: 2 x hidden)
    int nk1;                // k-blocks served by the first (A, B) pair; the rest come from the second
    int seq_len, head_dim, rot_cols;
    // ---- fused parameter all-gather (FSDP): the B operand is row-sharded over `world` ranks; this kernel pulls the
    // shards out of the owners' memory over NVLink into the local staging copy that its own TMA loads then read.
    // Work item = 16 rows of one 128-row half block; the epilogue warps of every CTA take items from a counter
    // while they wait for an accumulator, so the transfer rides along with the MMAs instead of preceding them.
    const __nv_bfloat16* const* gather_src;  // [world] base pointers of the owners' compact shard buffers (null = off)
    __nv_bfloat16* gather_dst[2];            // local [rows, K] destination(s): B (and B2 for the SwiGLU pair)
    int64_t gather_off[2];                   // element offset of this matrix' shard inside every owner's buffer
    int gather_rows_per_rank;                // rows of each matrix owned by one rank (multiple of 128)
    int gather_nmat;                         // 1, or 2 for SwiGLU (w1, w3)
    int rank, world;
    int* gather_next;                        // [1]  item dispenser      (zeroed before the launch)
    int* gather_done;                        // [nmat * N/128] slices finished per half block (zeroed before launch)
};

// Work items are SMALL (2 rows = 16 KB at K = 4096) on purpose.  All gathering warps of the GPU pull items off one
// in-order dispenser and share the NVLink bandwidth equally, so everything that is in flight completes at about the
// same time: with 16-row items ~1,200 warps held 150 MB in flight, whole waves of half blocks landed together and the
// MMAs waited for the end of each wave (0.46 ms instead of 0.33 ms for the gate|up GEMM at W = 2, whatever the number
// of gathering warps).  With 2-row items the window in flight is ~19 MB — the ~9 n-blocks the CTA pairs are working
// on — and it slides over the weight matrix in the order the producers need it.
constexpr int kGatherSliceRows = 2, kGatherSlices = 128 / kGatherSliceRows;
// The gather variants launch 4 more warps per CTA that do nothing but move weight slices: the epilogue warps alone only
// gather while they wait for an accumulator, and with a heavy epilogue (SwiGLU: three stores per element) they are
// rarely idle — the transfer then ran after the math instead of under it (0.498 ms fused vs 0.326 ms + 0.153 ms apart).
constexpr int kGatherWarpsMax = 8, kGemmThreadsGather = kGemmThreads + 32 * kGatherWarpsMax;

// Copy one work item (kGatherSliceRows rows of B) from its owner to the local staging matrix.  One warp per item.
__device__ __forceinline__ bool gather_one_item(const GemmArgs& g, int lane) {
    const int halves = (g.gather_nmat == 2 ? g.N / 2 : g.N) / 128;           // half blocks per matrix
    const int total = halves * g.gather_nmat * kGatherSlices;
    int it = 0;
    if (lane == 0) it = atomicAdd(g.gather_next, 1);
    it = __shfl_sync(0xffffffffu, it, 0);
    if (it >= total) return false;
    const int slice = it % kGatherSlices;
    const int mat = (it / kGatherSlices) % g.gather_nmat;
    const int hpr = g.gather_rows_per_rank / 128;                             // half blocks per rank
    const int hb = ((it / kGatherSlices) / g.gather_nmat + g.rank * hpr) % halves;   // own shard first, then rotate
    const int owner = hb / hpr;
    const int64_t row_in_shard = (int64_t)(hb % hpr) * 128 + slice * kGatherSliceRows;
    const int4* src = reinterpret_cast<const int4*>(g.gather_src[owner] + g.gather_off[mat] + row_in_shard * g.K);
    int4* dst = reinterpret_cast<int4*>(g.gather_dst[mat] + ((int64_t)hb * 128 + slice * kGatherSliceRows) * g.K);
    const int nvec = kGatherSliceRows * g.K / 8;
    int i = 0;
    // 8 independent 32-byte peer loads in flight per lane: 8 KB per warp.  NVLink needs ~770 GB/s x ~3 us of round trip
    // = 2.3 MB in flight per GPU just to stay busy; with 16-byte loads the ~600 gathering warps of a launch held 2.4 MB
    // and the fused gather ran at ~280 GB/s.  (Rows are 16-byte aligned by contract; 32-byte accesses need K % 16 == 0
    // and 32-byte aligned bases, otherwise the 16-byte loop below does all of it.)
    if ((g.K & 15) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 31) == 0) {
        const int n32 = nvec / 2;
        for (int k = lane; k + 7 * 32 < n32; k += 8 * 32) {
            uint32_t v[8][8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]), "=r"(v[u][4]), "=r"(v[u][5]),
                               "=r"(v[u][6]), "=r"(v[u][7]) : "l"(src + 2 * (k + u * 32)));
#pragma unroll
            for (int u = 0; u < 8; ++u)
                asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                             :: "l"(dst + 2 * (k + u * 32)), "r"(v[u][0]), "r"(v[u][1]), "r"(v[u][2]), "r"(v[u][3]), "r"(v[u][4]),
                                "r"(v[u][5]), "r"(v[u][6]), "r"(v[u][7]) : "memory");
        }
        i = (n32 / (8 * 32)) * (8 * 32) * 2;     // 16-byte vectors already moved
    }
    for (i += lane; i < nvec; i += 32) dst[i] = src[i];
    __threadfence();                           // the slice is visible device-wide before it is counted
    __syncwarp();
    if (lane == 0) atomicAdd(g.gather_done + mat * halves + hb, 1);
    return true;
}

// Producer side: block until every slice of half block `hb` of matrix `mat` has landed, then make those generic-proxy
// writes visible to the TMA (async proxy) loads that follow.
__device__ __forceinline__ void gather_wait(const GemmArgs& g, int mat, int hb) {
    const int halves = (g.gather_nmat == 2 ? g.N / 2 : g.N) / 128;
    const int* flag = g.gather_done + mat * halves + hb;
    int v;
    const uint64_t t0 = ptx::global_timer_ns();
    do {
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v < kGatherSlices && ptx::global_timer_ns() - t0 > 4000000000ull) __trap();
    } while (v < kGatherSlices);
    asm volatile("fence.proxy.async.global;" ::: "memory");
}

// One thread = one output row of the tile: TMEM (32 fp32 columns at a time) -> registers -> fused math -> 16-byte
// global stores.  taddr addresses this warp's TMEM lane quarter of the accumulator stage.
__device__ __forceinline__ void st16(__nv_bfloat16* p, const bf16x8& v) {
    const int4& r = *reinterpret_cast<const int4*>(&v);
    asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w) : "memory");
}

// Two adjacent 8-column groups of a row go out as ONE 32-byte store when the layout allows it (row stride and base
// 32-byte aligned, both groups inside N): a thread owns its row, so 16-byte stores issued by different instructions
// reach L2 as partial-sector writes (read-modify-write when the line is not resident) - the same pathology that cost
// the attention epilogues 9 us per 32 KB tile.
__device__ __forceinline__ void store_groups(__nv_bfloat16* base, bool wide, int64_t off, const bf16x8 (&o)[4], int col0, int n_cols) {
#pragma unroll
    for (int v = 0; v < 4; v += 2) {
        const int col = col0 + v * 8;
        if (wide && col + 16 <= n_cols) {
            st_row32(base + off + v * 8, o[v], o[v + 1]);
        } else {
            if (col < n_cols) st16(base + off + v * 8, o[v]);
            if (col + 8 < n_cols) st16(base + off + v * 8 + 8, o[v + 1]);
        }
    }
}
__device__ __forceinline__ bool wide_ok(const void* p, int64_t ld) {
    return p != nullptr && (ld & 15) == 0 && (reinterpret_cast<uintptr_t>(p) & 31) == 0;
}

template <int kEpi>
__device__ __forceinline__ void epilogue_tile(const GemmArgs& g, uint32_t taddr, int row, int n_blk, float& ss) {
                if (kEpi == EPI_SWIGLU_BWD) {
                    // accumulator = dact tile.  N % 16 == 0, ldc % 16 == 0 and 32-byte aligned bases (checked on the
                    // host): every 16-column group is read and written with 32-byte row accesses.
                    const int n0 = n_blk * BN;
    #pragma unroll 1
                    for (int c = 0; c < BN; c += 32) {
                        uint32_t gw[2][8], uw[2][8];
                        const bool live = row < g.M;
                        const int64_t off0 = (int64_t)row * g.ldc + n0 + c;
                        bool in[2];
    #pragma unroll
                        for (int h = 0; h < 2; ++h) {        // issue the gate / up loads before waiting on TMEM
                            in[h] = live && (n0 + c + h * 16 < g.N);
                            if (in[h]) {
                                ld_row32(g.in_gate + off0 + h * 16, gw[h]);
                                ld_row32(g.in_up + off0 + h * 16, uw[h]);
                            }
                        }
                        uint32_t r[32];
                        ptx::tmem_ld32(taddr + c, r);
                        ptx::tmem_ld_wait();
    #pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            if (!in[h]) continue;
                            bf16x8 og[2], ou[2];
    #pragma unroll
                            for (int v = 0; v < 2; ++v) {
                                float gt[8], up[8], a[8], b[8];
                                unpack8(*reinterpret_cast<const bf16x8*>(&gw[h][v * 4]), gt);
                                unpack8(*reinterpret_cast<const bf16x8*>(&uw[h][v * 4]), up);
    #pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    const float d = round_bf16(__uint_as_float(r[h * 16 + v * 8 + j]));   // dact in bf16, as autograd
                                    const float sig = 1.f / (1.f + __expf(-gt[j]));
                                    a[j] = d * up[j] * (sig * (1.f + gt[j] * (1.f - sig)));
                                    b[j] = d * (gt[j] * sig);
                                }
                                og[v] = pack8(a); ou[v] = pack8(b);
                            }
                            st_row32(g.C + off0 + h * 16, og[0], og[1]);
                            st_row32(g.C2 + off0 + h * 16, ou[0], ou[1]);
                        }
                    }
                } else if (kEpi == EPI_SWIGLU) {
                    const int h0 = n_blk * (BN / 2);   // first hidden column of this tile
                    const int hidden = g.N / 2;
                    const bool w_act = wide_ok(g.C, g.ldc), w_gate = wide_ok(g.C_gate, g.ldc), w_up = wide_ok(g.C_up, g.ldc);
    #pragma unroll 1
                    for (int c = 0; c < BN / 2; c += 32) {
                        uint32_t rg[32], ru[32];
                        ptx::tmem_ld32(taddr + c, rg);
                        ptx::tmem_ld32(taddr + BN / 2 + c, ru);
                        ptx::tmem_ld_wait();
                        if (row < g.M) {
                            bf16x8 oa[4], og[4], ou[4];
    #pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                float gt[8], up[8], act[8];
    #pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    gt[j] = round_bf16(__uint_as_float(rg[v * 8 + j]));
                                    up[j] = round_bf16(__uint_as_float(ru[v * 8 + j]));
                                    act[j] = round_bf16(gt[j] / (1.f + __expf(-gt[j]))) * up[j];
                                }
                                oa[v] = pack8(act); og[v] = pack8(gt); ou[v] = pack8(up);
                            }
                            const int64_t off = (int64_t)row * g.ldc + h0 + c;
                            store_groups(g.C, w_act, off, oa, h0 + c, hidden);
                            if (g.C_gate) store_groups(g.C_gate, w_gate, off, og, h0 + c, hidden);
                            if (g.C_up) store_groups(g.C_up, w_up, off, ou, h0 + c, hidden);
                        }
                    }
                } else {
                    const int n0 = n_blk * BN;
                    const bool w_c = wide_ok(g.C, g.ldc);
    #pragma unroll 1
                    for (int c = 0; c < BN; c += 32) {
                        uint32_t r[32];
                        ptx::tmem_ld32(taddr + c, r);
                        ptx::tmem_ld_wait();
                        if (row < g.M) {
                            bf16x8 o[4];
    #pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                const int col = n0 + c + v * 8;
                                float f[8];
    #pragma unroll
                                for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[v * 8 + j]);
                                if (kEpi == EPI_ROPE && col < g.rot_cols && col < g.N) {
                                    const int pos = row % g.seq_len;
                                    const int pair0 = (col % g.head_dim) >> 1;
                                    const int64_t to = (int64_t)pos * (g.head_dim >> 1) + pair0;
                                    const float4 cs = *reinterpret_cast<const float4*>(g.cos_t + to);
                                    const float4 sn = *reinterpret_cast<const float4*>(g.sin_t + to);
                                    const float cc[4] = {cs.x, cs.y, cs.z, cs.w}, ss[4] = {sn.x, sn.y, sn.z, sn.w};
    #pragma unroll
                                    for (int j = 0; j < 4; ++j) {
                                        const float a = round_bf16(f[2 * j]), b = round_bf16(f[2 * j + 1]);
                                        f[2 * j] = a * cc[j] - b * ss[j];
                                        f[2 * j + 1] = a * ss[j] + b * cc[j];
                                    }
                                }
                                if (kEpi == EPI_ACC32 && col < g.N) {
                                    // fp32 running sum over several launches (e.g. the vocabulary chunks of the LM head):
                                    // one 32-byte read-modify-write per 8 columns; bf16 is written by the last launch only
                                    float* p32 = g.C32 + (int64_t)row * g.ldc + col;
                                    if (g.acc_mode & 1) {
                                        uint32_t old[8];
                                        ld_row32(p32, old);
    #pragma unroll
                                        for (int j = 0; j < 8; ++j) f[j] += __uint_as_float(old[j]);
                                    }
                                    uint32_t nw[8];
    #pragma unroll
                                    for (int j = 0; j < 8; ++j) nw[j] = __float_as_uint(f[j]);
                                    st8f(p32, nw);
                                }
                                o[v] = pack8(f);
                                if (kEpi == EPI_PLAIN && g.sumsq != nullptr && col < g.N) {   // norm² of what is stored
                                    float q[8];
                                    unpack8(o[v], q);
    #pragma unroll
                                    for (int j = 0; j < 8; ++j) ss += q[j] * q[j];
                                }
                            }
                            if (kEpi != EPI_ACC32 || (g.acc_mode & 2))
                                store_groups(g.C, w_c, (int64_t)row * g.ldc + n0 + c, o, n0 + c, g.N);
                        }
                    }
                }
}

template <int kEpi, bool kAMn, bool kBMn, int kCluster>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
               const GemmArgs g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
    uint64_t* full = bars;                       // [kStages]     TMA -> MMA
    uint64_t* empty = bars + kStages;            // [kStages]     MMA -> TMA
    uint64_t* acc_full = bars + 2 * kStages;     // [kAccStages]  MMA -> epilogue
    uint64_t* acc_empty = acc_full + kAccStages; // [kAccStages]  epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + kAccStages);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = (g.M + BM - 1) / BM;
    const int n_tiles = (g.N + BN - 1) / BN;
    const int num_kb = (g.K + BK - 1) / BK;
    // A cluster of kCluster CTAs works on kCluster vertically adjacent tiles (same n block): they need the same
    // B tile, so each CTA fetches 1/kCluster of it and TMA-multicasts that part into every CTA of the cluster.
    // L2 -> SM traffic per CTA and k-block drops from 48 KB to 16 + 32/kCluster KB; at 48 KB the kernel is
    // bound by L2 bandwidth (profiles/gemm_tc_v1_ncu.md), not by the tensor pipe.
    const int crank = kCluster > 1 ? (int)ptx::cluster_ctarank() : 0;
    const int cluster_id = blockIdx.x / kCluster, num_clusters = gridDim.x / kCluster;
    const int m_groups = (m_tiles + kCluster - 1) / kCluster;
    const int num_units = m_groups * n_tiles;
    constexpr uint16_t kMask = (uint16_t)((1u << kCluster) - 1);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        for (int i = 0; i < kStages; ++i) {
            ptx::mbar_init(&full[i], 1);
            ptx::mbar_init(&empty[i], kCluster);   // every CTA of the cluster must have consumed the slot
        }
        for (int i = 0; i < kAccStages; ++i) {
            ptx::mbar_init(&acc_full[i], 1);
            ptx::mbar_init(&acc_empty[i], 4);  // one arrive per epilogue warp
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    if (kCluster > 1) ptx::cluster_sync();   // peers' barriers exist before anyone multicasts into them
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================= TMA producer =================
        if (ptx::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
                const int m0 = ((unit % m_groups) * kCluster + crank) * BM;
                const int n_blk = unit / m_groups;
                for (int kb = 0; kb < num_kb; ++kb) {
                    ptx::mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * kStageBytes;
                    uint8_t* sb = sa + kABytes;
                    ptx::mbar_arrive_expect_tx(&full[stage], kStageBytes);
                    const bool second = kb >= g.nk1;
                    const CUtensorMap* ma = second ? &tmA2 : &tmA;
                    const CUtensorMap* mb = second ? &tmB2 : &tmB;
                    const int r0 = (second ? kb - g.nk1 : kb) * BK;  // reduction coordinate
                    if (!kAMn) {
                        ptx::tma_load_2d(sa, ma, &full[stage], r0, m0);                 // box {64 r, 128 m}
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)                               // boxes {64 m, 64 r}
                            ptx::tma_load_2d(sa + j * (64 * BK * 2), ma, &full[stage], m0 + j * 64, r0);
                    }
                    if (kEpi == EPI_SWIGLU) {
                        // rows [0,128) of the B tile from w1, rows [128,256) from w3 (same row range of each)
                        if (kCluster == 1) {
                            ptx::tma_load_2d(sb, &tmB, &full[stage], r0, n_blk * (BN / 2));
                            ptx::tma_load_2d(sb + kBBytes / 2, &tmB2, &full[stage], r0, n_blk * (BN / 2));
                        } else {   // kCluster == 2: CTA 0 supplies the w1 half, CTA 1 the w3 half, to both CTAs
                            ptx::tma_load_2d_mcast(sb + crank * (kBBytes / 2), crank == 0 ? &tmB : &tmB2, &full[stage], r0,
                                                   n_blk * (BN / 2), kMask);
                        }
                    } else if (!kBMn) {
                        if (kCluster == 1) {
                            ptx::tma_load_2d(sb, mb, &full[stage], r0, n_blk * BN);     // box {64 r, 256 n}
                        } else {                                                        // box {64 r, 256/kCluster n}
                            ptx::tma_load_2d_mcast(sb + crank * (kBBytes / kCluster), mb, &full[stage], r0,
                                                   n_blk * BN + crank * (BN / kCluster), kMask);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j) {                             // boxes {64 n, 64 r}
                            if (kCluster == 1)
                                ptx::tma_load_2d(sb + j * (64 * BK * 2), mb, &full[stage], n_blk * BN + j * 64, r0);
                            else if (j % kCluster == crank)
                                ptx::tma_load_2d_mcast(sb + j * (64 * BK * 2), mb, &full[stage], n_blk * BN + j * 64, r0, kMask);
                        }
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        constexpr uint32_t idesc = ptx::umma_idesc_bf16(BM, BN, kAMn, kBMn);
        // K-major, 128B swizzle: 8-row groups are 1024 B apart (SBO); LBO unused.
        // MN-major, 128B swizzle: 64-element (128 B) columns blocks are one TMA box apart (LBO = 64 r x 128 B),
        // 8-row reduction groups 1024 B apart (SBO).
        constexpr uint32_t kLboMn = 64 * BK * 2, kSbo = 1024;
        int stage = 0, acc = 0;
        uint32_t phase = 0, acc_phase = 0;
        for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
            ptx::mbar_wait(&acc_empty[acc], acc_phase ^ 1);
            ptx::tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * BN;
            for (int kb = 0; kb < num_kb; ++kb) {
                ptx::mbar_wait(&full[stage], phase);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t sa = ptx::smem_u32(smem + stage * kStageBytes);
                    const uint32_t sb = sa + kABytes;
                    // one UMMA consumes 16 reduction elements: 32 B inside the swizzle row (K-major) or 16 rows of
                    // 128 B (MN-major).  The descriptor is built once per stage and stepped by adding a constant to
                    // its start-address field: the single issuing thread must not spend ~30 integer ops per UMMA.
                    const uint64_t da0 = kAMn ? ptx::umma_smem_desc(sa, kLboMn, kSbo) : ptx::umma_smem_desc(sa, 0, kSbo);
                    const uint64_t db0 = kBMn ? ptx::umma_smem_desc(sb, kLboMn, kSbo) : ptx::umma_smem_desc(sb, 0, kSbo);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                        ptx::umma_bf16(tmem_d, da0 + (uint64_t)((kAMn ? k * 16 * 128 : k * 32) >> 4),
                                       db0 + (uint64_t)((kBMn ? k * 16 * 128 : k * 32) >> 4), idesc, (kb | k) != 0);
                    // free the smem slot when the MMAs retire — in every CTA of the cluster, because the peers'
                    // multicast writes land in this CTA's slot too
                    if (kCluster == 1) ptx::umma_commit(&empty[stage]);
                    else ptx::umma_commit_mcast(&empty[stage], kMask);
                    if (kb == num_kb - 1) ptx::umma_commit(&acc_full[acc]);  // accumulator ready
                }
                __syncwarp();
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
    } else {
        // ================= epilogue: TMEM -> registers -> global =================
        const int q = warp & 3;                 // TMEM lane quarter this warp may read
        int acc = 0;
        uint32_t acc_phase = 0;
        float ss = 0.f;
        for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
            const int m0 = ((unit % m_groups) * kCluster + crank) * BM;
            const int n_blk = unit / m_groups;
            const int row = m0 + q * 32 + lane;
            ptx::mbar_wait(&acc_full[acc], acc_phase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
            epilogue_tile<kEpi>(g, taddr, row, n_blk, ss);
            // every TMEM read of this accumulator stage has completed (wait::ld above): hand it back
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&acc_empty[acc]);
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (g.sumsq != nullptr) {               // fixed tile -> CTA map and fixed shuffle tree: deterministic
            ss = warp_sum(ss);
            if (lane == 0) g.sumsq[blockIdx.x * 4 + q] = ss;
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (kCluster > 1) ptx::cluster_sync();   // no CTA leaves while a peer may still signal its barriers
    if (warp == 1) ptx::tmem_dealloc<512>(tmem_base);
}


// ------------------------------------------------------------------------------------------------------------
// CTA-pair version (tcgen05 cta_group::2).  Two CTAs on the two SMs of a TPC compute one 256 x 256 tile: each
// CTA keeps its own 128 rows of A and HALF of the B tile (128 of the 256 n-rows) in shared memory, the leader
// CTA issues one M=256 UMMA per 16 reduction elements that reads A from both CTAs' shared memory and both B
// halves, and each CTA's TMEM receives its 128 x 256 half of the accumulator.  Per CTA a k-block now costs
// 32 KB of TMA writes and 8 KB of operand reads per UMMA instead of 48 KB / 12 KB, which is what lifts the
// shared-memory-bandwidth ceiling of the single-CTA kernel; the stage is smaller, so the ring is 6 deep.
//   full[s]   lives in the leader; both producers arrive on it (the peer through a cluster address) and both
//             CTAs' TMA loads complete their bytes on it
//   empty[s], acc_full[a]   one copy per CTA; the leader's tcgen05.commit is multicast to both
//   acc_empty[a]            lives in the leader; 4 epilogue warps of each CTA arrive on it
constexpr int kStages2 = 6;
constexpr int kStageBytes2 = kABytes + kBBytes / 2;   // 32 KB
constexpr int kSmemBytes2 = kStages2 * kStageBytes2 + 1024 + 256;

template <int kEpi, bool kAMn, bool kBMn>
__global__ void __launch_bounds__(kGemmThreadsGather, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                const GemmArgs g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages2 * kStageBytes2);
    uint64_t* full = bars;
    uint64_t* empty = bars + kStages2;
    uint64_t* acc_full = bars + 2 * kStages2;
    uint64_t* acc_empty = acc_full + kAccStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + kAccStages);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int crank = (int)ptx::cluster_ctarank();
    const bool leader = crank == 0;
    const int m_pairs = (g.M + 2 * BM - 1) / (2 * BM);
    const int n_tiles = (g.N + BN - 1) / BN;
    const int num_units = m_pairs * n_tiles;
    const int num_kb = (g.K + BK - 1) / BK;
    const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
    // with a fused gather, walk the n blocks starting at this rank's own shard: local rows need no NVLink traffic, and
    // the ranks then pull from different owners at any moment
    const int n_rot = g.gather_src ? (g.rank * n_tiles) / g.world : 0;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        for (int i = 0; i < kStages2; ++i) {
            ptx::mbar_init(&full[i], 1);      // the leader's producer arms it for the bytes of BOTH CTAs' loads
            ptx::mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < kAccStages; ++i) {
            ptx::mbar_init(&acc_full[i], 1);
            ptx::mbar_init(&acc_empty[i], 8);  // 4 epilogue warps x 2 CTAs (used in the leader only)
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc_2sm<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================= TMA producer (both CTAs) =================
        if (ptx::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int unit = pair_id; unit < num_units; unit += num_pairs) {
                const int m0 = ((unit % m_pairs) * 2 + crank) * BM;
                const int n_blk = (unit / m_pairs + n_rot) % n_tiles;
                const int n0 = n_blk * BN + crank * (BN / 2);     // this CTA's half of the B tile
                if (g.gather_src) {
                    // this CTA's B half must have been gathered before TMA reads it
                    if (kEpi == EPI_SWIGLU) gather_wait(g, crank, n_blk);
                    else gather_wait(g, 0, n_blk * 2 + crank);
                }
                for (int kb = 0; kb < num_kb; ++kb) {
                    ptx::mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * kStageBytes2;
                    uint8_t* sb = sa + kABytes;
                    // Both CTAs' TMA loads complete their bytes on the LEADER's barrier.  Only the leader arrives on it
                    // (a plain CTA-scope arrive; a cluster-scope release here costs a MEMBAR.ALL.GPU per k-block).
                    // The peer's bytes may land before the leader has armed the phase: the transaction count just
                    // goes negative until expect_tx, and the phase cannot complete before the leader's arrive.
                    const uint32_t fb = ptx::mapa(ptx::smem_u32(&full[stage]), 0);
                    if (leader) ptx::mbar_arrive_expect_tx(&full[stage], 2 * kStageBytes2);
                    const bool second = kb >= g.nk1;
                    const CUtensorMap* ma = second ? &tmA2 : &tmA;
                    const CUtensorMap* mb = second ? &tmB2 : &tmB;
                    const int r0 = (second ? kb - g.nk1 : kb) * BK;
                    if (!kAMn) {
                        ptx::tma_load_2d_2sm(sa, ma, fb, r0, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)
                            ptx::tma_load_2d_2sm(sa + j * (64 * BK * 2), ma, fb, m0 + j * 64, r0);
                    }
                    if (kEpi == EPI_SWIGLU) {
                        // accumulator columns [0,128) = gate (w1 rows, held by the leader), [128,256) = up (w3, peer)
                        ptx::tma_load_2d_2sm(sb, leader ? &tmB : &tmB2, fb, r0, n_blk * (BN / 2));
                    } else if (!kBMn) {
                        ptx::tma_load_2d_2sm(sb, mb, fb, r0, n0);                        // box {64 r, 128 n}
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 2 / 64; ++j)                           // boxes {64 n, 64 r}
                            ptx::tma_load_2d_2sm(sb + j * (64 * BK * 2), mb, fb, n0 + j * 64, r0);
                    }
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (leader CTA only) =================
        if (leader) {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(2 * BM, BN, kAMn, kBMn);
            constexpr uint32_t kLboMn = 64 * BK * 2, kSbo = 1024;
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int unit = pair_id; unit < num_units; unit += num_pairs) {
                ptx::mbar_wait(&acc_empty[acc], acc_phase ^ 1);
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    ptx::mbar_wait(&full[stage], phase);
                    ptx::tc_fence_after();
                    if (ptx::elect_one()) {
                        const uint32_t sa = ptx::smem_u32(smem + stage * kStageBytes2);
                        const uint32_t sb = sa + kABytes;
                        const uint64_t da0 = kAMn ? ptx::umma_smem_desc(sa, kLboMn, kSbo) : ptx::umma_smem_desc(sa, 0, kSbo);
                        const uint64_t db0 = kBMn ? ptx::umma_smem_desc(sb, kLboMn, kSbo) : ptx::umma_smem_desc(sb, 0, kSbo);
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k)
                            ptx::umma_bf16_2sm(tmem_d, da0 + (uint64_t)((kAMn ? k * 16 * 128 : k * 32) >> 4),
                                               db0 + (uint64_t)((kBMn ? k * 16 * 128 : k * 32) >> 4), idesc, (kb | k) != 0);
                        ptx::umma_commit_2sm(&empty[stage], 3);
                        if (kb == num_kb - 1) ptx::umma_commit_2sm(&acc_full[acc], 3);
                    }
                    __syncwarp();
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
                if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 2 + 4) {
        // ================= dedicated gather warps (present only in the fused all-gather launches) =================
        while (gather_one_item(g, lane)) {}
    } else {
        // ================= epilogue (both CTAs, own 128 rows) =================
        const int q = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        bool gather_left = g.gather_src != nullptr;
        float ss = 0.f;
        for (int unit = pair_id; unit < num_units; unit += num_pairs) {
            const int m0 = ((unit % m_pairs) * 2 + crank) * BM;
            const int n_blk = (unit / m_pairs + n_rot) % n_tiles;
            const int row = m0 + q * 32 + lane;
            // while the accumulator is not ready, move parameter slices (fused all-gather)
            while (gather_left && !ptx::mbar_try_wait(&acc_full[acc], acc_phase)) gather_left = gather_one_item(g, lane);
            ptx::mbar_wait(&acc_full[acc], acc_phase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
            epilogue_tile<kEpi>(g, taddr, row, n_blk, ss);
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&acc_empty[acc]), 0));
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (g.sumsq != nullptr) {
            ss = warp_sum(ss);
            if (lane == 0) g.sumsq[blockIdx.x * 4 + q] = ss;
        }
        while (gather_left) gather_left = gather_one_item(g, lane);   // CTAs without tiles still help
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    if (warp == 1) ptx::tmem_dealloc_2sm<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ host
// Row-major bf16 matrix [rows, cols] (leading dimension ld elements); box = {box_cols (contiguous), box_rows}.
// Encoded once per (pointer, shape, box) and cached (common/tmap_cache.cuh).
static int make_tmap(CUtensorMap* out, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_cols,
                     int box_rows) {
    return g_tmaps.get(out, ptr, rows, cols, ld, box_cols, box_rows);
}

template <int kEpi, bool kAMn, bool kBMn, int kCluster>
static int launch_c(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& a2, const CUtensorMap& b2,
                    const GemmArgs& g, int max_ctas, cudaStream_t stream) {
    auto kern = gemm_tc_kernel<kEpi, kAMn, kBMn, kCluster>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return (int)e;
        attr_set = true;
    }
    const int m_tiles = (g.M + BM - 1) / BM, n_tiles = (g.N + BN - 1) / BN;
    const int units = ((m_tiles + kCluster - 1) / kCluster) * n_tiles;
    int grid = max_ctas > 0 ? max_ctas : kNumSMs;
    if (grid > units * kCluster) grid = units * kCluster;
    grid = grid / kCluster * kCluster;
    if (grid < kCluster) grid = kCluster;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return (int)cudaLaunchKernelEx(&cfg, kern, a, b, a2, b2, g);
}

// dedicated gather warps per CTA in the fused all-gather launches (FTL_GATHER_WARPS = 0..8, default 4)
static int gather_warps() {
    static int w = -1;
    if (w < 0) {
        const char* e = getenv("FTL_GATHER_WARPS");
        w = e ? atoi(e) : 4;
        if (w < 0 || w > kGatherWarpsMax) w = 4;
    }
    return w;
}

template <int kEpi, bool kAMn, bool kBMn>
static int launch_pair(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& a2, const CUtensorMap& b2,
                       const GemmArgs& g, int max_ctas, cudaStream_t stream) {
    auto kern = gemm_tc2_kernel<kEpi, kAMn, kBMn>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes2);
        if (e != cudaSuccess) return (int)e;
        attr_set = true;
    }
    const int units = ((g.M + 2 * BM - 1) / (2 * BM)) * ((g.N + BN - 1) / BN);
    int grid = max_ctas > 0 ? max_ctas : kNumSMs;
    if (grid > units * 2) grid = units * 2;
    grid = grid / 2 * 2;
    if (grid < 2) grid = 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(g.gather_src ? kGemmThreads + 32 * gather_warps() : kGemmThreads);
    cfg.dynamicSmemBytes = kSmemBytes2;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return (int)cudaLaunchKernelEx(&cfg, kern, a, b, a2, b2, g);
}

// cluster size asked for by the environment (FTL_GEMM_CLUSTER=1 falls back to independent CTAs)
static int gemm_cluster_env() {
    static int c = -1;
    if (c < 0) {
        const char* e = getenv("FTL_GEMM_CLUSTER");
        c = e ? atoi(e) : 0;   // 0 = CTA-pair kernel (cta_group::2); 1 / 2 / 4 = single-CTA MMA with that cluster size
        if (c != 0 && c != 1 && c != 2 && c != 4) c = 0;
    }
    return c;
}

// True when p is ordinary device memory allocated on the GPU that is current now.
static bool on_this_gpu(const void* p) {
    if (!p) return true;
    int dev = -1;
    cudaPointerAttributes at{};
    if (cudaGetDevice(&dev) != cudaSuccess || cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice && at.device == dev;
}

// Cluster size actually used for a call whose B operand(s) are b / b2.  The cluster kernels (2, 4) fetch B with TMA
// multicast loads.  TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
// another GPU's memory hung the GPU: any operand that is not plain memory of the current device is routed to the
// kernel without multicast.
static int gemm_cluster(const void* b, const void* b2) {
    const int c = gemm_cluster_env();
    if (c >= 2 && !(on_this_gpu(b) && on_this_gpu(b2))) return 1;
    return c;
}

template <int kEpi, bool kAMn, bool kBMn>
static int launch(int c, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& a2, const CUtensorMap& b2,
                  const GemmArgs& g, int max_ctas, cudaStream_t stream) {
    if (c == 0) return launch_pair<kEpi, kAMn, kBMn>(a, b, a2, b2, g, max_ctas, stream);
    if (c == 1) return launch_c<kEpi, kAMn, kBMn, 1>(a, b, a2, b2, g, max_ctas, stream);
    if (c == 4 && kEpi != EPI_SWIGLU) return launch_c<kEpi, kAMn, kBMn, 4>(a, b, a2, b2, g, max_ctas, stream);
    return launch_c<kEpi, kAMn, kBMn, 2>(a, b, a2, b2, g, max_ctas, stream);
}

}  // namespace ftl

using namespace ftl;

// C[M,N] = A·B with a, b row-major bf16 matrices and `a_mn` / `b_mn` selecting how they are read:
//   a_mn = 0: a is [M, K] (lda);  a_mn = 1: a is [K, M] (lda)  — i.e. C = a^T ...
//   b_mn = 0: b is [N, K] (ldb);  b_mn = 1: b is [K, N] (ldb)
// Optional second reduction segment (a2, b2, K2): C = A·B + A2·B2.  max_ctas = 0 uses all 148 SMs.
// sumsq (optional): fp32 [ftl_gemm_sumsq_slots()] receiving the per-(CTA, epilogue warp) sums of squares of the stored
// bf16 values; entries past the launched grid are left untouched (keep them zero).
FTL_API int ftl_gemm_sumsq_slots(void) { return kNumSMs * 4; }

FTL_API int ftl_gemm_bf16_ex(const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb, int b_mn, void* c,
                             int64_t ldc, int M, int N, int K, const void* a2, int64_t lda2, const void* b2, int64_t ldb2,
                             int K2, float* sumsq, int max_ctas, cudaStream_t stream) {
    if (M <= 0 || N <= 0 || K <= 0 || (N % 8) || (ldc % 8) || (lda % 8) || (ldb % 8) || (K % BK && K2 > 0))
        return (int)cudaErrorInvalidValue;
    CUtensorMap ta, tb, ta2, tb2;
    int e;
    e = a_mn ? make_tmap(&ta, a, K, M, lda, 64, BK) : make_tmap(&ta, a, M, K, lda, BK, BM);
    if (e) return e;
    const int csel = gemm_cluster(b, K2 > 0 ? b2 : nullptr);
    const int cl = csel == 0 ? 2 : csel;
    e = b_mn ? make_tmap(&tb, b, K, N, ldb, 64, BK) : make_tmap(&tb, b, N, K, ldb, BK, BN / cl);
    if (e) return e;
    ta2 = ta;
    tb2 = tb;
    if (K2 > 0) {
        e = a_mn ? make_tmap(&ta2, a2, K2, M, lda2, 64, BK) : make_tmap(&ta2, a2, M, K2, lda2, BK, BM);
        if (e) return e;
        e = b_mn ? make_tmap(&tb2, b2, K2, N, ldb2, 64, BK) : make_tmap(&tb2, b2, N, K2, ldb2, BK, BN / cl);
        if (e) return e;
    }
    GemmArgs g{};
    g.C = (__nv_bfloat16*)c;
    g.ldc = ldc;
    g.M = M; g.N = N; g.K = K + (K2 > 0 ? K2 : 0);
    g.nk1 = (K + BK - 1) / BK;
    g.sumsq = sumsq;
    if (!a_mn && !b_mn) return launch<EPI_PLAIN, false, false>(csel, ta, tb, ta2, tb2, g, max_ctas, stream);
    if (!a_mn && b_mn) return launch<EPI_PLAIN, false, true>(csel, ta, tb, ta2, tb2, g, max_ctas, stream);
    if (a_mn && b_mn) return launch<EPI_PLAIN, true, true>(csel, ta, tb, ta2, tb2, g, max_ctas, stream);
    return launch<EPI_PLAIN, true, false>(csel, ta, tb, ta2, tb2, g, max_ctas, stream);
}

// C32 (+)= A.B in fp32 across several calls.  mode bit 0: add to what c32 holds (else overwrite it); bit 1: also write
// the bf16 result to c (c may be NULL otherwise).  c32: fp32 [M, N], same leading dimension ldc (elements), 32-byte aligned.
// a: [M, K] row-major; b: [K, N] row-major (the dgrad form dY @ W).
FTL_API int ftl_gemm_bf16_acc32(const void* a, int64_t lda, const void* b, int64_t ldb, void* c, float* c32, int64_t ldc,
                                int M, int N, int K, int mode, int max_ctas, cudaStream_t stream) {
    if (M <= 0 || N <= 0 || K <= 0 || (N % 8) || (ldc % 8) || (lda % 8) || (ldb % 8) || mode < 0 || mode > 3 || !c32 ||
        (reinterpret_cast<uintptr_t>(c32) & 31) || ((mode & 2) && !c))
        return (int)cudaErrorInvalidValue;
    CUtensorMap ta, tb;
    int e;
    if ((e = make_tmap(&ta, a, M, K, lda, BK, BM))) return e;
    const int csel = gemm_cluster(b, nullptr);
    if ((e = make_tmap(&tb, b, K, N, ldb, 64, BK))) return e;
    GemmArgs g{};
    g.C = (__nv_bfloat16*)c;
    g.C32 = c32; g.acc_mode = mode;
    g.ldc = ldc;
    g.M = M; g.N = N; g.K = K;
    g.nk1 = (K + BK - 1) / BK;
    return launch<EPI_ACC32, false, true>(csel, ta, tb, ta, tb, g, max_ctas, stream);
}

FTL_API int ftl_gemm_bf16(const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb, int b_mn, void* c,
                          int64_t ldc, int M, int N, int K, const void* a2, int64_t lda2, const void* b2, int64_t ldb2,
                          int K2, int max_ctas, cudaStream_t stream) {
    return ftl_gemm_bf16_ex(a, lda, a_mn, b, ldb, b_mn, c, ldc, M, N, K, a2, lda2, b2, ldb2, K2, nullptr, max_ctas, stream);
}

// SwiGLU backward fused into the w2 dgrad GEMM: dact = dx [M, K] @ w2 [K, H] stays in TMEM; the epilogue reads the saved
// gate / up [M, H] and writes dgate, dup [M, H] (all four with leading dimension ldo).  H and ldo multiples of 16,
// 32-byte aligned bases.  Replaces a [M, H] bf16 round trip and one elementwise kernel (reference: autograd of
// model.py:254).
FTL_API int ftl_gemm_swiglu_bwd(const void* dx, int64_t lddx, const void* w2, int64_t ldw, const void* gate, const void* up,
                                void* dgate, void* dup, int64_t ldo, int M, int H, int K, int max_ctas, cudaStream_t stream) {
    auto al32 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; };
    if (M <= 0 || H <= 0 || K <= 0 || (H % 16) || (ldo % 16) || (lddx % 8) || (ldw % 8) || !al32(gate) || !al32(up) ||
        !al32(dgate) || !al32(dup))
        return (int)cudaErrorInvalidValue;
    CUtensorMap ta, tb;
    int e;
    if ((e = make_tmap(&ta, dx, M, K, lddx, BK, BM))) return e;
    if ((e = make_tmap(&tb, w2, K, H, ldw, 64, BK))) return e;     // MN-major B: boxes {64 n, 64 r}
    GemmArgs g{};
    g.C = (__nv_bfloat16*)dgate; g.C2 = (__nv_bfloat16*)dup;
    g.in_gate = (const __nv_bfloat16*)gate; g.in_up = (const __nv_bfloat16*)up;
    g.ldc = ldo;
    g.M = M; g.N = H; g.K = K; g.nk1 = (K + BK - 1) / BK;
    return launch<EPI_SWIGLU_BWD, false, true>(gemm_cluster(w2, nullptr), ta, tb, ta, tb, g, max_ctas, stream);
}

// gate|up dual GEMM with the SwiGLU epilogue: x [M, K], w1 / w3 [H, K]; writes act (and gate, up if non-null), [M, H].
FTL_API int ftl_gemm_swiglu(const void* x, int64_t ldx, const void* w1, const void* w3, int64_t ldw, void* act,
                            void* gate, void* up, int64_t ldo, int M, int H, int K, int max_ctas, cudaStream_t stream) {
    if (M <= 0 || H <= 0 || K <= 0 || (H % 8) || (ldo % 8) || (ldx % 8) || (ldw % 8)) return (int)cudaErrorInvalidValue;
    CUtensorMap ta, tb, tb2;
    int e;
    if ((e = make_tmap(&ta, x, M, K, ldx, BK, BM))) return e;
    if ((e = make_tmap(&tb, w1, H, K, ldw, BK, BN / 2))) return e;
    if ((e = make_tmap(&tb2, w3, H, K, ldw, BK, BN / 2))) return e;
    GemmArgs g{};
    g.C = (__nv_bfloat16*)act; g.C_gate = (__nv_bfloat16*)gate; g.C_up = (__nv_bfloat16*)up;
    g.ldc = ldo;
    g.M = M; g.N = 2 * H; g.K = K;   // accumulator columns: 128 gate + 128 up per tile; ceil(2H/256) = ceil(H/128) tiles
    g.nk1 = (K + BK - 1) / BK;
    return launch<EPI_SWIGLU, false, false>(gemm_cluster(w1, w3), ta, tb, ta, tb2, g, max_ctas, stream);
}

// fused QKV projection + RoPE on the first rot_cols output columns (q heads then k heads).
FTL_API int ftl_gemm_qkv_rope(const void* x, int64_t ldx, const void* wqkv, int64_t ldw, void* qkv, int64_t ldo,
                              const float* cos_t, const float* sin_t, int M, int N, int K, int seq_len, int head_dim,
                              int rot_cols, int max_ctas, cudaStream_t stream) {
    if (M <= 0 || N <= 0 || K <= 0 || (N % 8) || (ldo % 8) || (ldx % 8) || (ldw % 8) || (head_dim % 8) || (rot_cols % 8))
        return (int)cudaErrorInvalidValue;
    CUtensorMap ta, tb;
    int e;
    if ((e = make_tmap(&ta, x, M, K, ldx, BK, BM))) return e;
    const int csel = gemm_cluster(wqkv, nullptr);
    if ((e = make_tmap(&tb, wqkv, N, K, ldw, BK, BN / (csel == 0 ? 2 : csel)))) return e;
    GemmArgs g{};
    g.C = (__nv_bfloat16*)qkv; g.ldc = ldo;
    g.M = M; g.N = N; g.K = K; g.nk1 = (K + BK - 1) / BK;
    g.cos_t = cos_t; g.sin_t = sin_t; g.seq_len = seq_len; g.head_dim = head_dim; g.rot_cols = rot_cols;
    return launch<EPI_ROPE, false, false>(csel, ta, tb, ta, tb, g, max_ctas, stream);
}


// ---- FSDP: the same two fused GEMMs with the parameter all-gather of their weight operand pulled into the kernel ----
// peer_shards: device array [world] of the owners' compact shard buffers; shard_off*: element offset of the matrix'
// shard in each of them; rows_per_rank: rows owned by each rank (multiple of 128).  The full [rows, K] matrix is
// assembled in `wqkv` / `w1`, `w3` (local staging) while the MMAs run.  scratch: >= 1 + n_mat * rows/128 ints.
static int gather_setup(GemmArgs& g, const void* const* peer_shards, int nmat, void* dst0, void* dst1, int64_t off0,
                        int64_t off1, int rows, int rows_per_rank, int rank, int world, int* scratch, cudaStream_t stream) {
    if (gemm_cluster_env() != 0 || rows_per_rank % 128 || rows % 128 || rows_per_rank * world != rows) return (int)cudaErrorInvalidValue;
    const int halves = rows / 128;
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(int) * (1 + nmat * halves), stream);
    if (e != cudaSuccess) return (int)e;
    g.gather_src = (const __nv_bfloat16* const*)peer_shards;
    g.gather_dst[0] = (__nv_bfloat16*)dst0; g.gather_dst[1] = (__nv_bfloat16*)dst1;
    g.gather_off[0] = off0; g.gather_off[1] = off1;
    g.gather_rows_per_rank = rows_per_rank; g.gather_nmat = nmat; g.rank = rank; g.world = world;
    g.gather_next = scratch; g.gather_done = scratch + 1;
    return 0;
}

FTL_API int ftl_gemm_qkv_rope_gather(const void* x, int64_t ldx, void* wqkv, int64_t ldw, void* qkv, int64_t ldo,
                                     const float* cos_t, const float* sin_t, int M, int N, int K, int seq_len, int head_dim,
                                     int rot_cols, const void* const* peer_shards, int64_t shard_off, int rows_per_rank,
                                     int rank, int world, int* scratch, cudaStream_t stream) {
    if (M <= 0 || N <= 0 || K <= 0 || (N % 128) || (ldo % 8) || (ldx % 8) || ldw != K || (head_dim % 8) || (rot_cols % 8))
        return (int)cudaErrorInvalidValue;
    CUtensorMap ta, tb;
    int e;
    if ((e = make_tmap(&ta, x, M, K, ldx, BK, BM))) return e;
    if ((e = make_tmap(&tb, wqkv, N, K, ldw, BK, BN / 2))) return e;
    GemmArgs g{};
    g.C = (__nv_bfloat16*)qkv; g.ldc = ldo;
    g.M = M; g.N = N; g.K = K; g.nk1 = (K + BK - 1) / BK;
    g.cos_t = cos_t; g.sin_t = sin_t; g.seq_len = seq_len; g.head_dim = head_dim; g.rot_cols = rot_cols;
    if ((e = gather_setup(g, peer_shards, 1, wqkv, nullptr, shard_off, 0, N, rows_per_rank, rank, world, scratch, stream))) return e;
    return launch_pair<EPI_ROPE, false, false>(ta, tb, ta, tb, g, 0, stream);
}

FTL_API int ftl_gemm_swiglu_gather(const void* x, int64_t ldx, void* w1, void* w3, int64_t ldw, void* act, void* gate,
                                   void* up, int64_t ldo, int M, int H, int K, const void* const* peer_shards,
                                   int64_t shard_off_w1, int64_t shard_off_w3, int rows_per_rank, int rank, int world,
                                   int* scratch, cudaStream_t stream) {
    if (M <= 0 || H <= 0 || K <= 0 || (H % 128) || (ldo % 8) || (ldx % 8) || ldw != K) return (int)cudaErrorInvalidValue;
    CUtensorMap ta, tb, tb2;
    int e;
    if ((e = make_tmap(&ta, x, M, K, ldx, BK, BM))) return e;
    if ((e = make_tmap(&tb, w1, H, K, ldw, BK, BN / 2))) return e;
    if ((e = make_tmap(&tb2, w3, H, K, ldw, BK, BN / 2))) return e;
    GemmArgs g{};
    g.C = (__nv_bfloat16*)act; g.C_gate = (__nv_bfloat16*)gate; g.C_up = (__nv_bfloat16*)up;
    g.ldc = ldo;
    g.M = M; g.N = 2 * H; g.K = K; g.nk1 = (K + BK - 1) / BK;
    if ((e = gather_setup(g, peer_shards, 2, w1, w3, shard_off_w1, shard_off_w3, H, rows_per_rank, rank, world, scratch, stream)))
        return e;
    return launch_pair<EPI_SWIGLU, false, false>(ta, tb, ta, tb2, g, 0, stream);
}
