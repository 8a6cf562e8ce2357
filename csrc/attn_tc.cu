// Causal grouped-query flash attention on tcgen05 (replaces K5 + K6 and their backward).
//
// The reference materialises repeat_kv and calls F.scaled_dot_product_attention(is_causal=True)
// (/root/reference/model.py:129-138, 204-212).  Here q, k and v are read in place from the fused
// [tokens, (H + 2 KV) * hd] projection buffer by TMA (no transposes, no repeat: query head h uses kv head
// h / (H / KV)), scores and outputs are accumulated by tcgen05.mma in TMEM, and the backward writes dq | dk | dv
// straight into the matching columns of the d(qkv) buffer with the inverse rotary embedding fused in.
// Every reduction has a fixed order and no atomics are used, so forward AND backward are bitwise reproducible —
// the library backward that this replaces accumulates dQ with atomics at sequence lengths >= 4096
// (tools/determinism_check.py), which breaks the bitwise kill/resume guarantee.
//
//   attn_fwd2_kernel  (S % 256 == 0) CTA = (batch, q head, TWO 128-query tiles); two softmax warpgroups (thread = whole
//                     query row) ping-pong under the MMAs of the other tile, O accumulates in TMEM with a lazy rescale
//                     done by a correction warpgroup, P is a TMEM A operand; dynamic causal work queue.
//   attn_fwd_kernel   (other S) CTA = (batch, q head, 128-query tile); per 128-key tile: S = Q K^T (TMEM) -> online
//                     softmax in registers (two threads per query row) -> P (bf16, TMEM) -> O_tile = P V (TMEM) ->
//                     rescale and add in registers.  Both write O (bf16) and the row log-sum-exp (fp32, base 2).
//   attn_bwd_dq       CTA = (batch, q head, 128-query tile); Q and dO are written once per item into TMEM as bf16 A
//                     operands; loops over 64-key tiles: S = Q K^T, dP = dO V^T -> dS (thread = query row, written over
//                     its own dP columns) -> dQ += dS K.  All three MMAs take A from TMEM; shared memory carries only
//                     the K/V ring.  While staging it also computes delta[t, h] = sum_d dO * O (the softmax-backward
//                     row term) and leaves it in global memory for attn_bwd_dkv, which runs after it.
//   attn_bwd_dkv      CTA = (batch, kv head, 128-key tile) or a chunk of its loop over the group's query heads and
//                     64-query tiles: S^T = K Q^T, dP^T = V dO^T (TMEM) -> P^T, dS^T (thread = key row, written over
//                     its own score columns) -> dV += P^T dO, dK += dS^T Q with A from TMEM.  Chunks leave fp32
//                     partials that attn_dkv_reduce folds in chunk order.
// Work items of all three kernels come from atomic queues, longest first (causal items differ 10x in length).
// Warp roles of the backward kernels and attn_fwd_kernel: warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2-9 = 256 math threads: TMEM lane quarter warp % 4 (a hardware rule), column half (warp - 2) / 4.
// How these shapes were chosen and what limits them: profiles/attn_tc_notes.md, profiles/umma_shapes.md.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cudaTypedefs.h>
#include <stdio.h>
#include <stdlib.h>

#include <mutex>
#include <type_traits>

#include "common/ptx.cuh"
#include "common/tmap_cache.cuh"
#include "common/util.cuh"

namespace ftl {

constexpr int kAttnThreads = 320;   // warp 0 TMA, warp 1 MMA, warps 2-9: two math threads per tile row (each owns half the columns)
constexpr float kLog2e = 1.4426950408889634f;

// ---- swizzled shared-memory tiles -----------------------------------------------------------------------------
// A [rows x cols] bf16 tile is stored as cols/64 chunks; chunk c holds columns [64c, 64c+64) as `rows` lines of
// 128 bytes with the 128B swizzle that TMA and UMMA use: 16-byte unit u of line r sits at unit (u ^ (r & 7)).
__device__ __forceinline__ uint32_t sw128_addr(uint32_t base, int rows, int r, int col) {
    const int c = col >> 6, u = (col & 63) >> 3;
    return base + c * rows * 128 + r * 128 + ((u ^ (r & 7)) << 4) + ((col & 7) << 1);
}
__device__ __forceinline__ void st_shared16(uint32_t addr, const bf16x8& v) {
    const int4& r = *reinterpret_cast<const int4*>(&v);
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w) : "memory");
}

// The MMA warp is ONE thread issuing every tcgen05.mma of the CTA; a 128x64x16 UMMA occupies the tensor pipe for only
// ~32 cycles, so building each 64-bit descriptor from scratch (~15 dependent integer ops) made instruction issue, not
// the tensor pipe, the limiter of these kernels.  Descriptors are therefore built once per operand tile (k16 = 0) and
// stepped by adding a compile-time constant to the 14-bit start-address field (addresses are < 256 KB: no carry).
// K-major operand tile [rows x kdim] (kdim contiguous, 64-column chunks of rows*128 bytes): step k16 -> +32 B inside a
// chunk, next chunk every 4 steps.
__device__ __forceinline__ uint64_t desc_kmajor0(uint32_t base) { return ptx::umma_smem_desc(base, 0, 1024); }
__device__ __forceinline__ uint64_t desc_kmajor_step(uint64_t d0, int rows, int k16) {
    return d0 + (uint64_t)(((k16 >> 2) * rows * 128 + (k16 & 3) * 32) >> 4);
}
// MN-major operand: tile stored [krows x mn] (mn contiguous, chunked by 64): step k16 covers rows 16*k16..+15.
__device__ __forceinline__ uint64_t desc_mnmajor0(uint32_t base, int krows) { return ptx::umma_smem_desc(base, krows * 128, 1024); }
__device__ __forceinline__ uint64_t desc_mnmajor_step(uint64_t d0, int k16) { return d0 + (uint64_t)((k16 * 16 * 128) >> 4); }
// D[tmem] (+)= A * B over KSTEPS UMMA steps; A K-major with ROWS_A rows, B either K-major (ROWS_B rows) or MN-major.
template <int KSTEPS, int ROWS_A, int ROWS_B, bool B_MN>
__device__ __forceinline__ void umma_tile(uint32_t tmem_d, uint32_t a_base, uint32_t b_base, uint32_t idesc, bool accumulate) {
    const uint64_t da = desc_kmajor0(a_base);
    const uint64_t db = B_MN ? desc_mnmajor0(b_base, ROWS_B) : desc_kmajor0(b_base);
#pragma unroll
    for (int k = 0; k < KSTEPS; ++k)
        ptx::umma_bf16(tmem_d, desc_kmajor_step(da, ROWS_A, k), B_MN ? desc_mnmajor_step(db, k) : desc_kmajor_step(db, ROWS_B, k),
                       idesc, (accumulate || k != 0) ? 1u : 0u);
}

struct AttnArgs {
    const float* cos_t;            // backward: inverse RoPE on dq / dk (may be null)
    const float* sin_t;
    __nv_bfloat16* out;            // fwd: O [T, H*hd];  bwd: d(qkv) [T, ld_qkv]
    float* lse;                    // [B, H, S]  log2-domain log-sum-exp of the scaled scores
    const float* delta;            // [B, H, S]  (backward)
    int64_t ld_out;
    int B, S, H, KV, ld_qkv;
    float scale_log2;              // 1/sqrt(hd) * log2(e)
    float scale;                   // 1/sqrt(hd)
    // dK/dV kernel work split: an item (batch, kv head, key tile) loops over rep * (#query tiles) inner iterations, from
    // 8 (last key tile) to 4 * S/64 (first); items longer than chunk_len are cut into chunks that run as separate
    // units and leave fp32 partials in `partial`, which attn_dkv_reduce folds in chunk order (deterministic).
    int chunk_len, units;
    float* partial;                // [units][2][128][hd] fp32
    const __nv_bfloat16* qkv_in;   // backward: the forward's fused projection and dO, read directly by the dQ kernel
    const __nv_bfloat16* dout_in;
    const __nv_bfloat16* o_in;     // forward output (delta = rowsum(dO * O) is computed by the dQ kernel)
    float* delta_out;
    int64_t ld_do;
};

// unit -> (batch, kv head, key tile, inner range [n0, n1), whether the unit is the whole item)
__device__ __forceinline__ void dkv_decode(const AttnArgs& a, int u, int rep, int q_tiles64, int k_tiles, int& b, int& g,
                                           int& kt, int& n0, int& n1, bool& whole) {
    const int bg = u % (a.B * a.KV);
    int v = u / (a.B * a.KV);
    b = bg / a.KV; g = bg % a.KV;
    kt = 0; n0 = 0; n1 = 0; whole = true;
    for (int k = 0; k < k_tiles; ++k) {           // key tile 0 has the longest loop: heavy units come first
        const int inner = rep * (q_tiles64 - 2 * k);
        const int nc = (inner + a.chunk_len - 1) / a.chunk_len;
        if (v < nc) {
            kt = k; n0 = v * a.chunk_len; n1 = n0 + a.chunk_len < inner ? n0 + a.chunk_len : inner; whole = nc == 1;
            return;
        }
        v -= nc;
    }
}

// D[tmem] (+)= A[tmem, bf16 pairs: 8 columns per UMMA step] * B[smem]
template <int KSTEPS, int ROWS_B, bool B_MN>
__device__ __forceinline__ void umma_tile_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_base, uint32_t idesc, bool accumulate) {
    const uint64_t db = B_MN ? desc_mnmajor0(b_base, ROWS_B) : desc_kmajor0(b_base);
#pragma unroll
    for (int k = 0; k < KSTEPS; ++k)
        ptx::umma_bf16_ts(tmem_d, tmem_a + k * 8, B_MN ? desc_mnmajor_step(db, k) : desc_kmajor_step(db, ROWS_B, k), idesc,
                          (accumulate || k != 0) ? 1u : 0u);
}

#ifdef FTL_ATTN_DEBUG
// Timestamp probe (tools/attn_probe.py builds a separate library with this defined): CTA 0 records clock64 at fixed
// points of its first item, per key tile.  Slots 0-7 math thread (warp 2 lane 0), 8-15 MMA thread.
__device__ long long g_attn_dbg[64 * 16];
#ifdef FTL_PROBE_DQ
#define ATTN_DBG(cond, tile, slot) do { } while (0)
#else
#define ATTN_DBG(cond, tile, slot) do { if (blockIdx.x == 0 && (cond) && (tile) < 64) g_attn_dbg[(tile) * 16 + (slot)] = clock64(); } while (0)
#endif
#ifdef FTL_PROBE_DQ     // the dQ and dK/dV kernels run in one call and share the array: probe one of them per build
#define DQ_DBG(cond, tile, slot) do { if (blockIdx.x == 0 && (cond) && (tile) < 64) g_attn_dbg[(tile) * 16 + (slot)] = clock64(); } while (0)
#define KV_DBG(cond, tile, slot) do { } while (0)
#else
#define KV_DBG(cond, tile, slot) do { if (blockIdx.x == 0 && (cond) && (tile) < 64) g_attn_dbg[(tile) * 16 + (slot)] = clock64(); } while (0)
#define DQ_DBG(cond, tile, slot) do { } while (0)
#endif
#else
#define ATTN_DBG(cond, tile, slot) do { } while (0)
#define KV_DBG(cond, tile, slot) do { } while (0)
#define DQ_DBG(cond, tile, slot) do { } while (0)
#endif

// ================================================================================================= forward
// Software pipeline (per CTA): the MMA warp issues S(j+1) as soon as K(j+1) has landed, i.e. while the math threads
// are still in the softmax of tile j; P and the O tile are double-buffered, and the math threads fold O(j-1) into
// their registers only after they have published P(j), so P V (j) runs under that accumulation and under softmax(j+1).
template <int HD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnArgs a) {
    constexpr int BM = 128, BN = 128;
    constexpr int kQBytes = BM * HD * 2, kKBytes = BN * HD * 2;
    constexpr int NKV = 3;                            // K/V stages (P lives in TMEM, which freed the room for a third)
    constexpr int kTmemO = 2 * BN;                    // columns: S0 [0,128) S1 [128,256) O0 [256,384) O1 [384,512)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kQBytes;                       // NKV stages
    uint8_t* sV = sK + NKV * kKBytes;                 // NKV stages
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + NKV * kKBytes);
    float* sMax = reinterpret_cast<float*>(bars + 20);   // [2][2][BM] row-max / row-sum exchange between the column halves
    uint64_t* q_full = bars;          // 1
    uint64_t* q_empty = bars + 1;     // 1
    uint64_t* kv_full = bars + 2;     // NKV (<= 4)
    uint64_t* kv_empty = bars + 6;    // NKV
    uint64_t* s_full = bars + 10;     // 2
    uint64_t* p_full = bars + 12;     // 2
    uint64_t* o_full = bars + 14;     // 2
    uint64_t* o_empty = bars + 16;    // 2
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_tiles = a.S / BM;
    const int items = a.B * a.H * q_tiles;
    const int rep = a.H / a.KV;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::mbar_init(q_full, 1); ptx::mbar_init(q_empty, 1);
        for (int i = 0; i < NKV; ++i) { ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&s_full[i], 1);
            ptx::mbar_init(&p_full[i], 8); ptx::mbar_init(&o_full[i], 1); ptx::mbar_init(&o_empty[i], 8);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    // item -> (b, h, qt), heaviest (largest qt) first so the tail of the grid is made of short items
    auto decode = [&](int it, int& b, int& h, int& qt) {
        qt = q_tiles - 1 - it / (a.B * a.H);
        const int r = it % (a.B * a.H);
        b = r / a.H; h = r % a.H;
    };

    if (warp == 0) {
        if (ptx::elect_one()) {
            int kv_it = 0, q_it = 0;
            for (int it = blockIdx.x; it < items; it += gridDim.x, ++q_it) {
                int b, h, qt; decode(it, b, h, qt);
                const int g = h / rep;
                const int row0 = b * a.S + qt * BM;
                ptx::mbar_wait(q_empty, (q_it & 1) ^ 1);
                ptx::mbar_arrive_expect_tx(q_full, kQBytes);
#pragma unroll
                for (int c = 0; c < HD / 64; ++c) ptx::tma_load_2d(sQ + c * BM * 128, &tmQKV, q_full, h * HD + c * 64, row0);
                for (int j = 0; j <= qt; ++j, ++kv_it) {
                    const int st = kv_it % NKV;
                    ptx::mbar_wait(&kv_empty[st], ((kv_it / NKV) & 1) ^ 1);
                    ptx::mbar_arrive_expect_tx(&kv_full[st], 2 * kKBytes);
                    const int krow = b * a.S + j * BN;
#pragma unroll
                    for (int c = 0; c < HD / 64; ++c) {
                        ptx::tma_load_2d(sK + st * kKBytes + c * BN * 128, &tmQKV, &kv_full[st], (a.H + g) * HD + c * 64, krow);
                        ptx::tma_load_2d(sV + st * kKBytes + c * BN * 128, &tmQKV, &kv_full[st], (a.H + a.KV + g) * HD + c * 64, krow);
                    }
                }
            }
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(BM, BN, false, false);
        constexpr uint32_t idesc_o = ptx::umma_idesc_bf16(BM, HD, false, true);   // B = V is MN-major
        int t_it = 0, q_it = 0;                                                    // t_it: global tile counter
        for (int it = blockIdx.x; it < items; it += gridDim.x, ++q_it) {
            int b, h, qt; decode(it, b, h, qt);
            ptx::mbar_wait(q_full, q_it & 1);
            const int nkv = qt + 1;
            auto issue_s = [&](int t) {                                            // scores of global tile t
                const int st = t & 1, ks = t % NKV;
                ptx::mbar_wait(&kv_full[ks], (t / NKV) & 1);
                ptx::tc_fence_after();
                ATTN_DBG(q_it == 0 && lane == 0, t, 8);
                if (ptx::elect_one()) {
                    umma_tile<HD / 16, BM, BN, false>(tmem + st * BN, ptx::smem_u32(sQ), ptx::smem_u32(sK + ks * kKBytes), idesc_s, false);
                    ptx::umma_commit(&s_full[st]);
                }
                __syncwarp();
                ATTN_DBG(q_it == 0 && lane == 0, t, 9);
            };
            issue_s(t_it);
            for (int j = 0; j < nkv; ++j, ++t_it) {
                if (j + 1 < nkv) issue_s(t_it + 1);                   // next scores while the softmax of this tile runs
                const int st = t_it & 1, ph = (t_it >> 1) & 1, ks = t_it % NKV;
                ptx::mbar_wait(&p_full[st], ph);                      // P(t) is in smem, S(t) has been read
                ATTN_DBG(q_it == 0 && lane == 0, t_it, 10);
                ptx::mbar_wait(&o_empty[st], ph ^ 1);                 // O buffer st (tile t-2) has been folded
                ptx::tc_fence_after();
                ATTN_DBG(q_it == 0 && lane == 0, t_it, 11);
                if (ptx::elect_one()) {
                    // A = P, written by the math threads over the first 64 columns of the score buffer it came from:
                    // only V is fetched from shared memory, which halves the operand traffic of this MMA
                    umma_tile_ts<BN / 16, BN, true>(tmem + kTmemO + st * BN, tmem + st * BN, ptx::smem_u32(sV + ks * kKBytes),
                                                    idesc_o, false);
                    ptx::umma_commit(&o_full[st]);
                    ptx::umma_commit(&kv_empty[ks]);
                    if (j == nkv - 1) ptx::umma_commit(q_empty);
                }
                __syncwarp();
                ATTN_DBG(q_it == 0 && lane == 0, t_it, 12);
            }
        }
    } else {
        const int q = warp & 3, half = (warp - 2) >> 2;                // TMEM lane quarter, column half
        const int r = q * 32 + lane;                                   // query row inside the tile = TMEM lane
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        constexpr int CW = BN / 2, OW = HD / 2;                        // score / output columns owned by this thread
        const int c0 = half * CW, d0 = half * OW;
        int t_it = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x) {
            int b, h, qt; decode(it, b, h, qt);
            const int qi = qt * BM + r;
            float m = -INFINITY, l = 0.f, alpha_prev = 0.f;
#ifdef FTL_ATTN_DEBUG
            const bool first = it == (int)blockIdx.x;
#endif
            float acc[OW];
#pragma unroll
            for (int d = 0; d < OW; ++d) acc[d] = 0.f;
            auto fold = [&](int t, float alpha) {                      // acc = acc * alpha + O(t)[:, d0 : d0 + OW]
                const int st = t & 1;
                ptx::mbar_wait(&o_full[st], (t >> 1) & 1);
                ptx::tc_fence_after();
                ATTN_DBG(first && threadIdx.x == 64, t, 6);
                uint32_t v[OW];
#pragma unroll
                for (int c = 0; c < OW; c += 32)
                    ptx::tmem_ld32(tmem + kTmemO + st * BN + lane_addr + d0 + c, *reinterpret_cast<uint32_t(*)[32]>(&v[c]));
                ptx::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < OW; ++i) acc[i] = fmaf(acc[i], alpha, __uint_as_float(v[i]));
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&o_empty[st]);
                ATTN_DBG(first && threadIdx.x == 64, t, 7);
            };
            for (int j = 0; j <= qt; ++j, ++t_it) {
                const int st = t_it & 1;
                ATTN_DBG(first && threadIdx.x == 64, t_it, 0);
                ptx::mbar_wait(&s_full[st], (t_it >> 1) & 1);
                ptx::tc_fence_after();
                ATTN_DBG(first && threadIdx.x == 64, t_it, 1);
                const uint32_t ts = tmem + st * BN + lane_addr + c0;
                const bool diag = j == qt;
                const int lim = diag ? r - c0 : CW;                     // local columns <= lim are visible
                // this thread's half of the score row stays in registers between the two passes
                uint32_t v[CW];
#pragma unroll
                for (int c = 0; c < CW; c += 32) ptx::tmem_ld32(ts + c, *reinterpret_cast<uint32_t(*)[32]>(&v[c]));
                ptx::tmem_ld_wait();
                ATTN_DBG(first && threadIdx.x == 64, t_it, 2);
                float mx = -INFINITY;
                if (!diag) {
#pragma unroll
                    for (int i = 0; i < CW; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
                } else {
#pragma unroll
                    for (int i = 0; i < CW; ++i) if (i <= lim) mx = fmaxf(mx, __uint_as_float(v[i]));
                }
                // row maximum = max over the two halves: exchange through shared memory
                float* xch = sMax + (t_it & 1) * 2 * BM;                 // double-buffered: one barrier per tile
                xch[half * BM + r] = mx;
                asm volatile("bar.sync 1, 256;" ::: "memory");
                mx = fmaxf(mx, xch[(half ^ 1) * BM + r]);
                ATTN_DBG(first && threadIdx.x == 64, t_it, 3);
                const float m_new = fmaxf(m, mx * a.scale_log2);
                const float alpha = exp2f(m - m_new);                   // m = -inf on the first tile -> 0
                float sum = 0.f;
                uint32_t pk[CW / 2];                                    // this thread's P values, two bf16 per register
#pragma unroll
                for (int i = 0; i < CW; i += 2) {
                    float e0 = exp2f(fmaf(__uint_as_float(v[i]), a.scale_log2, -m_new));
                    float e1 = exp2f(fmaf(__uint_as_float(v[i + 1]), a.scale_log2, -m_new));
                    if (diag && i > lim) e0 = 0.f;
                    if (diag && i + 1 > lim) e1 = 0.f;
                    sum += e0 + e1;
                    pk[i >> 1] = ptx::pack_bf16x2(e0, e1);
                }
                // P overwrites the score buffer in place (both column halves loaded their scores before the exchange
                // barrier above): columns [c0/2, c0/2 + 32) of this thread's TMEM lane
                ATTN_DBG(first && threadIdx.x == 64, t_it, 4);
                ptx::tmem_st32(tmem + st * BN + lane_addr + (c0 >> 1), pk);
                ptx::tmem_st_wait();
                l = fmaf(l, alpha, sum);                                // partial over this thread's columns
                m = m_new;
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&p_full[st]);
                ATTN_DBG(first && threadIdx.x == 64, t_it, 5);
                if (j > 0) fold(t_it - 1, alpha_prev);                  // runs while the tensor core does P V of this tile
                alpha_prev = alpha;
            }
            fold(t_it - 1, alpha_prev);
            // finalize: the two halves add their partial row sums, then O / l and the lse
            float* xch = sMax + (t_it & 1) * 2 * BM;                     // the buffer the next tile would use: free
            xch[half * BM + r] = l;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            l += xch[(half ^ 1) * BM + r];
            const float inv = 1.f / l;
            __nv_bfloat16* orow = a.out + (int64_t)(b * a.S + qi) * a.ld_out + h * HD + d0;
#pragma unroll
            for (int d = 0; d < OW; d += 16) {
                float o[8], o2[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) { o[i] = acc[d + i] * inv; o2[i] = acc[d + 8 + i] * inv; }
                st_row32(orow + d, pack8(o), pack8(o2));                 // full 32-byte sectors (see util.cuh)
            }
            if (half == 0) a.lse[((int64_t)b * a.H + h) * a.S + qi] = m + log2f(l);
            asm volatile("bar.sync 1, 256;" ::: "memory");              // exchange buffer free again for the next item
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem);
}

// ================================================================================================= forward, two query tiles per CTA
// The timeline of attn_fwd_kernel (tools/attn_probe.py, profiles/attn_tc_notes.md) shows its math threads busy for
// ~2850 cycles per 128x128 tile while the tensor pipe needs ~1024: exponentials (MUFU, 16/clk/SM -> 1024 cycles per
// tile), the row-max exchange barrier between the two column halves and the per-tile fold of O into registers all sit
// on one serial chain.  This kernel removes the chain instead of shortening it:
//   * a CTA owns TWO consecutive query tiles A, B of one head (256 rows) which share every K/V tile; softmax warpgroup
//     A (warps 0-3) and B (warps 4-7) each own whole rows of their tile (no exchange) and run half a step out of
//     phase, so one group's MUFU phase overlaps the other's TMEM traffic and the tensor pipe alternates
//     P_A V, Q_A K^T(next), P_B V, Q_B K^T(next);
//   * O accumulates in TMEM across key tiles (accumulate flag); rescaling is LAZY: a row changes its reference maximum
//     only when the running maximum grew by more than 2^8, so the correction warpgroup (warps 8-11) almost never has
//     to touch O and the softmax warps never do.  P <= 2^8 is harmless in bf16 (same relative precision) and the row
//     sum uses the same reference, so lse = m_ref + log2(l) is exact;
//   * the correction warpgroup also does the epilogue (O / l -> bf16 -> global), off the softmax warps' path;
//   * work items are handed out by an atomic counter, heaviest (latest query tiles) first: causal items differ 10x in
//     length and a static round-robin leaves a third of the SMs idle at the end.
// TMEM columns: S_A [0,128) S_B [128,256) (P overwrites the first 64 columns of its S), O_A at 256, O_B at 256 + HD.
// The scheduler words are file-scope and reset by the last CTA to drain the queue, so launches of this kernel must be
// ordered on ONE stream per device (they are: attention runs on the compute stream only).
constexpr int kFwd2Threads = 512;   // warps 0-3 softmax A, 4-7 softmax B, 8-11 correction + epilogue, 12 TMA, 13 MMA, 14-15 idle
// Register budgets after setmaxnreg (per thread; 64K per SM = 512 x 128 at launch): the softmax warpgroups hold a whole
// score row, the others give registers up.  2 x 128 x 176 + 128 x 80 + 128 x 80 = 65536.
constexpr int kRegSoftmax = 176, kRegCorr = 80, kRegIssue = 80;
constexpr float kRescaleThreshold = 8.f;
constexpr int kPolyEvery = 2;       // one pair of exponentials in kPolyEvery goes to the FMA pipe
__device__ int g_fwd2_sched[2];     // [0] next work item, [1] CTAs that have seen the end of the queue

template <int HD>
__global__ void __launch_bounds__(kFwd2Threads, 1)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnArgs a) {
    constexpr int BM = 128, BN = 128;
    constexpr int kTile = BN * HD * 2;                // bytes of one Q, K or V tile
    constexpr int NST = 5;                            // K/V ring: K(0) V(0) K(1) V(1) ...
    constexpr int kTmemO = 2 * BN;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;                               // Q_A, Q_B
    uint8_t* sKV = sQ + 2 * kTile;                    // NST tiles
    uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + NST * kTile);
    uint64_t* q_full = bars;            // 1
    uint64_t* q_empty = bars + 1;       // 1
    uint64_t* kv_full = bars + 2;       // NST (<= 6)
    uint64_t* kv_empty = bars + 8;      // NST
    uint64_t* s_full = bars + 14;       // [2]  per query tile X
    uint64_t* p_full = bars + 16;       // [2]
    uint64_t* o_ready = bars + 18;      // [2]  O_X may be accumulated into (rescaled / drained)
    uint64_t* o_full = bars + 20;       // [2]  P_X V of a step has completed
    uint64_t* alpha_full = bars + 22;   // [2]
    uint64_t* l_full = bars + 24;       // [2]
    float* sAlpha = reinterpret_cast<float*>(bars + 26);   // [2][BM]
    float* sL = sAlpha + 2 * BM;                            // [2][BM]
    int* sItem = reinterpret_cast<int*>(sL + 2 * BM);       // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sItem + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_pairs = a.S / (2 * BM);
    const int items = a.B * a.H * q_pairs;
    const int rep = a.H / a.KV;

    if (warp == 13 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::mbar_init(q_full, 1); ptx::mbar_init(q_empty, 1);
        for (int i = 0; i < NST; ++i) { ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&p_full[i], 4); ptx::mbar_init(&o_ready[i], 4);
            ptx::mbar_init(&o_full[i], 1); ptx::mbar_init(&alpha_full[i], 4); ptx::mbar_init(&l_full[i], 4);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 12) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    // item -> (b, h, pair): latest pairs (longest loops) first; neighbouring items are the heads of one kv group
    auto decode = [&](int it, int& b, int& h, int& p) {
        p = q_pairs - 1 - it / (a.B * a.H);
        const int r = it % (a.B * a.H);
        b = r / a.H; h = r % a.H;
    };

    if (warp >= 12) {
        ptx::reg_dec<kRegIssue>();
        if (warp == 12) {
        // ------------------------------------------------------------------------------------------ TMA producer
        if (ptx::elect_one()) {
            int kv_n = 0;
            for (int q_it = 0;; ++q_it) {
                const int it = atomicAdd(&g_fwd2_sched[0], 1);
                ptx::mbar_wait_sleep(q_empty, (q_it & 1) ^ 1);           // Q tiles and the item slot of two items ago are free
                if (it >= items) { sItem[q_it & 1] = -1; ptx::mbar_arrive(q_full); break; }
                sItem[q_it & 1] = it;
                int b, h, p; decode(it, b, h, p);
                const int g = h / rep;
                const int row0 = b * a.S + p * 2 * BM;
                ptx::mbar_arrive_expect_tx(q_full, 2 * kTile);
#pragma unroll
                for (int x = 0; x < 2; ++x)
#pragma unroll
                    for (int c = 0; c < HD / 64; ++c)
                        ptx::tma_load_2d(sQ + x * kTile + c * BM * 128, &tmQKV, q_full, h * HD + c * 64, row0 + x * BM);
                const int nB = 2 * p + 2;
                for (int j = 0; j < nB; ++j) {
                    const int krow = b * a.S + j * BN;
#pragma unroll
                    for (int kv = 0; kv < 2; ++kv, ++kv_n) {       // K(j) then V(j)
                        const int st = kv_n % NST;
                        ptx::mbar_wait_sleep(&kv_empty[st], ((kv_n / NST) & 1) ^ 1);
                        ptx::mbar_arrive_expect_tx(&kv_full[st], kTile);
                        const int col = (a.H + kv * a.KV + g) * HD;
#pragma unroll
                        for (int c = 0; c < HD / 64; ++c)
                            ptx::tma_load_2d(sKV + st * kTile + c * BN * 128, &tmQKV, &kv_full[st], col + c * 64, krow);
                    }
                }
            }
            __threadfence();
            if (atomicAdd(&g_fwd2_sched[1], 1) == (int)gridDim.x - 1) {   // every CTA has fetched its last item
                g_fwd2_sched[0] = 0; g_fwd2_sched[1] = 0;
                __threadfence();
            }
        }
        } else if (warp == 13) {
        // ------------------------------------------------------------------------------------------ MMA issuer
        constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(BM, BN, false, false);
        constexpr uint32_t idesc_o = ptx::umma_idesc_bf16(BM, HD, false, true);   // B = V is MN-major
        int cx[2] = {0, 0};                                                        // tiles done per query tile (phases)
        int kv0 = 0;                                                               // ring index of K(0) of this item
        for (int q_it = 0;; ++q_it) {
            ptx::mbar_wait(q_full, q_it & 1);
            const int it = sItem[q_it & 1];
            if (it < 0) break;
            int b, h, p; decode(it, b, h, p);
            const int nB = 2 * p + 2;
            auto issue_s = [&](int x, int j) {                                     // S_X = Q_X K(j)^T
                const int kn = kv0 + 2 * j, ks = kn % NST;
                ptx::mbar_wait(&kv_full[ks], (kn / NST) & 1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    umma_tile<HD / 16, BM, BN, false>(tmem + x * BN, ptx::smem_u32(sQ + x * kTile), ptx::smem_u32(sKV + ks * kTile),
                                                      idesc_s, false);
                    ptx::umma_commit(&s_full[x]);
                    if (x == 1) {
                        ptx::umma_commit(&kv_empty[ks]);                           // B is the last reader of every K tile
                        if (j == nB - 1) ptx::umma_commit(q_empty);                // ... and of the Q tiles
                    }
                }
                __syncwarp();
            };
            issue_s(0, 0);
            issue_s(1, 0);
            for (int j = 0; j < nB; ++j) {
#pragma unroll
                for (int x = 0; x < 2; ++x) {
                    const int nx = nB - 1 + x;                                     // A: 2p + 1 key tiles, B: 2p + 2
                    if (j >= nx) continue;
                    const int ph = cx[x] & 1;
                    const int vn = kv0 + 2 * j + 1, vs = vn % NST;
                    ptx::mbar_wait(&kv_full[vs], (vn / NST) & 1);
                    ptx::mbar_wait(&o_ready[x], ph);                               // O_X rescaled (or drained, j = 0)
                    ptx::mbar_wait(&p_full[x], ph);                                // P_X(j) is in TMEM
                    ptx::tc_fence_after();
                    ATTN_DBG(q_it == 0 && lane == 0 && x == 0, j, 10);
                    if (ptx::elect_one()) {
                        umma_tile_ts<BN / 16, BN, true>(tmem + kTmemO + x * HD, tmem + x * BN, ptx::smem_u32(sKV + vs * kTile),
                                                        idesc_o, j > 0);
                        ptx::umma_commit(&o_full[x]);
                        if (x == 1) ptx::umma_commit(&kv_empty[vs]);
                    }
                    __syncwarp();
                    ++cx[x];
                    if (j + 1 < nx) issue_s(x, j + 1);                             // executes after P_X V: S_X may be overwritten
                    ATTN_DBG(q_it == 0 && lane == 0 && x == 0, j, 11);
                }
            }
            kv0 += 2 * nB;
        }
        }
    } else if (warp < 8) {
        // ------------------------------------------------------------------------------------------ softmax A / B
        ptx::reg_inc<kRegSoftmax>();
        const int x = warp >> 2, q = warp & 3;
        const int r = q * 32 + lane;                                   // query row inside the tile = TMEM lane
        const uint32_t ts = tmem + x * BN + ((uint32_t)(q * 32) << 16);
        int c = 0;
        for (int q_it = 0;; ++q_it) {
            ptx::mbar_wait(q_full, q_it & 1);
            const int it = sItem[q_it & 1];
            if (it < 0) break;
            int b, h, p; decode(it, b, h, p);
            const int qt = 2 * p + x;
            float m = -INFINITY, l = 0.f;
            for (int j = 0; j <= qt; ++j, ++c) {
                ATTN_DBG(q_it == 0 && threadIdx.x == 0, j, 0);
                ptx::mbar_wait(&s_full[x], c & 1);
                ptx::tc_fence_after();
                ATTN_DBG(q_it == 0 && threadIdx.x == 0, j, 1);
                ATTN_DBG(q_it == 0 && threadIdx.x == 128, j, 5);
                // The whole score row lives in registers (this warpgroup raised its register budget for it): one TMEM
                // round trip (~150 cycles) per tile instead of one per chunk and pass.  The diagonal tile (columns <= r
                // visible) is a separate instantiation so that the other tiles carry no compare / select per element.
                float alpha = 1.f, sum = 0.f;
                auto tile = [&](auto diag_c) {
                    constexpr bool DIAG = decltype(diag_c)::value;
                    uint32_t v[BN];
#pragma unroll
                    for (int cb = 0; cb < BN; cb += 32) ptx::tmem_ld32(ts + cb, *reinterpret_cast<uint32_t(*)[32]>(&v[cb]));
                    ptx::tmem_ld_wait();
                    // row maximum, four independent chains (one dependent chain of 128 costs ~500 cycles)
                    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
                    for (int i = 0; i < BN; ++i)
                        if (!DIAG || i <= r) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(v[i]));
                    const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
                    const float m_new = fmaxf(m, mx * a.scale_log2);
                    if (j == 0) m = m_new;                               // O is overwritten by the first P V: no rescale
                    else if (m_new - m > kRescaleThreshold) { alpha = ptx::ex2(m - m_new); m = m_new; }
                    if (j > 0) {
                        sAlpha[x * BM + r] = alpha;
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(&alpha_full[x]);
                    }
                    ATTN_DBG(q_it == 0 && threadIdx.x == 0, j, 2);
                    ATTN_DBG(q_it == 0 && threadIdx.x == 128, j, 6);
                    // P = exp2(s * scale - m): packed f32x2 multiply-add and add (two elements per issue slot), MUFU.EX2,
                    // bf16 pairs written over the scores (all of them are in registers by now)
                    const uint64_t sc2 = ptx::pack_f32x2(a.scale_log2, a.scale_log2), nm2 = ptx::pack_f32x2(-m, -m);
                    uint64_t sum2[4] = {0ull, 0ull, 0ull, 0ull};
                    // Every second pair of exponentials is evaluated on the FMA pipe (ptx::exp2_poly_x2) instead of the
                    // special-function unit: MUFU.EX2 delivers 16 results / clk / SM, which for a 128 x 128 tile is as
                    // long as its two MMAs (1024 cycles), and the polynomial pairs fill the issue slots in which a
                    // MUFU result is still in flight.
#pragma unroll
                    for (int cb = 0; cb < BN; cb += 32) {
                        uint32_t pk[16];
#pragma unroll
                        for (int i = 0; i < 32; i += 2) {
                            const uint64_t x2 = ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(v[cb + i]), __uint_as_float(v[cb + i + 1])), sc2, nm2);
                            float e0, e1;
                            if ((i >> 1) % kPolyEvery == kPolyEvery - 1) {
                                ptx::exp2_poly_x2(x2, e0, e1);
                            } else {
                                float x0, x1;
                                ptx::unpack_f32x2(x2, x0, x1);
                                e0 = ptx::ex2(x0); e1 = ptx::ex2(x1);
                            }
                            if (DIAG) {
                                if (cb + i > r) e0 = 0.f;
                                if (cb + i + 1 > r) e1 = 0.f;
                            }
                            sum2[(i >> 1) & 3] = ptx::add_f32x2(sum2[(i >> 1) & 3], ptx::pack_f32x2(e0, e1));
                            pk[i >> 1] = ptx::pack_bf16x2(e0, e1);
                        }
                        ptx::tmem_st16(ts + (cb >> 1), pk);
                    }
                    float s0, s1;
                    ptx::unpack_f32x2(ptx::add_f32x2(ptx::add_f32x2(sum2[0], sum2[1]), ptx::add_f32x2(sum2[2], sum2[3])), s0, s1);
                    sum = s0 + s1;
                };
                if (j == qt) tile(std::true_type{}); else tile(std::false_type{});
                ptx::tmem_st_wait();
                ATTN_DBG(q_it == 0 && threadIdx.x == 0, j, 3);
                ATTN_DBG(q_it == 0 && threadIdx.x == 128, j, 7);
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&p_full[x]);
                l = fmaf(l, alpha, sum);
                ATTN_DBG(q_it == 0 && threadIdx.x == 0, j, 4);
                ATTN_DBG(q_it == 0 && threadIdx.x == 128, j, 8);
            }
            sL[x * BM + r] = l;
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&l_full[x]);
            a.lse[((int64_t)b * a.H + h) * a.S + qt * BM + r] = m + log2f(l);
        }
    } else {
        // ------------------------------------------------------------------------------------------ correction + epilogue
        ptx::reg_dec<kRegCorr>();
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        int cx[2] = {0, 0}, ax[2] = {0, 0};
        for (int q_it = 0;; ++q_it) {
            ptx::mbar_wait_sleep(q_full, q_it & 1);
            const int it = sItem[q_it & 1];
            if (it < 0) break;
            int b, h, p; decode(it, b, h, p);
            const int nB = 2 * p + 2;
            for (int j = 0; j < nB; ++j) {
#pragma unroll
                for (int x = 0; x < 2; ++x) {
                    if (j >= nB - 1 + x) continue;
                    if (j > 0) {
                        ptx::mbar_wait_sleep(&alpha_full[x], ax[x] & 1);
                        ++ax[x];
                        const float alpha = sAlpha[x * BM + r];
                        if (__any_sync(0xffffffffu, alpha != 1.f)) {
                            ptx::mbar_wait_sleep(&o_full[x], (cx[x] - 1) & 1);            // P_X V (j-1) has completed
                            ptx::tc_fence_after();
                            const uint32_t to = tmem + kTmemO + x * HD + lane_addr;
#pragma unroll
                            for (int d = 0; d < HD; d += 32) {
                                uint32_t v[32];
                                ptx::tmem_ld32(to + d, v);
                                ptx::tmem_ld_wait();
#pragma unroll
                                for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
                                ptx::tmem_st32(to + d, v);
                            }
                            ptx::tmem_st_wait();
                        }
                    }
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(&o_ready[x]);
                    ++cx[x];
                }
            }
#pragma unroll
            for (int x = 0; x < 2; ++x) {
                ptx::mbar_wait_sleep(&o_full[x], (cx[x] - 1) & 1);
                ptx::mbar_wait_sleep(&l_full[x], q_it & 1);
                ptx::tc_fence_after();
                const float inv = 1.f / sL[x * BM + r];
                const uint32_t to = tmem + kTmemO + x * HD + lane_addr;
                __nv_bfloat16* orow = a.out + (int64_t)(b * a.S + (2 * p + x) * BM + r) * a.ld_out + h * HD;
#pragma unroll
                for (int d = 0; d < HD; d += 32) {
                    uint32_t v[32];
                    ptx::tmem_ld32(to + d, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; i += 16) {
                        float o[8], o2[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k) { o[k] = __uint_as_float(v[i + k]) * inv; o2[k] = __uint_as_float(v[i + 8 + k]) * inv; }
                        st_row32(orow + d + i, pack8(o), pack8(o2));     // full 32-byte sectors (see util.cuh)
                    }
                }
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 12) ptx::tmem_dealloc<512>(tmem);
}

// inverse rotary embedding of 8 consecutive gradient values (4 pairs) of a q / k row at position pos
// Same for 16 consecutive values (two groups) with ONE 32-byte load per table: the epilogue threads each own a row
// (= a position), so every table load of a warp touches 32 different lines and the number of load instructions is what
// the fused inverse RoPE costs (it added ~30 us per backward kernel at the model shape with 16-byte loads).
__device__ __forceinline__ void unrope16(float (&f)[8], float (&f2)[8], const AttnArgs& a, int pos, int d0, int hd) {
    if (!a.cos_t) return;
    const int64_t to = (int64_t)pos * (hd >> 1) + (d0 >> 1);
    uint32_t cu[8], su[8];
    ld_row32(a.cos_t + to, cu);
    ld_row32(a.sin_t + to, su);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float c0 = __uint_as_float(cu[j]), s0 = __uint_as_float(su[j]);
        const float c1 = __uint_as_float(cu[4 + j]), s1 = __uint_as_float(su[4 + j]);
        const float x = f[2 * j], y = f[2 * j + 1], x2 = f2[2 * j], y2 = f2[2 * j + 1];
        f[2 * j] = x * c0 + y * s0;                // rotation by -angle
        f[2 * j + 1] = -x * s0 + y * c0;
        f2[2 * j] = x2 * c1 + y2 * s1;
        f2[2 * j + 1] = -x2 * s1 + y2 * c1;
    }
}
__device__ __forceinline__ void unrope8(float (&f)[8], const AttnArgs& a, int pos, int d0, int hd) {
    if (!a.cos_t) return;
    const int64_t to = (int64_t)pos * (hd >> 1) + (d0 >> 1);
    const float4 cs = *reinterpret_cast<const float4*>(a.cos_t + to);
    const float4 sn = *reinterpret_cast<const float4*>(a.sin_t + to);
    const float cc[4] = {cs.x, cs.y, cs.z, cs.w}, ss[4] = {sn.x, sn.y, sn.z, sn.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float x = f[2 * j], y = f[2 * j + 1];
        f[2 * j] = x * cc[j] + y * ss[j];          // rotation by -angle
        f[2 * j + 1] = -x * ss[j] + y * cc[j];
    }
}

// Work queues of the two backward kernels (same scheme as g_fwd2_sched): [0..1] dK/dV units, [2..3] dQ items.
// The last CTA to see the end of a queue resets it, so launches must be stream-ordered per device (they are).
__device__ int g_bwd_sched[4];
__device__ __forceinline__ void sched_done(int* q) {
    __threadfence();
    if (atomicAdd(q + 1, 1) == (int)gridDim.x - 1) {
        q[0] = 0; q[1] = 0;
        __threadfence();
    }
}

// ================================================================================================= backward: dK, dV
template <int HD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmQKV64,
                    const __grid_constant__ CUtensorMap tmDO64, const AttnArgs a) {
    constexpr int BK = 128, BQ = 64;                   // key tile (UMMA M), query tile (UMMA N)
    constexpr int kKBytes = BK * HD * 2, kQBytes = BQ * HD * 2;
    constexpr int kTmDP = 64, kTmDV = 256, kTmDK = 256 + HD;   // S / dP of buffer b at b*128 and b*128 + 64
    constexpr int NQ = 4;      // Q / dO stages: the TMA loads run three iterations ahead of the MMAs that consume them
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sK = smem;
    uint8_t* sV = sK + kKBytes;
    uint8_t* sQ = sV + kKBytes;                        // NQ stages of [BQ x HD]
    uint8_t* sDO = sQ + NQ * kQBytes;                  // NQ stages
    float* sLse = reinterpret_cast<float*>(sDO + NQ * kQBytes);   // 2 x BQ (P^T and dS^T live in TMEM, not here)
    float* sDelta = sLse + 2 * BQ;                            // 2 x BQ
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDelta + 2 * BQ);
    uint64_t* kv_full = bars;        // 1
    uint64_t* kv_empty = bars + 1;   // 1
    uint64_t* q_full = bars + 2;     // NQ (<= 4)
    uint64_t* q_empty = bars + 6;    // NQ
    uint64_t* s_full = bars + 10;    // 2  (S^T and dP^T of buffer b ready)
    uint64_t* t_full = bars + 12;    // 2  (P^T, dS^T of buffer b written over its S^T / dP^T in TMEM)
    uint64_t* acc_full = bars + 16;  // 1
    uint64_t* acc_empty = bars + 17; // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);
    int* sItem = reinterpret_cast<int*>(bars + 19);    // [2] unit handed out by the scheduler, published with kv_full

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k_tiles = a.S / BK, rep = a.H / a.KV;
    const int items = a.units;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::mbar_init(kv_full, 1); ptx::mbar_init(kv_empty, 1);
        for (int i = 0; i < NQ; ++i) { ptx::mbar_init(&q_full[i], 1); ptx::mbar_init(&q_empty[i], 1); }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&t_full[i], 8);
        }
        ptx::mbar_init(acc_full, 1); ptx::mbar_init(acc_empty, 8);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    // inner iteration n of an item: query head g*rep + n / nq, 64-query tile 2*kt + n % nq, nq = S/64 - 2*kt
    const int q_tiles64 = a.S / BQ;

    if (warp == 0) {
        if (ptx::elect_one()) {
            // Units are handed out by an atomic counter, longest first: with a static round-robin some CTAs drew three
            // units and others two (ncu: SMs idle for a third of the kernel).  The unit index travels to the other
            // warps through sItem, published by the arrival on kv_full.
            int q_it = 0;
            for (int item_it = 0;; ++item_it) {
                const int it = atomicAdd(&g_bwd_sched[0], 1);
                ptx::mbar_wait(kv_empty, (item_it & 1) ^ 1);
                if (it >= items) { sItem[item_it & 1] = -1; ptx::mbar_arrive(kv_full); break; }
                sItem[item_it & 1] = it;
                int b, g, kt, n0, n1; bool whole;
                dkv_decode(a, it, rep, q_tiles64, k_tiles, b, g, kt, n0, n1, whole);
                ptx::mbar_arrive_expect_tx(kv_full, 2 * kKBytes);
                const int krow = b * a.S + kt * BK;
#pragma unroll
                for (int c = 0; c < HD / 64; ++c) {
                    ptx::tma_load_2d(sK + c * BK * 128, &tmQKV, kv_full, (a.H + g) * HD + c * 64, krow);
                    ptx::tma_load_2d(sV + c * BK * 128, &tmQKV, kv_full, (a.H + a.KV + g) * HD + c * 64, krow);
                }
                const int nq = q_tiles64 - 2 * kt;
                int hh = n0 / nq, qo = n0 % nq;
                for (int n = n0; n < n1; ++n, ++q_it) {
                    const int h = g * rep + hh, qt = 2 * kt + qo;
                    if (++qo == nq) { qo = 0; ++hh; }
                    const int st = q_it % NQ;
                    ptx::mbar_wait(&q_empty[st], ((q_it / NQ) & 1) ^ 1);
                    ptx::mbar_arrive_expect_tx(&q_full[st], 2 * kQBytes);
                    const int qrow = b * a.S + qt * BQ;
#pragma unroll
                    for (int c = 0; c < HD / 64; ++c) {
                        ptx::tma_load_2d(sQ + st * kQBytes + c * BQ * 128, &tmQKV64, &q_full[st], h * HD + c * 64, qrow);
                        ptx::tma_load_2d(sDO + st * kQBytes + c * BQ * 128, &tmDO64, &q_full[st], h * HD + c * 64, qrow);
                    }
                }
            }
            sched_done(&g_bwd_sched[0]);
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_st = ptx::umma_idesc_bf16(BK, BQ, false, false);   // S^T = K Q^T, dP^T = V dO^T
        constexpr uint32_t idesc_acc = ptx::umma_idesc_bf16(BK, HD, false, true);   // dV = P^T dO, dK = dS^T Q (B MN-major)
        int item_it = 0, n_it = 0;                                      // n_it: global inner-iteration counter
        const uint32_t ka = ptx::smem_u32(sK), va = ptx::smem_u32(sV);
        // S^T / dP^T of inner iteration t go to TMEM buffer t & 1 and read Q / dO stage t & 1
        auto issue_sdp = [&](int t) {
            const int st = t & 1, qs = t % NQ;
            ptx::mbar_wait(&q_full[qs], (t / NQ) & 1);
            ptx::tc_fence_after();
            KV_DBG(item_it == 0 && lane == 0, t, 8);
            if (ptx::elect_one()) {
                const uint32_t qb = ptx::smem_u32(sQ + qs * kQBytes), dob = ptx::smem_u32(sDO + qs * kQBytes);
                umma_tile<HD / 16, BK, BQ, false>(tmem + st * 128, ka, qb, idesc_st, false);
                umma_tile<HD / 16, BK, BQ, false>(tmem + st * 128 + kTmDP, va, dob, idesc_st, false);
                ptx::umma_commit(&s_full[st]);
            }
            __syncwarp();
        };
        for (;; ++item_it) {
            ptx::mbar_wait(kv_full, item_it & 1);
            const int it = sItem[item_it & 1];
            if (it < 0) break;
            int b, g, kt, n0, n1; bool whole;
            dkv_decode(a, it, rep, q_tiles64, k_tiles, b, g, kt, n0, n1, whole);
            ptx::mbar_wait(acc_empty, (item_it & 1) ^ 1);
            const int inner = n1 - n0;
            issue_sdp(n_it);
            for (int n = 0; n < inner; ++n, ++n_it) {
                if (n + 1 < inner) issue_sdp(n_it + 1);               // next scores run under the math of this iteration
                const int st = n_it & 1;
                KV_DBG(item_it == 0 && lane == 0, n, 9);
                ptx::mbar_wait(&t_full[st], (n_it >> 1) & 1);
                ptx::tc_fence_after();
                KV_DBG(item_it == 0 && lane == 0, n, 10);
                if (ptx::elect_one()) {
                    const int qs = n_it % NQ;
                    const uint32_t qb = ptx::smem_u32(sQ + qs * kQBytes), dob = ptx::smem_u32(sDO + qs * kQBytes);
                    // A = P^T / dS^T from TMEM: queries 0-31 sit in columns [0,16) of the score buffer, 32-63 in [32,48)
                    const uint64_t d_do = desc_mnmajor0(dob, BQ), d_q = desc_mnmajor0(qb, BQ);
#pragma unroll
                    for (int k = 0; k < BQ / 16; ++k)
                        ptx::umma_bf16_ts(tmem + kTmDV, tmem + st * 128 + (k >> 1) * 32 + (k & 1) * 8, desc_mnmajor_step(d_do, k), idesc_acc,
                                          (n != 0 || k != 0) ? 1u : 0u);
#pragma unroll
                    for (int k = 0; k < BQ / 16; ++k)
                        ptx::umma_bf16_ts(tmem + kTmDK, tmem + st * 128 + kTmDP + (k >> 1) * 32 + (k & 1) * 8, desc_mnmajor_step(d_q, k), idesc_acc,
                                          (n != 0 || k != 0) ? 1u : 0u);
                    ptx::umma_commit(&q_empty[qs]);
                    if (n == inner - 1) { ptx::umma_commit(acc_full); ptx::umma_commit(kv_empty); }
                }
                __syncwarp();
                KV_DBG(item_it == 0 && lane == 0, n, 11);
            }
        }
    } else {
        const int q4 = warp & 3;
        const int r = q4 * 32 + lane;                                  // key row inside the tile = TMEM lane
        const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
        const int tid = threadIdx.x - 64;                              // 0..255
        const int half = (warp - 2) >> 2;                              // which 32 of the 64 query columns / which half of hd
        const int c0 = half * 32;
        uint32_t st_off[4];                                            // swizzled offsets of this thread's four 16-byte stores
#pragma unroll
        for (int u = 0; u < 4; ++u) st_off[u] = r * 128 + (((half * 4 + u) ^ (r & 7)) << 4);
        const uint64_t scale2 = ptx::pack_f32x2(a.scale_log2, a.scale_log2);
        int item_it = 0, n_it = 0;
        for (;; ++item_it) {
            ptx::mbar_wait(kv_full, item_it & 1);
            const int it = sItem[item_it & 1];
            if (it < 0) break;
            int b, g, kt, n0, n1; bool whole;
            dkv_decode(a, it, rep, q_tiles64, k_tiles, b, g, kt, n0, n1, whole);
            const int kj = kt * BK + r;
            const int nq = q_tiles64 - 2 * kt;
            {
                // lse / delta of the NEXT iteration are fetched into registers while this one computes, so the
                // global-load latency is off the critical path of the staging barrier
                int hh = n0 / nq, qo = n0 % nq;                         // head within the group, query tile offset
                float lse_r = 0.f, del_r = 0.f;
                if (tid < BQ) {
                    const int64_t o = ((int64_t)b * a.H + g * rep + hh) * a.S + (2 * kt + qo) * BQ + tid;
                    lse_r = a.lse[o]; del_r = a.delta[o];
                }
                for (int n = n0; n < n1; ++n, ++n_it) {
                    const int qt = 2 * kt + qo;
                    float* lse_s = sLse + (n_it & 1) * BQ;
                    float* del_s = sDelta + (n_it & 1) * BQ;
                    if (tid < BQ) { lse_s[tid] = -lse_r; del_s[tid] = -del_r; }   // negated: operands of packed adds
                    if (++qo == nq) { qo = 0; ++hh; }
                    if (tid < BQ && n + 1 < n1) {
                        const int64_t o = ((int64_t)b * a.H + g * rep + hh) * a.S + (2 * kt + qo) * BQ + tid;
                        lse_r = a.lse[o]; del_r = a.delta[o];
                    }
                    KV_DBG(item_it == 0 && threadIdx.x == 64, n - n0, 0);
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    KV_DBG(item_it == 0 && threadIdx.x == 64, n - n0, 1);
                    const int st = n_it & 1;
                    ptx::mbar_wait(&s_full[st], (n_it >> 1) & 1);
                    KV_DBG(item_it == 0 && threadIdx.x == 64, n - n0, 2);
                    ptx::tc_fence_after();
                    KV_DBG(item_it == 0 && threadIdx.x == 64, n - n0, 3);
                    const bool diag = qt * BQ < kt * BK + BK;          // tile touches the diagonal: mask q < key
                    // P^T = exp2(S^T * scale - lse[q]), dS^T = P^T * (dP^T - delta[q]) on packed fp32 pairs: per two elements
                    // one FFMA2, two MUFU, one FADD2, one FMUL2, two F2FP (the scalar version spent ~18 issue slots per
                    // element and made instruction issue, not the tensor pipe, the limiter of this kernel).  -lse and
                    // -delta arrive from shared memory already paired; the diagonal block is a separate instantiation.
                    auto block = [&](auto diag_c) {
                        constexpr bool DIAG = decltype(diag_c)::value;
                        uint32_t sv[32], dv[32], pt[16], dst[16];
                        ptx::tmem_ld32(tmem + st * 128 + lane_addr + c0, sv);
                        ptx::tmem_ld32(tmem + st * 128 + kTmDP + lane_addr + c0, dv);
                        ptx::tmem_ld_wait();
                        KV_DBG(item_it == 0 && threadIdx.x == 64, n - n0, 4);
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const ulonglong2 la = *reinterpret_cast<const ulonglong2*>(lse_s + c0 + u * 8);
                            const ulonglong2 lb = *reinterpret_cast<const ulonglong2*>(lse_s + c0 + u * 8 + 4);
                            const ulonglong2 da = *reinterpret_cast<const ulonglong2*>(del_s + c0 + u * 8);
                            const ulonglong2 db = *reinterpret_cast<const ulonglong2*>(del_s + c0 + u * 8 + 4);
                            const uint64_t nl[4] = {la.x, la.y, lb.x, lb.y}, nd[4] = {da.x, da.y, db.x, db.y};
                            uint32_t pp[4], dd[4];
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const int i = u * 8 + 2 * k;
                                float x0, x1;
                                ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(sv[i]), __uint_as_float(sv[i + 1])), scale2, nl[k]), x0, x1);
                                float e0 = ptx::ex2(x0), e1 = ptx::ex2(x1);
                                if (DIAG) {
                                    if (qt * BQ + c0 + i < kj) e0 = 0.f;
                                    if (qt * BQ + c0 + i + 1 < kj) e1 = 0.f;
                                }
                                const uint64_t t2 = ptx::add_f32x2(ptx::pack_f32x2(__uint_as_float(dv[i]), __uint_as_float(dv[i + 1])), nd[k]);
                                float d0, d1;
                                ptx::unpack_f32x2(ptx::mul_f32x2(ptx::pack_f32x2(e0, e1), t2), d0, d1);
                                pp[k] = ptx::pack_bf16x2(e0, e1);
                                dd[k] = ptx::pack_bf16x2(d0, d1);
                            }
#pragma unroll
                            for (int k = 0; k < 4; ++k) { pt[u * 4 + k] = pp[k]; dst[u * 4 + k] = dd[k]; }
                        }
                        // bf16 pairs over the first half of this thread's own score columns (the other thread of the row
                        // owns the other 32 columns, so neither overwrites scores its partner has not read yet)
                        ptx::tmem_st16(tmem + st * 128 + lane_addr + c0, pt);
                        ptx::tmem_st16(tmem + st * 128 + kTmDP + lane_addr + c0, dst);
                        ptx::tmem_st_wait();
                    };
                    if (diag) block(std::true_type{}); else block(std::false_type{});
                    KV_DBG(item_it == 0 && threadIdx.x == 64, n - n0, 5);
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(&t_full[st]);
                    KV_DBG(item_it == 0 && threadIdx.x == 64, n - n0, 6);
                }
            }
            // epilogue of the item: dV, dK (scaled, inverse-rotated) -> d(qkv)
            ptx::mbar_wait(acc_full, item_it & 1);
            ptx::tc_fence_after();
            const int64_t trow = (int64_t)(b * a.S + kj) * a.ld_out;
#pragma unroll 1
            for (int c = half * (HD / 2); c < (half + 1) * (HD / 2); c += 32) {
                uint32_t v[32], kk[32];
                ptx::tmem_ld32(tmem + kTmDV + lane_addr + c, v);
                ptx::tmem_ld32(tmem + kTmDK + lane_addr + c, kk);
                ptx::tmem_ld_wait();
                if (!whole) {   // a chunk of a longer item: raw fp32 partial sums for attn_dkv_reduce
                    float* pv = a.partial + (((int64_t)it * 2 + 0) * BK + r) * HD + c;
                    float* pk = a.partial + (((int64_t)it * 2 + 1) * BK + r) * HD + c;
#pragma unroll
                    for (int i = 0; i < 32; i += 8) { st8f(pv + i, v + i); st8f(pk + i, kk + i); }   // 32-byte stores
                    continue;
                }
#pragma unroll
                for (int u = 0; u < 4; u += 2) {
                    float fv[8], fk[8], fv2[8], fk2[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        fv[i] = __uint_as_float(v[u * 8 + i]);
                        fk[i] = __uint_as_float(kk[u * 8 + i]) * a.scale;
                        fv2[i] = __uint_as_float(v[u * 8 + 8 + i]);
                        fk2[i] = __uint_as_float(kk[u * 8 + 8 + i]) * a.scale;
                    }
                    unrope16(fk, fk2, a, kj, c + u * 8, HD);
                    st_row32(a.out + trow + (a.H + a.KV + g) * HD + c + u * 8, pack8(fv), pack8(fv2));
                    st_row32(a.out + trow + (a.H + g) * HD + c + u * 8, pack8(fk), pack8(fk2));
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(acc_empty);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem);
}

// Fold the chunk partials of every (batch, kv head, key tile) that was split, in chunk order; scale and un-rotate dK.
// Block = 8 key rows x (hd / 8) column groups, so a warp reads whole 128-byte lines of the partials.
template <int HD>
__global__ void __launch_bounds__(8 * (HD / 8))
attn_dkv_reduce_kernel(const AttnArgs a) {
    const int k_tiles = a.S / 128, q_tiles64 = a.S / 64, rep = a.H / a.KV;
    const int rg = blockIdx.x % 16, item = blockIdx.x / 16;            // 16 groups of 8 rows per 128-key tile
    const int kt = item % k_tiles, bg = item / k_tiles;
    const int b = bg / a.KV, g = bg % a.KV;
    const int r = rg * 8 + threadIdx.x / (HD / 8), d = (threadIdx.x % (HD / 8)) * 8;
    int prefix = 0;
    for (int k = 0; k < kt; ++k) prefix += (rep * (q_tiles64 - 2 * k) + a.chunk_len - 1) / a.chunk_len;
    const int nc = (rep * (q_tiles64 - 2 * kt) + a.chunk_len - 1) / a.chunk_len;
    if (nc == 1) return;                       // written directly by the main kernel
    const int kj = kt * 128 + r;
    const int64_t trow = (int64_t)(b * a.S + kj) * a.ld_out;
    float fv[8] = {0, 0, 0, 0, 0, 0, 0, 0}, fk[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int c = 0; c < nc; ++c) {
        const int64_t u = (int64_t)(prefix + c) * (a.B * a.KV) + bg;
        const float4* pv = reinterpret_cast<const float4*>(a.partial + ((u * 2 + 0) * 128 + r) * HD + d);
        const float4* pk = reinterpret_cast<const float4*>(a.partial + ((u * 2 + 1) * 128 + r) * HD + d);
        const float4 v0 = pv[0], v1 = pv[1], k0 = pk[0], k1 = pk[1];
        fv[0] += v0.x; fv[1] += v0.y; fv[2] += v0.z; fv[3] += v0.w; fv[4] += v1.x; fv[5] += v1.y; fv[6] += v1.z; fv[7] += v1.w;
        fk[0] += k0.x; fk[1] += k0.y; fk[2] += k0.z; fk[3] += k0.w; fk[4] += k1.x; fk[5] += k1.y; fk[6] += k1.z; fk[7] += k1.w;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) fk[i] *= a.scale;
    unrope8(fk, a, kj, d, HD);
    st8(a.out + trow + (a.H + a.KV + g) * HD + d, pack8(fv));
    st8(a.out + trow + (a.H + g) * HD + d, pack8(fk));
}

// ================================================================================================= backward: dQ
template <int HD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV64, const AttnArgs a) {
    // Every A operand of this kernel lives in TMEM: Q and dO of the query tile (bf16, written once per item by the math
    // threads straight from global memory) and dS (written over the first half of each thread's dP columns).  Shared
    // memory then only carries the K / V ring, i.e. 48 KB of operand reads per key tile instead of 144 KB: the previous
    // version, with Q, dO and dS in shared memory, was bound by shared-memory bandwidth (tools/umma_bench.cu: an SS
    // 128x64x16 MMA takes 48 cycles, its TS form 32; and every st.shared / ld.shared of the math warps queued behind
    // the tensor core's operand fetches).
    constexpr int BM = 128, BN = 64;                   // query tile (UMMA M), key tile (UMMA N)
    constexpr int kKBytes = BN * HD * 2;
    constexpr int kTmDP = 64, kTmDQ = 256;            // S / dP of buffer b at b*128 and b*128 + 64
    constexpr int kTmQ = 384, kTmDO = 384 + HD / 2;   // bf16 A operands: HD/2 columns each
    constexpr int NK = 6;      // K / V stages: loads run ahead of the MMAs
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sK = smem;                                // NK stages [BN x HD]
    uint8_t* sV = sK + NK * kKBytes;                   // NK stages
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + NK * kKBytes);
    uint64_t* qd_full = bars;        // 1: Q, dO of the item are in TMEM (8 math warps)
    uint64_t* kv_full = bars + 2;    // NK (<= 8)
    uint64_t* kv_empty = bars + 10;  // NK
    uint64_t* s_full = bars + 18;    // 2
    uint64_t* t_full = bars + 20;    // 2
    uint64_t* acc_full = bars + 22;  // 1
    uint64_t* acc_empty = bars + 23; // 1
    uint64_t* item_full = bars + 24; // 2: the scheduler's next item is in sItem[slot]
    uint64_t* item_empty = bars + 26;// 2: MMA warp + 8 math warps have read it
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 28);
    int* sItem = reinterpret_cast<int*>(bars + 29);    // [2]
    float* sDelta = reinterpret_cast<float*>(bars + 32);   // [2][2][BM] halves of the row dot products dO . O

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_tiles = a.S / BM, rep = a.H / a.KV;
    const int items = a.B * a.H * q_tiles;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV64);
        ptx::mbar_init(qd_full, 8);
        for (int i = 0; i < NK; ++i) { ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&t_full[i], 8);
            ptx::mbar_init(&item_full[i], 1); ptx::mbar_init(&item_empty[i], 9);
        }
        ptx::mbar_init(acc_full, 1); ptx::mbar_init(acc_empty, 8);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    auto decode = [&](int it, int& b, int& h, int& qt) {
        qt = q_tiles - 1 - it / (a.B * a.H);
        const int r = it % (a.B * a.H);
        b = r / a.H; h = r % a.H;
    };

    if (warp == 0) {
        if (ptx::elect_one()) {
            // items come from an atomic counter, heaviest (latest query tiles) first: a static round-robin gave the
            // first CTAs 76 key tiles and the last ones 44 (ncu: SMs idle for 23% of the kernel)
            int kv_it = 0;
            for (int item_it = 0;; ++item_it) {
                const int it = atomicAdd(&g_bwd_sched[2], 1);
                const int sl = item_it & 1;
                ptx::mbar_wait(&item_empty[sl], ((item_it >> 1) & 1) ^ 1);
                sItem[sl] = it < items ? it : -1;
                ptx::mbar_arrive(&item_full[sl]);
                if (it >= items) break;
                int b, h, qt; decode(it, b, h, qt);
                const int g = h / rep;
                const int nkv = (qt + 1) * (BM / BN);
                for (int j = 0; j < nkv; ++j, ++kv_it) {
                    const int st = kv_it % NK;
                    ptx::mbar_wait(&kv_empty[st], ((kv_it / NK) & 1) ^ 1);
                    ptx::mbar_arrive_expect_tx(&kv_full[st], 2 * kKBytes);
                    const int krow = b * a.S + j * BN;
#pragma unroll
                    for (int c = 0; c < HD / 64; ++c) {
                        ptx::tma_load_2d(sK + st * kKBytes + c * BN * 128, &tmQKV64, &kv_full[st], (a.H + g) * HD + c * 64, krow);
                        ptx::tma_load_2d(sV + st * kKBytes + c * BN * 128, &tmQKV64, &kv_full[st], (a.H + a.KV + g) * HD + c * 64, krow);
                    }
                }
            }
            sched_done(&g_bwd_sched[2]);
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(BM, BN, false, false);
        constexpr uint32_t idesc_q = ptx::umma_idesc_bf16(BM, HD, false, true);    // dQ = dS K, B = K MN-major
        int item_it = 0, n_it = 0;                                      // n_it: global key-tile counter
        auto issue_sdp = [&](int t) {                                   // S, dP of tile t -> TMEM buffer t & 1
            const int st = t & 1, ks = t % NK;
            ptx::mbar_wait(&kv_full[ks], (t / NK) & 1);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
                const uint64_t dk = desc_kmajor0(ptx::smem_u32(sK + ks * kKBytes)), dv = desc_kmajor0(ptx::smem_u32(sV + ks * kKBytes));
#pragma unroll
                for (int k = 0; k < HD / 16; ++k)
                    ptx::umma_bf16_ts(tmem + st * 128, tmem + kTmQ + k * 8, desc_kmajor_step(dk, BN, k), idesc_s, k != 0 ? 1u : 0u);
#pragma unroll
                for (int k = 0; k < HD / 16; ++k)
                    ptx::umma_bf16_ts(tmem + st * 128 + kTmDP, tmem + kTmDO + k * 8, desc_kmajor_step(dv, BN, k), idesc_s, k != 0 ? 1u : 0u);
                ptx::umma_commit(&s_full[st]);
            }
            __syncwarp();
        };
        for (;; ++item_it) {
            const int sl = item_it & 1;
            ptx::mbar_wait(&item_full[sl], (item_it >> 1) & 1);
            const int it = sItem[sl];
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&item_empty[sl]);
            if (it < 0) break;
            int b, h, qt; decode(it, b, h, qt);
            ptx::mbar_wait(qd_full, item_it & 1);
            ptx::tc_fence_after();
            const int nkv = (qt + 1) * (BM / BN);
            issue_sdp(n_it);
            ptx::mbar_wait(acc_empty, (item_it & 1) ^ 1);               // the epilogue of the previous item has read dQ
            for (int j = 0; j < nkv; ++j, ++n_it) {
                if (j + 1 < nkv) issue_sdp(n_it + 1);
                const int st = n_it & 1;
                ptx::mbar_wait(&t_full[st], (n_it >> 1) & 1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const int ks = n_it % NK;
                    // A = dS from TMEM: keys 0-31 of the tile sit in columns [0,16) of the dP buffer, keys 32-63 in [32,48)
                    const uint64_t dkm = desc_mnmajor0(ptx::smem_u32(sK + ks * kKBytes), BN);
#pragma unroll
                    for (int k = 0; k < BN / 16; ++k)
                        ptx::umma_bf16_ts(tmem + kTmDQ, tmem + st * 128 + kTmDP + (k >> 1) * 32 + (k & 1) * 8, desc_mnmajor_step(dkm, k), idesc_q,
                                          (j != 0 || k != 0) ? 1u : 0u);
                    ptx::umma_commit(&kv_empty[ks]);
                    if (j == nkv - 1) ptx::umma_commit(acc_full);
                }
                __syncwarp();
            }
        }
    } else {
        const int q4 = warp & 3, half = (warp - 2) >> 2;
        const int r = q4 * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
        const int c0 = half * 32;
        const uint64_t scale2 = ptx::pack_f32x2(a.scale_log2, a.scale_log2);
        // This thread's half of the Q and dO rows of an item: global -> registers (load_rows, issued BEFORE the wait for
        // the previous item's last MMAs so that the load latency hides behind it) -> TMEM as bf16 pairs, HD/4 columns each
        // (publish_rows, once the previous item no longer reads these columns).
        uint32_t vq[HD / 4], vd[HD / 4];
        float dot_half = 0.f, delta_staged = 0.f;                      // half of / whole delta of row r of the staged item
        int64_t delta_idx = 0;
        int stage_n = 0;
        auto load_rows = [&](int it2) {
            int b2, h2, qt2; decode(it2, b2, h2, qt2);
            const int64_t tok = (int64_t)b2 * a.S + qt2 * BM + r;
            const uint4* qrow = reinterpret_cast<const uint4*>(a.qkv_in + tok * a.ld_qkv + h2 * HD + half * (HD / 2));
            const uint4* drow = reinterpret_cast<const uint4*>(a.dout_in + tok * a.ld_do + h2 * HD + half * (HD / 2));
            const uint4* orow = reinterpret_cast<const uint4*>(a.o_in + tok * a.ld_do + h2 * HD + half * (HD / 2));
            delta_idx = ((int64_t)b2 * a.H + h2) * a.S + qt2 * BM + r;
            // one row per thread: every load instruction of a warp touches 32 different lines, so the count of
            // instructions is what this costs (~5 900 cycles per item with 16-byte loads): 32-byte loads halve it
            uint32_t vo[HD / 4];
#pragma unroll
            for (int i = 0; i < HD / 32; ++i) {
                ld_row32(qrow + 2 * i, vq + 8 * i);
                ld_row32(drow + 2 * i, vd + 8 * i);
                ld_row32(orow + 2 * i, vo + 8 * i);
            }
            float dot = 0.f;                                           // this half of delta = sum_d dO * O of the row
#pragma unroll
            for (int i = 0; i < HD / 16; ++i) {
                float fo[8], fd[8];
                unpack8(*reinterpret_cast<const bf16x8*>(&vo[4 * i]), fo);
                unpack8(*reinterpret_cast<const bf16x8*>(&vd[4 * i]), fd);
#pragma unroll
                for (int k = 0; k < 8; ++k) dot = fmaf(fo[k], fd[k], dot);
            }
            dot_half = dot;
        };
        auto publish_rows = [&]() {
            if constexpr (HD == 128) {
                ptx::tmem_st32(tmem + kTmQ + lane_addr + half * (HD / 4), vq);
                ptx::tmem_st32(tmem + kTmDO + lane_addr + half * (HD / 4), vd);
            } else {
                ptx::tmem_st16(tmem + kTmQ + lane_addr + half * (HD / 4), vq);
                ptx::tmem_st16(tmem + kTmDO + lane_addr + half * (HD / 4), vd);
            }
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(qd_full);
            // the two threads of a row add their halves in a fixed order; the row's delta also goes to global memory for
            // the dK/dV kernel, which runs after this one (this replaced a separate kernel that re-read O and dO)
            float* xd = sDelta + (stage_n & 1) * 2 * BM;
            xd[half * BM + r] = dot_half;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            delta_staged = xd[r] + xd[BM + r];
            if (half == 0) a.delta_out[delta_idx] = delta_staged;
            ++stage_n;
        };
        int item_it = 0, n_it = 0;
        auto next_item = [&](int idx) {                                // idx-th item of this CTA (-1: queue drained)
            const int sl = idx & 1;
            ptx::mbar_wait(&item_full[sl], (idx >> 1) & 1);
            const int v = sItem[sl];
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&item_empty[sl]);
            return v;
        };
        int it = next_item(0);
        if (it >= 0) { load_rows(it); publish_rows(); }
        for (; it >= 0; ++item_it) {
            int b, h, qt; decode(it, b, h, qt);
            DQ_DBG(threadIdx.x == 64, item_it, 0);
            const int qi = qt * BM + r;
            const int64_t so = ((int64_t)b * a.H + h) * a.S + qi;
            const float lse = a.lse[so], delta = delta_staged;         // staged before the next item's staging overwrites it
            const uint64_t nlse2 = ptx::pack_f32x2(-lse, -lse), ndel2 = ptx::pack_f32x2(-delta, -delta);
            const int nkv = (qt + 1) * (BM / BN);
            for (int j = 0; j < nkv; ++j, ++n_it) {
                const int st = n_it & 1;
                ptx::mbar_wait(&s_full[st], (n_it >> 1) & 1);
                ptx::tc_fence_after();
                if (j == 0) DQ_DBG(threadIdx.x == 64, item_it, 1);
                const bool diag = j * BN + BN > qt * BM;
                // dS = exp2(S * scale - lse) * (dP - delta) on packed fp32 pairs (see attn_bwd_dkv_kernel)
                auto block = [&](auto diag_c) {
                    constexpr bool DIAG = decltype(diag_c)::value;
                    uint32_t sv[32], dv[32], dst[16];
                    ptx::tmem_ld32(tmem + st * 128 + lane_addr + c0, sv);
                    ptx::tmem_ld32(tmem + st * 128 + kTmDP + lane_addr + c0, dv);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        float x0, x1;
                        ptx::unpack_f32x2(ptx::fma_f32x2(ptx::pack_f32x2(__uint_as_float(sv[i]), __uint_as_float(sv[i + 1])), scale2, nlse2), x0, x1);
                        float e0 = ptx::ex2(x0), e1 = ptx::ex2(x1);
                        if (DIAG) {
                            if (j * BN + c0 + i > qi) e0 = 0.f;
                            if (j * BN + c0 + i + 1 > qi) e1 = 0.f;
                        }
                        const uint64_t t2 = ptx::add_f32x2(ptx::pack_f32x2(__uint_as_float(dv[i]), __uint_as_float(dv[i + 1])), ndel2);
                        float d0, d1;
                        ptx::unpack_f32x2(ptx::mul_f32x2(ptx::pack_f32x2(e0, e1), t2), d0, d1);
                        dst[i >> 1] = ptx::pack_bf16x2(d0, d1);
                    }
                    // bf16 pairs over the first half of this thread's own dP columns (its partner owns the other 32)
                    ptx::tmem_st16(tmem + st * 128 + kTmDP + lane_addr + c0, dst);
                    ptx::tmem_st_wait();
                };
                if (diag) block(std::true_type{}); else block(std::false_type{});
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&t_full[st]);
            }
            DQ_DBG(threadIdx.x == 64, item_it, 2);
            const int it_next = next_item(item_it + 1);
            if (it_next >= 0) load_rows(it_next);                        // in flight while the last MMAs of this item drain
            DQ_DBG(threadIdx.x == 64, item_it, 3);
            ptx::mbar_wait(acc_full, item_it & 1);                       // every MMA of the item has completed
            ptx::tc_fence_after();
            DQ_DBG(threadIdx.x == 64, item_it, 4);
            if (it_next >= 0) publish_rows();
            DQ_DBG(threadIdx.x == 64, item_it, 5);                            // next Q, dO: S, dP of its first tile start under the epilogue
            __nv_bfloat16* orow = a.out + (int64_t)(b * a.S + qi) * a.ld_out + h * HD;
#pragma unroll 1
            for (int c = half * (HD / 2); c < (half + 1) * (HD / 2); c += 32) {
                uint32_t v[32];
                ptx::tmem_ld32(tmem + kTmDQ + lane_addr + c, v);
                ptx::tmem_ld_wait();
                DQ_DBG(threadIdx.x == 64, item_it, c == 0 ? 8 : 10);
#pragma unroll
                for (int u = 0; u < 4; u += 2) {
                    float f[8], f2[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        f[i] = __uint_as_float(v[u * 8 + i]) * a.scale;
                        f2[i] = __uint_as_float(v[u * 8 + 8 + i]) * a.scale;
                    }
                    unrope16(f, f2, a, qi, c + u * 8, HD);
                    st_row32(orow + c + u * 8, pack8(f), pack8(f2));        // full 32-byte sectors
                }
                DQ_DBG(threadIdx.x == 64, item_it, c == 0 ? 9 : 11);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(acc_empty);
            DQ_DBG(threadIdx.x == 64, item_it, 6);
#if defined(FTL_ATTN_DEBUG) && defined(FTL_PROBE_DQ)
            if (blockIdx.x == 0 && threadIdx.x == 64 && item_it < 64) g_attn_dbg[item_it * 16 + 7] = nkv;
#endif
            it = it_next;
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------------------------------------------- host
// bf16 [rows, cols] (ld elements), box {64 columns, box_rows}; encoded once per key and cached
static int attn_tmap(CUtensorMap* out, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    return g_tmaps.get(out, ptr, rows, cols, ld, 64, box_rows);
}

template <typename K>
static int set_smem(K kern, int bytes) {
    return (int)cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

}  // namespace ftl

using namespace ftl;

// qkv: [B*S, ld_qkv] fused projection (q heads | k heads | v heads), RoPE already applied to q and k.
// out: [B*S, H*hd]; lse: fp32 [B, H, S].  S must be a multiple of 128; hd in {64, 128}.
#ifdef FTL_ATTN_DEBUG
FTL_API int ftl_attn_dbg_read(long long* host) {
    return (int)cudaMemcpyFromSymbol(host, g_attn_dbg, sizeof(g_attn_dbg));
}
#endif

FTL_API int ftl_attn_fwd(const void* qkv, int64_t ld_qkv, void* out, int64_t ld_out, float* lse, int B, int S, int H, int KV,
                         int hd, cudaStream_t stream) {
    // rows are written with 32-byte stores: strides in multiples of 16 elements, 32-byte aligned bases
    if (S % 128 || (hd != 64 && hd != 128) || H % KV || ld_qkv % 16 || ld_out % 16 || ((uintptr_t)out & 31))
        return (int)cudaErrorInvalidValue;
    CUtensorMap tm;
    int e = attn_tmap(&tm, qkv, (int64_t)B * S, (int64_t)(H + 2 * KV) * hd, ld_qkv, 128);
    if (e) return e;
    AttnArgs a{};
    a.out = (__nv_bfloat16*)out; a.ld_out = ld_out; a.lse = lse;
    a.B = B; a.S = S; a.H = H; a.KV = KV; a.ld_qkv = (int)ld_qkv;
    a.scale = 1.f / sqrtf((float)hd); a.scale_log2 = a.scale * kLog2e;
    static const bool use_v1 = [] { const char* v = getenv("FTL_ATTN_FWD"); return v && v[0] == '1'; }();
    if (S % 256 == 0 && !use_v1) {      // two query tiles per CTA, dynamic scheduler (see attn_fwd2_kernel)
        const int items2 = B * H * (S / 256);
        const int grid2 = items2 < kNumSMs ? items2 : kNumSMs;
        if (hd == 128) {
            constexpr int smem = 128 * 128 * 2 * 7 + 26 * 8 + 2048 + 16;
            static bool once = false; if (!once) { if ((e = set_smem(attn_fwd2_kernel<128>, smem))) return e; once = true; }
            attn_fwd2_kernel<128><<<grid2, kFwd2Threads, smem, stream>>>(tm, a);
        } else {
            constexpr int smem = 128 * 64 * 2 * 7 + 26 * 8 + 2048 + 16;
            static bool once = false; if (!once) { if ((e = set_smem(attn_fwd2_kernel<64>, smem))) return e; once = true; }
            attn_fwd2_kernel<64><<<grid2, kFwd2Threads, smem, stream>>>(tm, a);
        }
        return (int)cudaGetLastError();
    }
    const int items = B * H * (S / 128);
    const int grid = items < kNumSMs ? items : kNumSMs;
    if (hd == 128) {
        constexpr int smem = 128 * 128 * 2 * 7 + 160 + 2048 + 512;   // Q + 3 x (K, V) = 224 KB, barriers, exchange (base is 1024-aligned)
        static bool once = false; if (!once) { if ((e = set_smem(attn_fwd_kernel<128>, smem))) return e; once = true; }
        attn_fwd_kernel<128><<<grid, kAttnThreads, smem, stream>>>(tm, a);
    } else {
        constexpr int smem = 128 * 64 * 2 * 7 + 160 + 2048 + 512;
        static bool once = false; if (!once) { if ((e = set_smem(attn_fwd_kernel<64>, smem))) return e; once = true; }
        attn_fwd_kernel<64><<<grid, kAttnThreads, smem, stream>>>(tm, a);
    }
    return (int)cudaGetLastError();
}

static void dkv_plan(int B, int S, int H, int KV, int* chunk_len, int* units) {
    const int rep = H / KV, q64 = S / 64, kt_n = S / 128;
    int64_t total = 0;
    for (int k = 0; k < kt_n; ++k) total += (int64_t)rep * (q64 - 2 * k);
    total *= (int64_t)B * KV;
    // Units per SM: every unit pays a fill/drain of the MMA pipeline and, when an item is split, 128 KB of fp32
    // partials; units are handed out dynamically (longest first), so a few long units per SM already balance.
    // Measured (profiles/attn_tc_notes.md): 2 per SM is best at the model shape (59 inner iterations per SM),
    // 3 at seq 8192 (900 per SM).
    static const int forced = [] { const char* v = getenv("FTL_DKV_UNITS_PER_SM"); return v ? atoi(v) : 0; }();
    int per_sm = total / kNumSMs >= 256 ? 3 : 2;
    if (forced > 0) per_sm = forced;
    int lc = (int)((total + per_sm * kNumSMs - 1) / (per_sm * kNumSMs));
    if (lc < 8) lc = 8;
    int u = 0;
    for (int k = 0; k < kt_n; ++k) u += (rep * (q64 - 2 * k) + lc - 1) / lc;
    *chunk_len = lc;
    *units = u * B * KV;
}

// bytes of fp32 scratch ftl_attn_bwd needs (partials of the split dK/dV items)
FTL_API int64_t ftl_attn_bwd_scratch_bytes(int B, int S, int H, int KV, int hd) {
    int lc, units;
    dkv_plan(B, S, H, KV, &lc, &units);
    return (int64_t)units * 2 * 128 * hd * 4;
}

// d_out: [B*S, H*hd] gradient of the attention output; out: the forward output; writes dq | dk | dv into dqkv
// ([B*S, ld_qkv]) with the inverse RoPE applied to dq and dk when cos/sin are given.  delta: fp32 scratch [B, H, S].
FTL_API int ftl_attn_bwd(const void* qkv, int64_t ld_qkv, const void* out, const void* d_out, int64_t ld_out, const float* lse,
                         float* delta, void* dqkv, const float* cos_t, const float* sin_t, float* scratch, int B, int S, int H,
                         int KV, int hd, cudaStream_t stream) {
    if (S % 128 || (hd != 64 && hd != 128) || H % KV || ld_qkv % 16 || ld_out % 16 || ((uintptr_t)dqkv & 31) ||
        ((uintptr_t)scratch & 31) || ((uintptr_t)qkv & 31) || ((uintptr_t)out & 31) || ((uintptr_t)d_out & 31) ||
        ((uintptr_t)cos_t & 31) || ((uintptr_t)sin_t & 31))
        return (int)cudaErrorInvalidValue;
    const int64_t T = (int64_t)B * S;
    CUtensorMap tq128, tq64, tdo128, tdo64;
    int e;
    if ((e = attn_tmap(&tq128, qkv, T, (int64_t)(H + 2 * KV) * hd, ld_qkv, 128))) return e;
    if ((e = attn_tmap(&tq64, qkv, T, (int64_t)(H + 2 * KV) * hd, ld_qkv, 64))) return e;
    if ((e = attn_tmap(&tdo128, d_out, T, (int64_t)H * hd, ld_out, 128))) return e;
    if ((e = attn_tmap(&tdo64, d_out, T, (int64_t)H * hd, ld_out, 64))) return e;
    AttnArgs a{};
    a.cos_t = cos_t; a.sin_t = sin_t;
    a.out = (__nv_bfloat16*)dqkv; a.ld_out = ld_qkv; a.lse = const_cast<float*>(lse); a.delta = delta;
    a.B = B; a.S = S; a.H = H; a.KV = KV; a.ld_qkv = (int)ld_qkv;
    a.scale = 1.f / sqrtf((float)hd); a.scale_log2 = a.scale * kLog2e;
    dkv_plan(B, S, H, KV, &a.chunk_len, &a.units);
    a.partial = scratch;
    a.qkv_in = (const __nv_bfloat16*)qkv; a.dout_in = (const __nv_bfloat16*)d_out; a.ld_do = ld_out;
    a.o_in = (const __nv_bfloat16*)out; a.delta_out = delta;
    const int items_kv = a.units, items_q = B * H * (S / 128), items_red = B * KV * (S / 128);
    if (hd == 128) {
        constexpr int smem_kv = 2 * 128 * 128 * 2 + 8 * 64 * 128 * 2 + 4 * 64 * 4 + 1024 + 256;   // K, V, 4 x (Q, dO)
        constexpr int smem_q = 12 * 64 * 128 * 2 + 256 + 2048 + 256;               // K / V ring, barriers, delta exchange
        static bool once = false;
        if (!once) {
            if ((e = set_smem(attn_bwd_dkv_kernel<128>, smem_kv))) return e;
            if ((e = set_smem(attn_bwd_dq_kernel<128>, smem_q))) return e;
            once = true;
        }
        // dQ first: it also produces delta = rowsum(dO * O), which the dK/dV kernel reads
        attn_bwd_dq_kernel<128><<<items_q < kNumSMs ? items_q : kNumSMs, kAttnThreads, smem_q, stream>>>(tq64, a);
        FTL_LAUNCH_CHECK();
        attn_bwd_dkv_kernel<128><<<items_kv < kNumSMs ? items_kv : kNumSMs, kAttnThreads, smem_kv, stream>>>(tq128, tq64, tdo64, a);
        FTL_LAUNCH_CHECK();
        attn_dkv_reduce_kernel<128><<<items_red * 16, 128, 0, stream>>>(a);
    } else {
        constexpr int smem_kv = 2 * 128 * 64 * 2 + 8 * 64 * 64 * 2 + 4 * 64 * 4 + 1024 + 256;
        constexpr int smem_q = 12 * 64 * 64 * 2 + 256 + 2048 + 256;
        static bool once = false;
        if (!once) {
            if ((e = set_smem(attn_bwd_dkv_kernel<64>, smem_kv))) return e;
            if ((e = set_smem(attn_bwd_dq_kernel<64>, smem_q))) return e;
            once = true;
        }
        // dQ first: it also produces delta = rowsum(dO * O), which the dK/dV kernel reads
        attn_bwd_dq_kernel<64><<<items_q < kNumSMs ? items_q : kNumSMs, kAttnThreads, smem_q, stream>>>(tq64, a);
        FTL_LAUNCH_CHECK();
        attn_bwd_dkv_kernel<64><<<items_kv < kNumSMs ? items_kv : kNumSMs, kAttnThreads, smem_kv, stream>>>(tq128, tq64, tdo64, a);
        FTL_LAUNCH_CHECK();
        attn_dkv_reduce_kernel<64><<<items_red * 16, 64, 0, stream>>>(a);
    }
    return (int)cudaGetLastError();
}
