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
//   attn_fwd_kernel   CTA = (batch, q head, 128-query tile); per 128-key tile: S = Q K^T (TMEM) -> online softmax in
//                     registers (thread = query row) -> P (bf16) to swizzled smem -> O_tile = P V (TMEM) -> rescale
//                     and add in registers.  Writes O (bf16) and the row log-sum-exp (fp32, base 2 domain).
//   attn_bwd_dkv      CTA = (batch, kv head, 128-key tile); loops over the group's query heads and 64-query tiles:
//                     S^T = K Q^T, dP^T = V dO^T (TMEM) -> P^T, dS^T (thread = key row) to smem ->
//                     dV += P^T dO, dK += dS^T Q accumulated in TMEM over the whole loop.
//   attn_bwd_dq       CTA = (batch, q head, 128-query tile); loops over 64-key tiles: S = Q K^T, dP = dO V^T ->
//                     dS (thread = query row) to smem -> dQ += dS K accumulated in TMEM.
//   attn_delta        delta[t, h] = sum_d dO * O   (fp32), the softmax-backward row term.
// Warp roles as in the GEMM: warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2-5 = 128 math threads.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cudaTypedefs.h>
#include <stdio.h>

#include <mutex>

#include "common/ptx.cuh"
#include "common/util.cuh"

namespace ftl {

constexpr int kAttnThreads = 192;
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

// K-major operand tile [rows x kdim] (kdim contiguous): descriptor for the k16-th UMMA step.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t base, int rows, int k16) {
    return ptx::umma_smem_desc(base + (k16 >> 2) * rows * 128 + (k16 & 3) * 32, 0, 1024);
}
// MN-major operand: tile stored [krows x mn] (mn contiguous, chunked by 64): step k16 covers rows 16*k16..+15.
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int krows, int k16) {
    return ptx::umma_smem_desc(base + k16 * 16 * 128, krows * 128, 1024);
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
};

// ================================================================================================= forward
template <int HD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnArgs a) {
    constexpr int BM = 128, BN = 128;
    constexpr int kQBytes = BM * HD * 2, kKBytes = BN * HD * 2, kPBytes = BM * BN * 2;
    constexpr int kTmemO = 2 * BN;                    // columns: S0 [0,128) S1 [128,256) O [256, 256+HD)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kQBytes;                       // 2 stages
    uint8_t* sV = sK + 2 * kKBytes;                   // 2 stages
    uint8_t* sP = sV + 2 * kKBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kPBytes);
    uint64_t* q_full = bars;          // 1
    uint64_t* q_empty = bars + 1;     // 1
    uint64_t* kv_full = bars + 2;     // 2
    uint64_t* kv_empty = bars + 4;    // 2
    uint64_t* s_full = bars + 6;      // 2
    uint64_t* p_full = bars + 8;      // 1
    uint64_t* o_full = bars + 9;      // 1
    uint64_t* o_empty = bars + 10;    // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_tiles = a.S / BM;
    const int items = a.B * a.H * q_tiles;
    const int rep = a.H / a.KV;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::mbar_init(q_full, 1); ptx::mbar_init(q_empty, 1);
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); ptx::mbar_init(&s_full[i], 1); }
        ptx::mbar_init(p_full, 4); ptx::mbar_init(o_full, 1); ptx::mbar_init(o_empty, 4);
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
                    const int st = kv_it & 1;
                    ptx::mbar_wait(&kv_empty[st], ((kv_it >> 1) & 1) ^ 1);
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
        int kv_it = 0, q_it = 0, s_it = 0, p_it = 0, o_it = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x, ++q_it) {
            int b, h, qt; decode(it, b, h, qt);
            ptx::mbar_wait(q_full, q_it & 1);
            const int nkv = qt + 1;
            auto issue_s = [&](int kv_index, int s_index) {
                const int st = kv_index & 1;
                ptx::mbar_wait(&kv_full[st], (kv_index >> 1) & 1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t qa = ptx::smem_u32(sQ), kb = ptx::smem_u32(sK + st * kKBytes);
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        ptx::umma_bf16(tmem + (s_index & 1) * BN, desc_kmajor(qa, BM, k), desc_kmajor(kb, BN, k), idesc_s, k != 0);
                    ptx::umma_commit(&s_full[s_index & 1]);
                }
                __syncwarp();
            };
            issue_s(kv_it, s_it);
            for (int j = 0; j < nkv; ++j) {
                if (j + 1 < nkv) issue_s(kv_it + 1, s_it + 1);       // next scores while the softmax of this tile runs
                ptx::mbar_wait(p_full, p_it & 1); ++p_it;             // P(j) is in smem, S(j) has been read
                ptx::mbar_wait(o_empty, (o_it & 1) ^ 1);              // O_tile(j-1) has been read
                ptx::tc_fence_after();
                const int st = kv_it & 1;
                if (ptx::elect_one()) {
                    const uint32_t pa = ptx::smem_u32(sP), vb = ptx::smem_u32(sV + st * kKBytes);
#pragma unroll
                    for (int k = 0; k < BN / 16; ++k)
                        ptx::umma_bf16(tmem + kTmemO, desc_kmajor(pa, BM, k), desc_mnmajor(vb, BN, k), idesc_o, k != 0);
                    ptx::umma_commit(o_full);
                    ptx::umma_commit(&kv_empty[st]);
                    if (j == nkv - 1) ptx::umma_commit(q_empty);
                }
                __syncwarp();
                ++o_it; ++kv_it; ++s_it;
            }
        }
    } else {
        const int q = warp & 3;
        const int r = q * 32 + lane;                                   // query row inside the tile = TMEM lane
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        int s_it = 0, o_it = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x) {
            int b, h, qt; decode(it, b, h, qt);
            const int qi = qt * BM + r;
            float m = -INFINITY, l = 0.f;
            float acc[HD];
#pragma unroll
            for (int d = 0; d < HD; ++d) acc[d] = 0.f;
            for (int j = 0; j <= qt; ++j, ++s_it, ++o_it) {
                ptx::mbar_wait(&s_full[s_it & 1], (s_it >> 1) & 1);
                ptx::tc_fence_after();
                const uint32_t ts = tmem + (s_it & 1) * BN + lane_addr;
                const bool diag = j == qt;
                // pass 1: row maximum
                float mx = -INFINITY;
#pragma unroll 1
                for (int c = 0; c < BN; c += 32) {
                    uint32_t v[32];
                    ptx::tmem_ld32(ts + c, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float s = __uint_as_float(v[i]);
                        if (!diag || j * BN + c + i <= qi) mx = fmaxf(mx, s);
                    }
                }
                const float m_new = fmaxf(m, mx * a.scale_log2);
                const float alpha = exp2f(m - m_new);                   // m = -inf on the first tile -> 0
                // pass 2: probabilities -> bf16 -> swizzled smem (A operand of P V), row sum
                float sum = 0.f;
                const uint32_t pbase = ptx::smem_u32(sP);
#pragma unroll 1
                for (int c = 0; c < BN; c += 32) {
                    uint32_t v[32];
                    ptx::tmem_ld32(ts + c, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        float p[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int col = c + u * 8 + i;
                            float e = exp2f(__uint_as_float(v[u * 8 + i]) * a.scale_log2 - m_new);
                            if (diag && j * BN + col > qi) e = 0.f;
                            p[i] = round_bf16(e);
                            sum += p[i];
                        }
                        st_shared16(sw128_addr(pbase, BM, r, c + u * 8), pack8(p));
                    }
                }
                l = l * alpha + sum;
                m = m_new;
                ptx::tc_fence_before();
                ptx::fence_proxy_async();                               // generic-proxy smem writes -> UMMA (async proxy)
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(p_full);
                // O += P V
                ptx::mbar_wait(o_full, o_it & 1);
                ptx::tc_fence_after();
#pragma unroll
                for (int c = 0; c < HD; c += 32) {
                    uint32_t v[32];
                    ptx::tmem_ld32(tmem + kTmemO + lane_addr + c, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) acc[c + i] = acc[c + i] * alpha + __uint_as_float(v[i]);
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(o_empty);
            }
            // finalize: O / l, lse
            const float inv = 1.f / l;
            __nv_bfloat16* orow = a.out + (int64_t)(b * a.S + qi) * a.ld_out + h * HD;
#pragma unroll
            for (int d = 0; d < HD; d += 8) {
                float o[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = acc[d + i] * inv;
                st8(orow + d, pack8(o));
            }
            a.lse[((int64_t)b * a.H + h) * a.S + qi] = m + log2f(l);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem);
}

// ================================================================================================= delta
// delta[b, h, s] = sum_d dO[t, h, d] * O[t, h, d]; one warp per (token, head).
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                  float* __restrict__ delta, int B, int S, int H, int HD, int64_t ld) {
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= (int64_t)B * S * H) return;
    const int64_t t = w / H;
    const int h = (int)(w % H);
    float s = 0.f;
    for (int i = lane * 8; i < HD; i += 256) {
        float x[8], y[8];
        unpack8(ld8(o + t * ld + h * HD + i), x);
        unpack8(ld8(d_o + t * ld + h * HD + i), y);
#pragma unroll
        for (int k = 0; k < 8; ++k) s += x[k] * y[k];
    }
    s = warp_sum(s);
    if (lane == 0) delta[((int64_t)(t / S) * H + h) * S + (t % S)] = s;
}

// inverse rotary embedding of 8 consecutive gradient values (4 pairs) of a q / k row at position pos
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

// ================================================================================================= backward: dK, dV
template <int HD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmQKV64,
                    const __grid_constant__ CUtensorMap tmDO64, const AttnArgs a) {
    constexpr int BK = 128, BQ = 64;                   // key tile (UMMA M), query tile (UMMA N)
    constexpr int kKBytes = BK * HD * 2, kQBytes = BQ * HD * 2, kTBytes = BK * BQ * 2;
    constexpr int kTmS = 0, kTmDP = 64, kTmDV = 128, kTmDK = 128 + HD;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sK = smem;
    uint8_t* sV = sK + kKBytes;
    uint8_t* sQ = sV + kKBytes;                        // 2 stages of [BQ x HD]
    uint8_t* sDO = sQ + 2 * kQBytes;                   // 2 stages
    uint8_t* sPT = sDO + 2 * kQBytes;                  // P^T  [BK x BQ]
    uint8_t* sDST = sPT + kTBytes;                     // dS^T [BK x BQ]
    float* sLse = reinterpret_cast<float*>(sDST + kTBytes);   // 2 x BQ
    float* sDelta = sLse + 2 * BQ;                            // 2 x BQ
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDelta + 2 * BQ);
    uint64_t* kv_full = bars;        // 1
    uint64_t* kv_empty = bars + 1;   // 1
    uint64_t* q_full = bars + 2;     // 2
    uint64_t* q_empty = bars + 4;    // 2
    uint64_t* s_full = bars + 6;     // 1  (S^T and dP^T ready)
    uint64_t* t_full = bars + 7;     // 1  (P^T, dS^T in smem; S/dP TMEM read)
    uint64_t* t_empty = bars + 8;    // 1  (dV/dK MMAs that read P^T/dS^T have retired)
    uint64_t* acc_full = bars + 9;   // 1
    uint64_t* acc_empty = bars + 10; // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k_tiles = a.S / BK, rep = a.H / a.KV;
    const int items = a.B * a.KV * k_tiles;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::mbar_init(kv_full, 1); ptx::mbar_init(kv_empty, 1);
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&q_full[i], 1); ptx::mbar_init(&q_empty[i], 1); }
        ptx::mbar_init(s_full, 1); ptx::mbar_init(t_full, 4); ptx::mbar_init(t_empty, 1);
        ptx::mbar_init(acc_full, 1); ptx::mbar_init(acc_empty, 4);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    auto decode = [&](int it, int& b, int& g, int& kt) {     // small kt (long query loops) first
        kt = it / (a.B * a.KV);
        const int r = it % (a.B * a.KV);
        b = r / a.KV; g = r % a.KV;
    };
    // inner iterations of an item: for hh in 0..rep-1, for qtile in first..last   (BQ = 64: two per 128-key tile row)
    const int q_tiles64 = a.S / BQ;

    if (warp == 0) {
        if (ptx::elect_one()) {
            int item_it = 0, q_it = 0;
            for (int it = blockIdx.x; it < items; it += gridDim.x, ++item_it) {
                int b, g, kt; decode(it, b, g, kt);
                ptx::mbar_wait(kv_empty, (item_it & 1) ^ 1);
                ptx::mbar_arrive_expect_tx(kv_full, 2 * kKBytes);
                const int krow = b * a.S + kt * BK;
#pragma unroll
                for (int c = 0; c < HD / 64; ++c) {
                    ptx::tma_load_2d(sK + c * BK * 128, &tmQKV, kv_full, (a.H + g) * HD + c * 64, krow);
                    ptx::tma_load_2d(sV + c * BK * 128, &tmQKV, kv_full, (a.H + a.KV + g) * HD + c * 64, krow);
                }
                for (int hh = 0; hh < rep; ++hh) {
                    const int h = g * rep + hh;
                    for (int qt = kt * (BK / BQ); qt < q_tiles64; ++qt, ++q_it) {
                        const int st = q_it & 1;
                        ptx::mbar_wait(&q_empty[st], ((q_it >> 1) & 1) ^ 1);
                        ptx::mbar_arrive_expect_tx(&q_full[st], 2 * kQBytes);
                        const int qrow = b * a.S + qt * BQ;
#pragma unroll
                        for (int c = 0; c < HD / 64; ++c) {
                            ptx::tma_load_2d(sQ + st * kQBytes + c * BQ * 128, &tmQKV64, &q_full[st], h * HD + c * 64, qrow);
                            ptx::tma_load_2d(sDO + st * kQBytes + c * BQ * 128, &tmDO64, &q_full[st], h * HD + c * 64, qrow);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_st = ptx::umma_idesc_bf16(BK, BQ, false, false);   // S^T = K Q^T, dP^T = V dO^T
        constexpr uint32_t idesc_acc = ptx::umma_idesc_bf16(BK, HD, false, true);   // dV = P^T dO, dK = dS^T Q (B MN-major)
        int item_it = 0, q_it = 0, n_it = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x, ++item_it) {
            int b, g, kt; decode(it, b, g, kt);
            ptx::mbar_wait(kv_full, item_it & 1);
            ptx::mbar_wait(acc_empty, (item_it & 1) ^ 1);
            const int inner = rep * (q_tiles64 - kt * (BK / BQ));
            for (int n = 0; n < inner; ++n, ++q_it, ++n_it) {
                const int st = q_it & 1;
                ptx::mbar_wait(&q_full[st], (q_it >> 1) & 1);
                ptx::mbar_wait(t_empty, (n_it & 1) ^ 1);   // previous dV/dK MMAs retired: S/dP TMEM and P^T/dS^T smem are free
                ptx::tc_fence_after();
                const uint32_t ka = ptx::smem_u32(sK), va = ptx::smem_u32(sV);
                const uint32_t qb = ptx::smem_u32(sQ + st * kQBytes), dob = ptx::smem_u32(sDO + st * kQBytes);
                if (ptx::elect_one()) {
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        ptx::umma_bf16(tmem + kTmS, desc_kmajor(ka, BK, k), desc_kmajor(qb, BQ, k), idesc_st, k != 0);
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        ptx::umma_bf16(tmem + kTmDP, desc_kmajor(va, BK, k), desc_kmajor(dob, BQ, k), idesc_st, k != 0);
                    ptx::umma_commit(s_full);
                }
                __syncwarp();
                ptx::mbar_wait(t_full, n_it & 1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t pt = ptx::smem_u32(sPT), dst = ptx::smem_u32(sDST);
#pragma unroll
                    for (int k = 0; k < BQ / 16; ++k)
                        ptx::umma_bf16(tmem + kTmDV, desc_kmajor(pt, BK, k), desc_mnmajor(dob, BQ, k), idesc_acc, (n | k) != 0);
#pragma unroll
                    for (int k = 0; k < BQ / 16; ++k)
                        ptx::umma_bf16(tmem + kTmDK, desc_kmajor(dst, BK, k), desc_mnmajor(qb, BQ, k), idesc_acc, (n | k) != 0);
                    ptx::umma_commit(t_empty);
                    ptx::umma_commit(&q_empty[st]);
                    if (n == inner - 1) { ptx::umma_commit(acc_full); ptx::umma_commit(kv_empty); }
                }
                __syncwarp();
            }
        }
    } else {
        const int q4 = warp & 3;
        const int r = q4 * 32 + lane;                                  // key row inside the tile = TMEM lane
        const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
        const int tid = threadIdx.x - 64;                              // 0..127
        int item_it = 0, n_it = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x, ++item_it) {
            int b, g, kt; decode(it, b, g, kt);
            const int kj = kt * BK + r;
            for (int hh = 0; hh < rep; ++hh) {
                const int h = g * rep + hh;
                for (int qt = kt * (BK / BQ); qt < q_tiles64; ++qt, ++n_it) {
                    // stage this query tile's lse / delta (read by every thread below)
                    float* lse_s = sLse + (n_it & 1) * BQ;
                    float* del_s = sDelta + (n_it & 1) * BQ;
                    if (tid < BQ) {
                        const int64_t o = ((int64_t)b * a.H + h) * a.S + qt * BQ + tid;
                        lse_s[tid] = a.lse[o];
                        del_s[tid] = a.delta[o];
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    ptx::mbar_wait(s_full, n_it & 1);
                    ptx::tc_fence_after();
                    const bool diag = qt * BQ < kt * BK + BK;          // tile touches the diagonal: mask q < key
                    const uint32_t ptb = ptx::smem_u32(sPT), dstb = ptx::smem_u32(sDST);
#pragma unroll 1
                    for (int c = 0; c < BQ; c += 32) {
                        uint32_t sv[32], dv[32];
                        ptx::tmem_ld32(tmem + kTmS + lane_addr + c, sv);
                        ptx::tmem_ld32(tmem + kTmDP + lane_addr + c, dv);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            float p[8], ds[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int col = c + u * 8 + i;
                                float e = exp2f(__uint_as_float(sv[u * 8 + i]) * a.scale_log2 - lse_s[col]);
                                if (diag && qt * BQ + col < kj) e = 0.f;
                                p[i] = e;
                                ds[i] = e * (__uint_as_float(dv[u * 8 + i]) - del_s[col]);
                            }
                            st_shared16(sw128_addr(ptb, BK, r, c + u * 8), pack8(p));
                            st_shared16(sw128_addr(dstb, BK, r, c + u * 8), pack8(ds));
                        }
                    }
                    ptx::tc_fence_before();
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(t_full);
                }
            }
            // epilogue of the item: dV, dK (scaled, inverse-rotated) -> d(qkv)
            ptx::mbar_wait(acc_full, item_it & 1);
            ptx::tc_fence_after();
            const int64_t trow = (int64_t)(b * a.S + kj) * a.ld_out;
#pragma unroll 1
            for (int c = 0; c < HD; c += 32) {
                uint32_t v[32], kk[32];
                ptx::tmem_ld32(tmem + kTmDV + lane_addr + c, v);
                ptx::tmem_ld32(tmem + kTmDK + lane_addr + c, kk);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    float fv[8], fk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        fv[i] = __uint_as_float(v[u * 8 + i]);
                        fk[i] = __uint_as_float(kk[u * 8 + i]) * a.scale;
                    }
                    unrope8(fk, a, kj, c + u * 8, HD);
                    st8(a.out + trow + (a.H + a.KV + g) * HD + c + u * 8, pack8(fv));
                    st8(a.out + trow + (a.H + g) * HD + c + u * 8, pack8(fk));
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

// ================================================================================================= backward: dQ
template <int HD>
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmQKV64,
                   const __grid_constant__ CUtensorMap tmDO, const AttnArgs a) {
    constexpr int BM = 128, BN = 64;                   // query tile (UMMA M), key tile (UMMA N)
    constexpr int kQBytes = BM * HD * 2, kKBytes = BN * HD * 2, kSBytes = BM * BN * 2;
    constexpr int kTmS = 0, kTmDP = 64, kTmDQ = 128;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sDO = sQ + kQBytes;
    uint8_t* sK = sDO + kQBytes;                       // 2 stages [BN x HD]
    uint8_t* sV = sK + 2 * kKBytes;                    // 2 stages
    uint8_t* sDS = sV + 2 * kKBytes;                   // dS [BM x BN]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + kSBytes);
    uint64_t* q_full = bars;         // 1
    uint64_t* q_empty = bars + 1;    // 1
    uint64_t* kv_full = bars + 2;    // 2
    uint64_t* kv_empty = bars + 4;   // 2
    uint64_t* s_full = bars + 6;     // 1
    uint64_t* t_full = bars + 7;     // 1
    uint64_t* t_empty = bars + 8;    // 1
    uint64_t* acc_full = bars + 9;   // 1
    uint64_t* acc_empty = bars + 10; // 1
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_tiles = a.S / BM, rep = a.H / a.KV;
    const int items = a.B * a.H * q_tiles;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::mbar_init(q_full, 1); ptx::mbar_init(q_empty, 1);
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); }
        ptx::mbar_init(s_full, 1); ptx::mbar_init(t_full, 4); ptx::mbar_init(t_empty, 1);
        ptx::mbar_init(acc_full, 1); ptx::mbar_init(acc_empty, 4);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<256>(tmem_slot);
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
            int item_it = 0, kv_it = 0;
            for (int it = blockIdx.x; it < items; it += gridDim.x, ++item_it) {
                int b, h, qt; decode(it, b, h, qt);
                const int g = h / rep, row0 = b * a.S + qt * BM;
                ptx::mbar_wait(q_empty, (item_it & 1) ^ 1);
                ptx::mbar_arrive_expect_tx(q_full, 2 * kQBytes);
#pragma unroll
                for (int c = 0; c < HD / 64; ++c) {
                    ptx::tma_load_2d(sQ + c * BM * 128, &tmQKV, q_full, h * HD + c * 64, row0);
                    ptx::tma_load_2d(sDO + c * BM * 128, &tmDO, q_full, h * HD + c * 64, row0);
                }
                const int nkv = (qt + 1) * (BM / BN);
                for (int j = 0; j < nkv; ++j, ++kv_it) {
                    const int st = kv_it & 1;
                    ptx::mbar_wait(&kv_empty[st], ((kv_it >> 1) & 1) ^ 1);
                    ptx::mbar_arrive_expect_tx(&kv_full[st], 2 * kKBytes);
                    const int krow = b * a.S + j * BN;
#pragma unroll
                    for (int c = 0; c < HD / 64; ++c) {
                        ptx::tma_load_2d(sK + st * kKBytes + c * BN * 128, &tmQKV64, &kv_full[st], (a.H + g) * HD + c * 64, krow);
                        ptx::tma_load_2d(sV + st * kKBytes + c * BN * 128, &tmQKV64, &kv_full[st], (a.H + a.KV + g) * HD + c * 64, krow);
                    }
                }
            }
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(BM, BN, false, false);
        constexpr uint32_t idesc_q = ptx::umma_idesc_bf16(BM, HD, false, true);    // dQ = dS K, B = K MN-major
        int item_it = 0, kv_it = 0, n_it = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x, ++item_it) {
            int b, h, qt; decode(it, b, h, qt);
            ptx::mbar_wait(q_full, item_it & 1);
            ptx::mbar_wait(acc_empty, (item_it & 1) ^ 1);
            const int nkv = (qt + 1) * (BM / BN);
            for (int j = 0; j < nkv; ++j, ++kv_it, ++n_it) {
                const int st = kv_it & 1;
                ptx::mbar_wait(&kv_full[st], (kv_it >> 1) & 1);
                ptx::mbar_wait(t_empty, (n_it & 1) ^ 1);
                ptx::tc_fence_after();
                const uint32_t qa = ptx::smem_u32(sQ), doa = ptx::smem_u32(sDO);
                const uint32_t kb = ptx::smem_u32(sK + st * kKBytes), vb = ptx::smem_u32(sV + st * kKBytes);
                if (ptx::elect_one()) {
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        ptx::umma_bf16(tmem + kTmS, desc_kmajor(qa, BM, k), desc_kmajor(kb, BN, k), idesc_s, k != 0);
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        ptx::umma_bf16(tmem + kTmDP, desc_kmajor(doa, BM, k), desc_kmajor(vb, BN, k), idesc_s, k != 0);
                    ptx::umma_commit(s_full);
                }
                __syncwarp();
                ptx::mbar_wait(t_full, n_it & 1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t dsa = ptx::smem_u32(sDS);
#pragma unroll
                    for (int k = 0; k < BN / 16; ++k)
                        ptx::umma_bf16(tmem + kTmDQ, desc_kmajor(dsa, BM, k), desc_mnmajor(kb, BN, k), idesc_q, (j | k) != 0);
                    ptx::umma_commit(t_empty);
                    ptx::umma_commit(&kv_empty[st]);
                    if (j == nkv - 1) { ptx::umma_commit(acc_full); ptx::umma_commit(q_empty); }
                }
                __syncwarp();
            }
        }
    } else {
        const int q4 = warp & 3;
        const int r = q4 * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
        int item_it = 0, n_it = 0;
        for (int it = blockIdx.x; it < items; it += gridDim.x, ++item_it) {
            int b, h, qt; decode(it, b, h, qt);
            const int qi = qt * BM + r;
            const int64_t so = ((int64_t)b * a.H + h) * a.S + qi;
            const float lse = a.lse[so], delta = a.delta[so];
            const int nkv = (qt + 1) * (BM / BN);
            for (int j = 0; j < nkv; ++j, ++n_it) {
                ptx::mbar_wait(s_full, n_it & 1);
                ptx::tc_fence_after();
                const bool diag = j * BN + BN > qt * BM;
                const uint32_t dsb = ptx::smem_u32(sDS);
#pragma unroll 1
                for (int c = 0; c < BN; c += 32) {
                    uint32_t sv[32], dv[32];
                    ptx::tmem_ld32(tmem + kTmS + lane_addr + c, sv);
                    ptx::tmem_ld32(tmem + kTmDP + lane_addr + c, dv);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        float ds[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int col = c + u * 8 + i;
                            float e = exp2f(__uint_as_float(sv[u * 8 + i]) * a.scale_log2 - lse);
                            if (diag && j * BN + col > qi) e = 0.f;
                            ds[i] = e * (__uint_as_float(dv[u * 8 + i]) - delta);
                        }
                        st_shared16(sw128_addr(dsb, BM, r, c + u * 8), pack8(ds));
                    }
                }
                ptx::tc_fence_before();
                ptx::fence_proxy_async();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(t_full);
            }
            ptx::mbar_wait(acc_full, item_it & 1);
            ptx::tc_fence_after();
            __nv_bfloat16* orow = a.out + (int64_t)(b * a.S + qi) * a.ld_out + h * HD;
#pragma unroll 1
            for (int c = 0; c < HD; c += 32) {
                uint32_t v[32];
                ptx::tmem_ld32(tmem + kTmDQ + lane_addr + c, v);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    float f[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[u * 8 + i]) * a.scale;
                    unrope8(f, a, qi, c + u * 8, HD);
                    st8(orow + c + u * 8, pack8(f));
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(acc_empty);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<256>(tmem);
}

// ------------------------------------------------------------------------------------------------------- host
static PFN_cuTensorMapEncodeTiled_v12000 g_attn_encode = nullptr;
static std::once_flag g_attn_once;
static void attn_load_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
        g_attn_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
}
static int attn_tmap(CUtensorMap* out, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    std::call_once(g_attn_once, attn_load_encode);
    if (!g_attn_encode) return (int)cudaErrorNotSupported;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t es[2] = {1, 1};
    CUresult r = g_attn_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, es,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

template <typename K>
static int set_smem(K kern, int bytes) {
    return (int)cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

}  // namespace ftl

using namespace ftl;

// qkv: [B*S, ld_qkv] fused projection (q heads | k heads | v heads), RoPE already applied to q and k.
// out: [B*S, H*hd]; lse: fp32 [B, H, S].  S must be a multiple of 128; hd in {64, 128}.
FTL_API int ftl_attn_fwd(const void* qkv, int64_t ld_qkv, void* out, int64_t ld_out, float* lse, int B, int S, int H, int KV,
                         int hd, cudaStream_t stream) {
    if (S % 128 || (hd != 64 && hd != 128) || H % KV || ld_qkv % 8 || ld_out % 8) return (int)cudaErrorInvalidValue;
    CUtensorMap tm;
    int e = attn_tmap(&tm, qkv, (int64_t)B * S, (int64_t)(H + 2 * KV) * hd, ld_qkv, 128);
    if (e) return e;
    AttnArgs a{};
    a.out = (__nv_bfloat16*)out; a.ld_out = ld_out; a.lse = lse;
    a.B = B; a.S = S; a.H = H; a.KV = KV; a.ld_qkv = (int)ld_qkv;
    a.scale = 1.f / sqrtf((float)hd); a.scale_log2 = a.scale * kLog2e;
    const int items = B * H * (S / 128);
    const int grid = items < kNumSMs ? items : kNumSMs;
    if (hd == 128) {
        constexpr int smem = 128 * 128 * 2 * 6 + 1024 + 256;
        static bool once = false; if (!once) { if ((e = set_smem(attn_fwd_kernel<128>, smem))) return e; once = true; }
        attn_fwd_kernel<128><<<grid, kAttnThreads, smem, stream>>>(tm, a);
    } else {
        constexpr int smem = 128 * 64 * 2 * 5 + 128 * 128 * 2 + 1024 + 256;
        static bool once = false; if (!once) { if ((e = set_smem(attn_fwd_kernel<64>, smem))) return e; once = true; }
        attn_fwd_kernel<64><<<grid, kAttnThreads, smem, stream>>>(tm, a);
    }
    return (int)cudaGetLastError();
}

// d_out: [B*S, H*hd] gradient of the attention output; out: the forward output; writes dq | dk | dv into dqkv
// ([B*S, ld_qkv]) with the inverse RoPE applied to dq and dk when cos/sin are given.  delta: fp32 scratch [B, H, S].
FTL_API int ftl_attn_bwd(const void* qkv, int64_t ld_qkv, const void* out, const void* d_out, int64_t ld_out, const float* lse,
                         float* delta, void* dqkv, const float* cos_t, const float* sin_t, int B, int S, int H, int KV, int hd,
                         cudaStream_t stream) {
    if (S % 128 || (hd != 64 && hd != 128) || H % KV || ld_qkv % 8 || ld_out % 8) return (int)cudaErrorInvalidValue;
    const int64_t T = (int64_t)B * S;
    CUtensorMap tq128, tq64, tdo128, tdo64;
    int e;
    if ((e = attn_tmap(&tq128, qkv, T, (int64_t)(H + 2 * KV) * hd, ld_qkv, 128))) return e;
    if ((e = attn_tmap(&tq64, qkv, T, (int64_t)(H + 2 * KV) * hd, ld_qkv, 64))) return e;
    if ((e = attn_tmap(&tdo128, d_out, T, (int64_t)H * hd, ld_out, 128))) return e;
    if ((e = attn_tmap(&tdo64, d_out, T, (int64_t)H * hd, ld_out, 64))) return e;
    {
        const int64_t warps = T * H;
        attn_delta_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, stream>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)d_out,
                                                                                   delta, B, S, H, hd, ld_out);
        FTL_LAUNCH_CHECK();
    }
    AttnArgs a{};
    a.cos_t = cos_t; a.sin_t = sin_t;
    a.out = (__nv_bfloat16*)dqkv; a.ld_out = ld_qkv; a.lse = const_cast<float*>(lse); a.delta = delta;
    a.B = B; a.S = S; a.H = H; a.KV = KV; a.ld_qkv = (int)ld_qkv;
    a.scale = 1.f / sqrtf((float)hd); a.scale_log2 = a.scale * kLog2e;
    const int items_kv = B * KV * (S / 128), items_q = B * H * (S / 128);
    if (hd == 128) {
        constexpr int smem_kv = 2 * 128 * 128 * 2 + 4 * 64 * 128 * 2 + 2 * 128 * 64 * 2 + 4 * 64 * 4 + 1024 + 256;
        constexpr int smem_q = 2 * 128 * 128 * 2 + 4 * 64 * 128 * 2 + 128 * 64 * 2 + 1024 + 256;
        static bool once = false;
        if (!once) {
            if ((e = set_smem(attn_bwd_dkv_kernel<128>, smem_kv))) return e;
            if ((e = set_smem(attn_bwd_dq_kernel<128>, smem_q))) return e;
            once = true;
        }
        attn_bwd_dkv_kernel<128><<<items_kv < kNumSMs ? items_kv : kNumSMs, kAttnThreads, smem_kv, stream>>>(tq128, tq64, tdo64, a);
        FTL_LAUNCH_CHECK();
        attn_bwd_dq_kernel<128><<<items_q < kNumSMs ? items_q : kNumSMs, kAttnThreads, smem_q, stream>>>(tq128, tq64, tdo128, a);
    } else {
        constexpr int smem_kv = 2 * 128 * 64 * 2 + 4 * 64 * 64 * 2 + 2 * 128 * 64 * 2 + 4 * 64 * 4 + 1024 + 256;
        constexpr int smem_q = 2 * 128 * 64 * 2 + 4 * 64 * 64 * 2 + 128 * 64 * 2 + 1024 + 256;
        static bool once = false;
        if (!once) {
            if ((e = set_smem(attn_bwd_dkv_kernel<64>, smem_kv))) return e;
            if ((e = set_smem(attn_bwd_dq_kernel<64>, smem_q))) return e;
            once = true;
        }
        attn_bwd_dkv_kernel<64><<<items_kv < kNumSMs ? items_kv : kNumSMs, kAttnThreads, smem_kv, stream>>>(tq128, tq64, tdo64, a);
        FTL_LAUNCH_CHECK();
        attn_bwd_dq_kernel<64><<<items_q < kNumSMs ? items_q : kNumSMs, kAttnThreads, smem_q, stream>>>(tq128, tq64, tdo128, a);
    }
    return (int)cudaGetLastError();
}
