// Fused attention forward, second generation (head_dim 64: the CLIP blocks) for sm_100a.
// Replaces the materialised softmax(q k^T) v of clip/myAtt.py:53-64; same interface and results as attn_fwd_kernel<64>
// (attention.cu), which stays for head_dim 32 and as the A/B reference (WC_ATTN2=0).
//
// Why a second kernel: the first one splits every 128-key tile between two softmax warps per query row (64 keys each), so the
// fixed per-tile work (running maximum, vote, barrier probes, publish: ~850 of ~1750 clk per tile) is paid per 64 keys, P
// goes through a swizzled shared-memory tile (st.shared + fence.proxy.async + 32 KB of smem traffic per tile) and the two
// halves need two accumulators and a merge.  Measured against the library: F.scaled_dot_product_attention runs the same
// shapes 1.4x faster (profiles/r02_library_cmp.txt).  Here:
//   * one CTA owns TWO 128-row query tiles of one (image, head); softmax warp-group t (4 warps, thread = one query row,
//     ALL 128 keys of a tile) serves tile t, so the per-tile fixed cost is paid once per 128 keys and no merge exists;
//   * P never touches shared memory: it is written as packed fp16 into TMEM columns of its own (tcgen05.st) and the P V
//     product takes its A operand from there (tcgen05.mma with [a_tmem]).  P does not alias S, so the issuer puts the NEXT
//     scores of a tile in front of its P V product: the softmax group's wait is one Q K^T, not P V + Q K^T;
//   * one sweep over the score row per tile: the exponentials use the running maximum of the earlier tiles while the tile's
//     own maximum is collected on the side, and the sweep is repeated only when that maximum would take P out of fp16 range
//     (first tile, or a row whose maximum grows by more than 2^14) — S stays in tensor memory until the hand-over.  A thread
//     never holds its whole row (10 warps share 64 K registers: 168 each): it streams it 32 columns at a time;
//   * the two warp-groups ping-pong: while group 0 exponentiates tile (0, j) the tensor core runs the products of tile 1,
//     and one warp of each group shares an SM sub-partition, so its MUFU always has a customer;
//   * the running maximum is lazy as before (rescale own accumulator row only when it grows by more than 2^8).
// Measured floors (tools/micro/mma_rate.cu, B200): Q K^T 128x128x16 issues every 96 clk and P V 128x64x16 every 62 clk (operand
// fetch, not the 64 / 32 clk math floor): 745 clk of tensor pipe per 128x128 tile; MUFU.EX2 at 16 / clk / SM: 1024 clk per tile.
// Tensor memory (all 512 columns): S0 @0, S1 @128 (fp32 scores), P0 @256, P1 @320 (packed fp16), O0 @384, O1 @448 (fp32 [128 x 64]).
#include "attention.cuh"
#include "gemm.cuh"
#include "tc_ptx.cuh"

#include <stdlib.h>
#include <type_traits>

namespace wc {

constexpr int A2_BM = 128, A2_BN = 128, A2_HD = 64;
constexpr int A2_KST = 4, A2_VST = 4;                     // K / V ring depth
constexpr int A2_TILE = 128 * 64 * 2;                     // one [128][64] fp16 tile
constexpr int A2_THREADS = 320;                           // warps 0-3 / 4-7: softmax groups 0 / 1; warp 8 producer, 9 MMA
constexpr float A2_LOG2E = 1.4426950408889634f;
constexpr float A2_RESCALE_THR = 8.f;                     // lazy running maximum: a row rescales only when its maximum grows by > 2^8
constexpr float A2_REDO_THR = 14.f;                       // ... and a sweep is repeated only if P would leave 2^14 (fp16: 2^16)
constexpr uint32_t A2_TM_S = 0, A2_TM_P = 256, A2_TM_O = 384;     // TMEM columns: S0 S1 (fp32 128 each) | P0 P1 (64) | O0 O1 (64)

struct A2Smem {
    static constexpr int Q = 0;                                       // [2 item buffers][2 tiles]
    static constexpr int K = Q + 4 * A2_TILE;
    static constexpr int V = K + A2_KST * A2_TILE;
    static constexpr int BARS = V + A2_VST * A2_TILE;
    static constexpr int TOTAL = BARS + 512 + 1024;
};

struct A2Args {
    int L, heads, n_kt, n_qt, n_pairs, n_items, n_bh, n_full;     // n_full: items that own two query tiles (they come first)
    h16* o;
    float* lse2;
};

__device__ __forceinline__ float a2_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// D[tmem] (+)= A[tmem] * B[smem]: the A operand (P, packed fp16, K-major) comes from tensor memory
__device__ __forceinline__ void a2_umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

__global__ void __launch_bounds__(A2_THREADS, 1) attn_fwd2_kernel(const __grid_constant__ CUtensorMap mapQ,
                                                                   const __grid_constant__ CUtensorMap mapK,
                                                                   const __grid_constant__ CUtensorMap mapV,
                                                                   const __grid_constant__ A2Args a) {
    extern __shared__ uint8_t a2_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(a2_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A2Smem::BARS);
    uint64_t* q_full = bars;                    // [2]
    uint64_t* q_empty = q_full + 2;             // [2]
    uint64_t* k_full = q_empty + 2;             // [KST]
    uint64_t* k_empty = k_full + A2_KST;        // [KST]
    uint64_t* v_full = k_empty + A2_KST;        // [VST]
    uint64_t* v_empty = v_full + A2_VST;        // [VST]
    uint64_t* s_full = v_empty + A2_VST;        // [2]  S_t computed
    uint64_t* p_full = s_full + 2;              // [2]  P_t written to TMEM (and S_t read out) by the 4 warps of group t
    uint64_t* pv_done = p_full + 2;             // [2]  P V of (t, j) complete: P_t may be overwritten, O_t is quiescent
    uint64_t* o_empty = pv_done + 2;             // [2]  O_t read out by group t
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_kt = a.n_kt;

    if (warp == 8 && lane == 0) {
        prefetch_tmap(&mapQ); prefetch_tmap(&mapK); prefetch_tmap(&mapV);
        for (int b = 0; b < 2; ++b) { mbar_init(&q_full[b], 1); mbar_init(&q_empty[b], 1); }
        for (int s = 0; s < A2_KST; ++s) { mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); }
        for (int s = 0; s < A2_VST; ++s) { mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1); }
        for (int t = 0; t < 2; ++t) {
            mbar_init(&s_full[t], 1); mbar_init(&p_full[t], 4); mbar_init(&pv_done[t], 1); mbar_init(&o_empty[t], 4);
        }
        mbar_fence_init();
    }
    if (warp == 9) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // (10 warps = up to 3 per SM sub-partition of 16 K registers: 168 registers per thread.  A softmax thread therefore never
    // holds its whole 128-key score row: it streams the row from tensor memory twice, 32 columns at a time.)

    // item w of the launch = (pair of query tiles, head*image).  The two-tile items come first (pair fastest, so concurrently
    // running CTAs share K / V of one head in L2), the one-tile items of an odd tile count last: they fill the tail of the
    // static round-robin schedule at half the granularity.
    auto item_of = [&](int w, int& pair, int& bh) {
        if (w < a.n_full) { const int fp = a.n_qt >> 1; pair = w % fp; bh = w / fp; }
        else { pair = a.n_pairs - 1; bh = w - a.n_full; }
    };
    if (warp == 8) {
        // ---------------------------------------------------------------- TMA producer
        if (lane == 0) {
            int i = 0;
            for (int w = blockIdx.x; w < a.n_items; w += gridDim.x, ++i) {
                int pair, bh; item_of(w, pair, bh);
                const int qb = i & 1, g0 = i * n_kt;
                const int nt = (2 * pair + 1 < a.n_qt) ? 2 : 1;
                mbar_wait(&q_empty[qb], ((i >> 1) & 1) ^ 1);
                mbar_expect_tx(&q_full[qb], nt * A2_TILE);
                for (int t = 0; t < nt; ++t)
                    tma_load_4d(smem + A2Smem::Q + (qb * 2 + t) * A2_TILE, &mapQ, &q_full[qb], 0, (2 * pair + t) * A2_BM, bh, 0);
                for (int j = 0; j < n_kt; ++j) {
                    const int g = g0 + j, sk = g % A2_KST, sv = g % A2_VST;
                    mbar_wait(&k_empty[sk], ((g / A2_KST) & 1) ^ 1);
                    mbar_expect_tx(&k_full[sk], A2_TILE);
                    tma_load_4d(smem + A2Smem::K + sk * A2_TILE, &mapK, &k_full[sk], 0, j * A2_BN, bh, 0);
                    mbar_wait(&v_empty[sv], ((g / A2_VST) & 1) ^ 1);
                    mbar_expect_tx(&v_full[sv], A2_TILE);
                    tma_load_4d(smem + A2Smem::V + sv * A2_TILE, &mapV, &v_full[sv], 0, j * A2_BN, bh, 0);
                }
            }
        }
    } else if (warp == 9) {
        // ---------------------------------------------------------------- MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc_s = make_idesc(A2_BM, A2_BN);
            constexpr uint32_t idesc_o = make_idesc(A2_BM, A2_HD) | (1u << 16);      // B = V is MN-major (dims contiguous)
            constexpr uint32_t V_KSTEP = (16 * 128) >> 4;                            // 16 keys = 16 rows of 128 bytes
            int i = 0;
            uint32_t cp[2] = {0, 0}, co[2] = {0, 0};                                 // uses of p_full[t] / o_empty[t] so far
            for (int w = blockIdx.x; w < a.n_items; w += gridDim.x, ++i) {
                int pair, bh; item_of(w, pair, bh);
                const int qb = i & 1, g0 = i * n_kt;
                const int nt = (2 * pair + 1 < a.n_qt) ? 2 : 1;
                mbar_wait(&q_full[qb], (i >> 1) & 1);
                for (int t = 0; t < nt; ++t) { mbar_wait(&o_empty[t], (co[t] & 1) ^ 1); ++co[t]; }   // previous O_t read out
                tc_fence_after();
                auto issue_s = [&](int t, int sk) {
                    const uint64_t dq = make_smem_desc(smem_u32(smem + A2Smem::Q + (qb * 2 + t) * A2_TILE));
                    const uint64_t dk = make_smem_desc(smem_u32(smem + A2Smem::K + sk * A2_TILE));
#pragma unroll
                    for (int k = 0; k < A2_HD / 16; ++k) umma_f16(tmem_base + A2_TM_S + t * A2_BN, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                    umma_commit(&s_full[t]);
                };
                {   // scores of key tile 0
                    const int sk = g0 % A2_KST;
                    mbar_wait(&k_full[sk], (g0 / A2_KST) & 1);
                    tc_fence_after();
                    for (int t = 0; t < nt; ++t) issue_s(t, sk);
                    umma_commit(&k_empty[sk]);
                    if (n_kt == 1) umma_commit(&q_empty[qb]);
                }
                for (int j = 0; j < n_kt; ++j) {
                    const int g = g0 + j, sv = g % A2_VST, skn = (g + 1) % A2_KST;
                    const bool more = j + 1 < n_kt;
                    const uint64_t dv = make_smem_desc(smem_u32(smem + A2Smem::V + sv * A2_TILE));
                    for (int t = 0; t < nt; ++t) {
#ifdef A2_TRACE
                        uint32_t* trm = reinterpret_cast<uint32_t*>(a.lse2) + (size_t)a.n_bh * a.L + 8192 + (j * 2 + t) * 4;
                        const bool trm_on = blockIdx.x == 0 && i == 1;
                        if (trm_on) trm[0] = (uint32_t)clock64();
#endif
                        // P_t(j) is in its own TMEM columns and S_t(j) has been read out: the next scores go first (the softmax
                        // group is idle until they exist), the P V product of this tile runs behind them
                        mbar_wait(&p_full[t], cp[t] & 1); ++cp[t];
#ifdef A2_TRACE
                        if (trm_on) trm[1] = (uint32_t)clock64();
#endif
                        if (more) {
                            if (t == 0) mbar_wait(&k_full[skn], ((g + 1) / A2_KST) & 1);
                            tc_fence_after();
                            issue_s(t, skn);
                        }
#ifdef A2_TRACE
                        if (trm_on) trm[3] = (uint32_t)clock64();
#endif
                        if (t == 0) mbar_wait(&v_full[sv], (g / A2_VST) & 1);
                        tc_fence_after();
                        const uint32_t tp = tmem_base + A2_TM_P + t * (A2_BN / 2);   // P_t: packed fp16, 64 columns
                        const uint32_t to = tmem_base + A2_TM_O + t * A2_HD;
#pragma unroll
                        for (int k = 0; k < A2_BN / 16; ++k) a2_umma_ts(to, tp + 8 * k, dv + V_KSTEP * k, idesc_o, (j | k) != 0);
                        umma_commit(&pv_done[t]);
#ifdef A2_TRACE
                        if (trm_on) trm[2] = (uint32_t)clock64();
#endif
                    }
                    umma_commit(&v_empty[sv]);
                    if (more) {
                        umma_commit(&k_empty[skn]);
                        if (j + 2 == n_kt) umma_commit(&q_empty[qb]);                // the item's last scores have been issued
                    }
                }
            }
        }
    } else {
        // ---------------------------------------------------------------- softmax warp-groups
        const int t = warp >> 2;                            // query tile of the pair this group serves
        const int q = warp & 3;                             // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;
        const uint32_t lanes = (uint32_t)(q * 32) << 16;
        const uint32_t ts = tmem_base + lanes + A2_TM_S + t * A2_BN;        // S_t
        const uint32_t tp = tmem_base + lanes + A2_TM_P + t * (A2_BN / 2);  // P_t
        const uint32_t to = tmem_base + lanes + A2_TM_O + t * A2_HD;        // O_t
        uint32_t cs = 0, cpv = 0;                           // uses of s_full[t] / pv_done[t] so far
        for (int w = blockIdx.x; w < a.n_items; w += gridDim.x) {
            int pair, bh; item_of(w, pair, bh);
            const int head = bh % a.heads, img = bh / a.heads;
            if (2 * pair + t >= a.n_qt) continue;           // an odd last pair: group 1 has no tile
            const int grow = (2 * pair + t) * A2_BM + row;
            float m_used = -INFINITY, l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
#ifdef A2_TRACE
            // developer builds only (tools/attn_trace.py): clock stamps of warps 0 / 4 for the second item of CTA 0
            const bool tr_on = blockIdx.x == 0 && w == (int)gridDim.x && q == 0 && lane == 0;
            uint32_t* tr = reinterpret_cast<uint32_t*>(a.lse2) + (size_t)a.n_bh * a.L + t * 4096;
#define A2_STAMP(i) if (tr_on) tr[j * 8 + (i)] = (uint32_t)clock64();
#else
#define A2_STAMP(i)
#endif
            for (int j = 0; j < n_kt; ++j) {
                uint32_t ra[32], rb[32], pk[32];
                A2_STAMP(0)
                mbar_wait(&s_full[t], cs & 1); ++cs;
                A2_STAMP(1)
                tc_fence_after();
                const int nv = a.L - j * A2_BN;             // valid key columns of this tile (>= 1)
                bool pv_pending = j > 0;                    // P V of (t, j-1) still reads P_t and accumulates into O_t
                // One sweep over the score row with the running maximum of the EARLIER tiles: P = exp2(S log2e - m_used) while the
                // tile's own maximum is collected on the side.  Only if that maximum would take P towards the fp16 range
                // (> 2^14; always on the first tile, where m_used = -inf) is the sweep repeated with the new maximum — S_t stays
                // in tensor memory until the hand-over.  (Probabilities are relative to m_used, not to the true maximum: the
                // normalisation by l at the end is the same.)
                float t0, t1, t2, t3;
#ifndef A2_SKIP_SOFTMAX      // (developer experiment: tensor / TMA floor of the pipeline without the softmax arithmetic)
                for (int attempt = 0;; ++attempt) {
                    const float nm = -m_used;
                    float c0 = -INFINITY, c1 = -INFINITY, c2 = -INFINITY, c3 = -INFINITY;
                    t0 = t1 = t2 = t3 = 0.f;
                    auto chunk = [&](const uint32_t (&r)[32], int base, int pofs, auto masked) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            float s0 = __uint_as_float(r[2 * e]), s1 = __uint_as_float(r[2 * e + 1]);
                            if (decltype(masked)::value) {
                                if (base + 2 * e >= nv) s0 = -INFINITY;
                                if (base + 2 * e + 1 >= nv) s1 = -INFINITY;
                            }
                            const float p0 = a2_ex2(fmaf(s0, A2_LOG2E, nm)), p1 = a2_ex2(fmaf(s1, A2_LOG2E, nm));
                            if (e & 1) { c2 = fmaxf(c2, s0); c3 = fmaxf(c3, s1); t2 += p0; t3 += p1; }
                            else { c0 = fmaxf(c0, s0); c1 = fmaxf(c1, s1); t0 += p0; t1 += p1; }
                            const h16x2 h = __floats2half2_rn(p0, p1);
                            pk[pofs + e] = *reinterpret_cast<const uint32_t*>(&h);
                        }
                    };
                    auto do_chunk = [&](const uint32_t (&r)[32], int base, int pofs) {
                        if (base + 32 <= nv) chunk(r, base, pofs, std::false_type{});
                        else chunk(r, base, pofs, std::true_type{});
                    };
                    tmem_ld32_issue(ts, ra);
                    tmem_ld_wait(); tmem_ld32_issue(ts + 32, rb); do_chunk(ra, 0, 0);
                    tmem_ld_wait(); tmem_ld32_issue(ts + 64, ra); do_chunk(rb, 32, 16);
                    if (pv_pending) {                       // first write into P_t: the previous tile's P V must be through with it
                        mbar_wait(&pv_done[t], cpv & 1); ++cpv;
                        tc_fence_after();
                        pv_pending = false;
                    }
                    tmem_st32(tp, pk);
                    tmem_ld_wait(); tmem_ld32_issue(ts + 96, rb); do_chunk(ra, 64, 0);
                    tmem_ld_wait();                               do_chunk(rb, 96, 16);
                    tmem_st32(tp + 32, pk);
                    const float m_new = fmaxf(fmaxf(c0, c1), fmaxf(c2, c3)) * A2_LOG2E;
                    if (attempt || !__any_sync(0xffffffffu, m_new > m_used + A2_REDO_THR)) break;
                    const bool need = m_new > m_used + A2_RESCALE_THR;
                    if (j > 0) {                            // O_t is quiescent (pv_done waited above): rescale the own row
                        const float f = need ? a2_ex2(m_used - m_new) : 1.f;
#pragma unroll
                        for (int c = 0; c < A2_HD / 32; ++c) {
                            tmem_ld32_issue(to + c * 32, rb);
                            tmem_ld_wait();
#pragma unroll
                            for (int e = 0; e < 32; ++e) rb[e] = __float_as_uint(__uint_as_float(rb[e]) * f);
                            tmem_st32(to + c * 32, rb);
                        }
                        l0 *= f; l1 *= f; l2 *= f; l3 *= f;
                    }
                    if (need) m_used = m_new;
                }
                l0 += t0; l1 += t1; l2 += t2; l3 += t3;
#endif
                A2_STAMP(4)
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[t]);
                A2_STAMP(5)
            }
            // ---- epilogue: O_t / l -> fp16, head-interleaved [N*L, heads*64]; row log-sum-exp (log2 domain)
            const float l = (l0 + l1) + (l2 + l3);
            const float inv_l = 1.f / l;
            mbar_wait(&pv_done[t], cpv & 1); ++cpv;         // the item's last P V
            tc_fence_after();
            uint32_t r0[32], r1[32];
            tmem_ld32_issue(to, r0);
            tmem_ld32_issue(to + 32, r1);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&o_empty[t]);
            if (grow < a.L) {
                if (a.lse2) a.lse2[(size_t)bh * a.L + grow] = m_used + __log2f(l);
                h16* orow = a.o + ((size_t)img * a.L + grow) * (size_t)(a.heads * A2_HD) + head * A2_HD;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const uint32_t* r = c ? r1 : r0;
#pragma unroll
                    for (int v4 = 0; v4 < 4; ++v4) {
                        uint32_t wv[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const h16x2 h = __floats2half2_rn(__uint_as_float(r[8 * v4 + 2 * e]) * inv_l,
                                                              __uint_as_float(r[8 * v4 + 2 * e + 1]) * inv_l);
                            wv[e] = *reinterpret_cast<const uint32_t*>(&h);
                        }
                        *reinterpret_cast<uint4*>(orow + 32 * c + 8 * v4) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
                    }
                }
            }
        }
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// q, k, v: [n_img*heads, L, 64] fp16 (q pre-scaled); maps built by the caller (attention.cu: attn_maps)
int launch_attention_fwd2(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, h16* o, float* lse2, int n_img,
                          int heads, int L, int sms, cudaStream_t st) {
    WC_SET_MAX_SMEM(attn_fwd2_kernel, A2Smem::TOTAL);
    const int n_qt = ceil_div(L, A2_BM), n_pairs = (n_qt + 1) / 2;
    const long n_items = (long)n_pairs * heads * n_img;
    WC_CHECK_ARG(n_items < (1L << 30));
    const int n_bh = heads * n_img;
    A2Args a{L, heads, ceil_div(L, A2_BN), n_qt, n_pairs, (int)n_items, n_bh, (n_qt >> 1) * n_bh, o, lse2};
    const unsigned grid = (unsigned)(n_items < sms ? n_items : sms);
    attn_fwd2_kernel<<<grid, A2_THREADS, A2Smem::TOTAL, st>>>(mq, mk, mv, a);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

}  // namespace wc
