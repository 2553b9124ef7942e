// Fused attention forward, third variant (head_dim 64): 64-key tiles, TWO CTAs resident per SM (WC_ATTN2=2).
// Same interface and results as attn_fwd_kernel<64> (attention.cu); replaces the materialised softmax(q k^T) v of
// clip/myAtt.py:53-64.
//
// Why: neither of the two one-CTA-per-SM kernels saturates a unit (ncu r02: MUFU 47 %, tensor pipe 23 %, issue slots 46 %);
// they are bound by the latency of their barrier / TMEM / MUFU chains.  The same diagnosis on the PAR step and on the head-mean
// kernel was cured by residency (three / two CTAs per SM), so this kernel is built to be small: one 128-row query tile per
// CTA, 64-key tiles, thread = query row x 32 keys (8 softmax warps, 32 score registers), 256 TMEM columns, 96 KB of shared
// memory, <= 102 registers — two CTAs per SM, 16 softmax warps per SM, and one CTA's prologue / epilogue / barrier waits run
// under the other's arithmetic.
//   TMEM: S0 @0, S1 @64 (fp32 scores of a 64-key tile, double-buffered), O_A @128, O_B @192 (fp32 [128 x 64] accumulators of the
//         two key halves of every tile, merged in the epilogue as in attention.cu: each half keeps its own lazy running maximum).
//   smem: Q [128][64], K ring 3 x [64][64], V ring 3 x [64][64] (MN-major B operand), P 2 x [128][64] fp16 (128-byte swizzled
//         rows: half h writes the 16-byte chunks 4h .. 4h+3 of its rows).
#include "attention.cuh"
#include "gemm.cuh"
#include "tc_ptx.cuh"

namespace wc {

constexpr int A3_BM = 128, A3_BN = 64, A3_HD = 64, A3_ST = 3;
constexpr int A3_QBYTES = 128 * 64 * 2, A3_KBYTES = 64 * 64 * 2, A3_PBYTES = 128 * 64 * 2;
constexpr int A3_THREADS = 320;                       // warp 0 TMA, warp 1 MMA (+ TMEM), warps 2-9 softmax
constexpr float A3_LOG2E = 1.4426950408889634f, A3_RESCALE_THR = 8.f;

struct A3Smem {
    static constexpr int Q = 0;
    static constexpr int K = Q + A3_QBYTES;
    static constexpr int V = K + A3_ST * A3_KBYTES;
    static constexpr int P = V + A3_ST * A3_KBYTES;
    static constexpr int BARS = P + 2 * A3_PBYTES;
    static constexpr int STATS = BARS + 256;                 // [m | l][2 halves][128 rows] floats
    static constexpr int TOTAL = STATS + 2 * 2 * 128 * 4 + 1024;
};

struct A3Args {
    int L, heads, n_kt, n_qt;
    h16* o;
    float* lse2;
};

__device__ __forceinline__ float a3_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__global__ void __launch_bounds__(A3_THREADS, 2) attn_fwd3_kernel(const __grid_constant__ CUtensorMap mapQ,
                                                                   const __grid_constant__ CUtensorMap mapK,
                                                                   const __grid_constant__ CUtensorMap mapV,
                                                                   const __grid_constant__ A3Args a) {
    extern __shared__ uint8_t a3_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(a3_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A3Smem::BARS);
    uint64_t* q_full = bars;                    // [1]
    uint64_t* k_full = q_full + 1;              // [ST]
    uint64_t* k_empty = k_full + A3_ST;         // [ST]
    uint64_t* v_full = k_empty + A3_ST;         // [ST]
    uint64_t* v_empty = v_full + A3_ST;         // [ST]
    uint64_t* s_full = v_empty + A3_ST;         // [2]  S[b] computed             (MMA -> softmax)
    uint64_t* s_empty = s_full + 2;             // [2]  S[b] read into registers  (softmax -> MMA), 8 warps
    uint64_t* p_full = s_empty + 2;             // [2]  P[b] written              (softmax -> MMA), 8 warps
    uint64_t* p_empty = p_full + 2;             // [2]  P V of the tile complete  (MMA -> softmax): P[b] free, accumulators quiescent
    uint64_t* o_full = p_empty + 2;             // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);
    float* stats = reinterpret_cast<float*>(smem + A3Smem::STATS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = blockIdx.x % a.n_qt, bh = blockIdx.x / a.n_qt;      // query tile fastest: co-running CTAs share K / V in L2
    const int head = bh % a.heads, img = bh / a.heads;
    const int n_kt = a.n_kt;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&mapQ); prefetch_tmap(&mapK); prefetch_tmap(&mapV);
        mbar_init(q_full, 1);
        for (int s = 0; s < A3_ST; ++s) {
            mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&s_full[b], 1); mbar_init(&s_empty[b], 8); mbar_init(&p_full[b], 8); mbar_init(&p_empty[b], 1);
        }
        mbar_init(o_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tm_s = tmem_base, tm_oa = tmem_base + 128, tm_ob = tmem_base + 192;

    if (warp == 0) {
        // ---------------------------------------------------------------- TMA producer
        if (lane == 0) {
            mbar_expect_tx(q_full, A3_QBYTES);
            tma_load_4d(smem + A3Smem::Q, &mapQ, q_full, 0, qt * A3_BM, bh, 0);
            for (int j = 0; j < n_kt; ++j) {
                const int s = j % A3_ST, ph = ((j / A3_ST) & 1) ^ 1;
                mbar_wait(&k_empty[s], ph);
                mbar_expect_tx(&k_full[s], A3_KBYTES);
                tma_load_4d(smem + A3Smem::K + s * A3_KBYTES, &mapK, &k_full[s], 0, j * A3_BN, bh, 0);
                mbar_wait(&v_empty[s], ph);
                mbar_expect_tx(&v_full[s], A3_KBYTES);
                tma_load_4d(smem + A3Smem::V + s * A3_KBYTES, &mapV, &v_full[s], 0, j * A3_BN, bh, 0);
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer: S(j) one tile ahead of P V(j-1)
        if (lane == 0) {
            constexpr uint32_t idesc_s = make_idesc(A3_BM, A3_BN);
            constexpr uint32_t idesc_o = make_idesc(A3_BM, A3_HD) | (1u << 16);      // B = V is MN-major (dims contiguous)
            constexpr uint32_t V_KSTEP = (16 * 128) >> 4;                            // 16 keys = 16 rows of 128 bytes
            const uint64_t dq = make_smem_desc(smem_u32(smem + A3Smem::Q));
            auto pv = [&](int t) {
                const int bt = t & 1, st = t % A3_ST;
                mbar_wait(&p_full[bt], (t >> 1) & 1);
                mbar_wait(&v_full[st], (t / A3_ST) & 1);
                tc_fence_after();
                const uint64_t dp = make_smem_desc(smem_u32(smem + A3Smem::P + bt * A3_PBYTES));
                const uint64_t dv = make_smem_desc(smem_u32(smem + A3Smem::V + st * A3_KBYTES));
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)       // keys 0-31 of the tile -> O_A, 32-63 -> O_B
                    umma_f16(ks < 2 ? tm_oa : tm_ob, dp + 2 * ks, dv + V_KSTEP * ks, idesc_o, (t > 0) || (ks & 1));
                umma_commit(&p_empty[bt]);
                umma_commit(&v_empty[st]);
            };
            mbar_wait(q_full, 0);
            for (int j = 0; j < n_kt; ++j) {
                const int s = j % A3_ST, b = j & 1;
                mbar_wait(&k_full[s], (j / A3_ST) & 1);
                mbar_wait(&s_empty[b], ((j >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint64_t dk = make_smem_desc(smem_u32(smem + A3Smem::K + s * A3_KBYTES));
#pragma unroll
                for (int k = 0; k < A3_HD / 16; ++k) umma_f16(tm_s + b * A3_BN, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                umma_commit(&s_full[b]);
                umma_commit(&k_empty[s]);
                if (j >= 1) pv(j - 1);
            }
            pv(n_kt - 1);
            umma_commit(o_full);
        }
    } else {
        // ---------------------------------------------------------------- softmax warps: thread = query row x 32 keys of a tile
        const int q = warp & 3;                              // TMEM lane quarter this warp may access
        const int half = (warp - 2) >> 2;                    // key half of every tile (0: keys 0-31 -> O_A, 1: keys 32-63 -> O_B)
        const int row = q * 32 + lane, grow = qt * A3_BM + row;
        const uint32_t lanes = (uint32_t)(q * 32) << 16;
        const uint32_t ts = tm_s + lanes + half * 32;
        const uint32_t to = (half ? tm_ob : tm_oa) + lanes;
        float m_used = -INFINITY, l0 = 0.f, l1 = 0.f;
        for (int j = 0; j < n_kt; ++j) {
            const int b = j & 1;
            uint32_t r[32];
            mbar_wait(&s_full[b], (j >> 1) & 1);
            tc_fence_after();
            tmem_ld32_issue(ts + b * A3_BN, r);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[b]);         // S[b] is in registers
            const int nv = a.L - j * A3_BN - half * 32;      // this warp's columns of the tile that exist (may be <= 0)
            float c0 = -INFINITY, c1 = -INFINITY, c2 = -INFINITY, c3 = -INFINITY;
            if (nv >= 32) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    c0 = fmaxf(c0, __uint_as_float(r[i])); c1 = fmaxf(c1, __uint_as_float(r[i + 1]));
                    c2 = fmaxf(c2, __uint_as_float(r[i + 2])); c3 = fmaxf(c3, __uint_as_float(r[i + 3]));
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) if (i < nv) c0 = fmaxf(c0, __uint_as_float(r[i]));
            }
            const float m_new = fmaxf(fmaxf(c0, c1), fmaxf(c2, c3)) * A3_LOG2E;     // -inf when no column exists
            const bool need = m_new > m_used + A3_RESCALE_THR;                      // first valid tile: m_used = -inf
            if (__any_sync(0xffffffffu, need)) {
                if (j > 0) {
                    // the accumulator must be quiescent: P V of tile j-1 complete
                    mbar_wait(&p_empty[(j - 1) & 1], ((j - 1) >> 1) & 1);
                    tc_fence_after();
                    const float f = need ? a3_ex2(m_used - m_new) : 1.f;            // (m_used = -inf: the row is still all zero)
                    uint32_t t32[32];
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        tmem_ld32_issue(to + c * 32, t32);
                        tmem_ld_wait();
#pragma unroll
                        for (int e = 0; e < 32; ++e) t32[e] = __float_as_uint(__uint_as_float(t32[e]) * f);
                        tmem_st32(to + c * 32, t32);
                    }
                    tmem_st_wait();
                    tc_fence_before();
                    l0 *= f; l1 *= f;
                }
                if (need) m_used = m_new;
            }
            // P[b] is free once P V of tile j-2 has run
            mbar_wait(&p_empty[b], ((j >> 1) & 1) ^ 1);
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                float p0 = a3_ex2(fmaf(__uint_as_float(r[2 * i]), A3_LOG2E, -m_used));
                float p1 = a3_ex2(fmaf(__uint_as_float(r[2 * i + 1]), A3_LOG2E, -m_used));
                if (nv < 32) {
                    if (2 * i >= nv) p0 = 0.f;
                    if (2 * i + 1 >= nv) p1 = 0.f;
                }
                l0 += p0; l1 += p1;
                const h16x2 h = __floats2half2_rn(p0, p1);
                pk[i] = *reinterpret_cast<const uint32_t*>(&h);
            }
            uint8_t* prow = smem + A3Smem::P + b * A3_PBYTES + row * 128;
#pragma unroll
            for (int t = 0; t < 4; ++t)
                *reinterpret_cast<uint4*>(prow + (((4 * half + t) ^ (row & 7)) << 4)) =
                    make_uint4(pk[4 * t], pk[4 * t + 1], pk[4 * t + 2], pk[4 * t + 3]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[b]);
        }
        // ---- epilogue: merge the two key halves.  O = (c_A O_A + c_B O_B) / (c_A l_A + c_B l_B), c_h = 2^(m_h - max(m_A, m_B))
        const float l_own = l0 + l1;
        stats[(0 * 2 + half) * 128 + row] = m_used;
        stats[(1 * 2 + half) * 128 + row] = l_own;
        asm volatile("bar.sync 1, 256;" ::: "memory");       // the 8 softmax warps
        const float mA = stats[(0 * 2 + 0) * 128 + row], mB = stats[(0 * 2 + 1) * 128 + row];
        const float lA = stats[(1 * 2 + 0) * 128 + row], lB = stats[(1 * 2 + 1) * 128 + row];
        const float m = fmaxf(mA, mB);                       // finite: half A always sees key 0
        const float cA = a3_ex2(mA - m), cB = (mB == -INFINITY) ? 0.f : a3_ex2(mB - m);
        const float lsum = cA * lA + cB * lB;
        const float inv_l = 1.f / lsum;
        mbar_wait(o_full, 0);
        tc_fence_after();
        uint32_t ra[32], rb[32];
        tmem_ld32_issue(tm_oa + lanes + half * 32, ra);      // this warp stores head dims [32 half, 32 half + 32) of its rows
        tmem_ld32_issue(tm_ob + lanes + half * 32, rb);
        tmem_ld_wait();
        if (grow < a.L) {
            if (a.lse2 && half == 0) a.lse2[(size_t)bh * a.L + grow] = m + __log2f(lsum);
            h16* orow = a.o + ((size_t)img * a.L + grow) * (size_t)(a.heads * A3_HD) + head * A3_HD + half * 32;
            const float wa = cA * inv_l, wb = cB * inv_l;
#pragma unroll
            for (int v4 = 0; v4 < 4; ++v4) {
                uint32_t wv[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int i = 8 * v4 + 2 * e;
                    const h16x2 h = __floats2half2_rn(wa * __uint_as_float(ra[i]) + wb * __uint_as_float(rb[i]),
                                                      wa * __uint_as_float(ra[i + 1]) + wb * __uint_as_float(rb[i + 1]));
                    wv[e] = *reinterpret_cast<const uint32_t*>(&h);
                }
                *reinterpret_cast<uint4*>(orow + 8 * v4) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
            }
        }
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 256);
    }
}

// q, k, v: [n_img*heads, L, 64] fp16 (q pre-scaled); the Q map (128-row box) comes from the caller, K / V maps with 64-row
// boxes are built here
int launch_attention_fwd3(const CUtensorMap& mq, const h16* k, const h16* v, h16* o, float* lse2, int n_img, int heads, int L,
                          cudaStream_t st) {
    WC_SET_MAX_SMEM(attn_fwd3_kernel, A3Smem::TOTAL);
    CUtensorMap mk, mv;
    int u0, u1, rc;
    GemmOperand ok{k, L, A3_HD, (long)L * A3_HD, 0}, ov{v, L, A3_HD, (long)L * A3_HD, 0};
    if ((rc = make_tma_map(&mk, ok, A3_HD, n_img * heads, 1, A3_BN, &u0, &u1))) return rc;
    if ((rc = make_tma_map(&mv, ov, A3_HD, n_img * heads, 1, A3_BN, &u0, &u1))) return rc;
    const int n_qt = ceil_div(L, A3_BM);
    const long n_cta = (long)n_qt * heads * n_img;
    WC_CHECK_ARG(n_cta < (1L << 31));
    A3Args a{L, heads, ceil_div(L, A3_BN), n_qt, o, lse2};
    attn_fwd3_kernel<<<(unsigned)n_cta, A3_THREADS, A3Smem::TOTAL, st>>>(mq, mk, mv, a);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

}  // namespace wc
