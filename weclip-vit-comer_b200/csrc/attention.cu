// Fused multi-head attention forward (head_dim 64) for sm_100a.  See attention.cuh for the contract.
//
// attn_fwd_kernel — one CTA per (128-query tile, head, image), 320 threads, TWO passes over the key tiles:
//   pass 1  S = Q K^T per 128-key tile (tcgen05.mma, fp32 in TMEM) -> running row max / sum of exp2 (registers)
//   pass 2  S again, P = exp2(S*log2e - lse2) is the EXACT softmax probability (no rescaling of the output accumulator),
//           rounded to fp16 into a 128B-swizzled shared-memory tile, O += P V (tcgen05.mma, O stays in TMEM)
//   warp 0     TMA producer: Q once, then a 4-stage ring of K (pass 1) / K + V^T (pass 2) tiles, OOB rows zero-filled
//   warp 1     MMA issuer (one lane): S double-buffered in TMEM so QK^T of tile j+1 overlaps the softmax of tile j
//   warps 2-9  softmax: thread = query row (TMEM lane), two warps per lane quarter (64 key columns each);
//              tcgen05.ld 32 columns at a time, software-pipelined against the ex2.approx work (MUFU is the bound:
//              16 exp2/clk/SM measured, tools/micro/mufu_bench.cu)
// The score matrix never leaves the SM: HBM traffic is q, k, v once per query tile (L2-resident) + o.
//
// attn_headmean_kernel — the head-averaged probability map the CAM refinement consumes (myAtt.py:323-326): one CTA per
//   (key tile, query tile, image) loops over the heads, recomputes S = Q_h K_h^T on the tensor cores and accumulates
//   exp2(S*log2e - lse2_h) in registers; each [N,L,L] element is written once (4 L^2 bytes per image, not 12x).
#include "attention.cuh"
#include "gemm.cuh"
#include "tc_ptx.cuh"

#include <stdlib.h>

namespace wc {

constexpr int AT_BM = 128, AT_BN = 128, AT_HD = 64;
constexpr int AT_STAGES = 4;
constexpr int AT_SBUF = 3;             // score tiles buffered in tensor memory (3 x 128 columns + the output accumulator <= 512)
constexpr int AT_SM_WARPS = 8;          // softmax warps: two per TMEM lane quarter, each owning 64 of a tile's 128 key columns
constexpr int AT_THREADS = 64 + 32 * AT_SM_WARPS;
constexpr int FW_PV_WARP = 2 + AT_SM_WARPS;          // forward kernel: one more warp, the P V issuer (after the softmax warps)
constexpr int FW_THREADS = AT_THREADS + 32;
constexpr int AT_TILE_BYTES = 128 * 64 * 2;          // one [128][64] fp16 operand tile (Q, K, V^T pair, half of P)
constexpr float AT_LOG2E = 1.4426950408889634f;

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// 2^x on the FMA / ALU pipes instead of the MUFU (the softmax phase issues one MUFU.EX2 per score at 16 / clk / SM): x = n + f with
// n = round(x) taken from the low mantissa bits of x + 1.5 * 2^23, 2^f on [-0.5, 0.5] by the degree-4 Taylor polynomial (relative
// error < 4.3e-5, an order below the fp16 rounding P gets anyway), scaled by adding n into the exponent field.  x <= 127.
__device__ __forceinline__ float ex2_poly(float x) {
    x = fmaxf(x, -125.f);
    const float t = x + 12582912.f;
    const float f = x - (t - 12582912.f);
    float p = fmaf(f, 0.0096181291f, 0.055504109f);
    p = fmaf(p, f, 0.24022651f);
    p = fmaf(p, f, 0.69314718f);
    p = fmaf(p, f, 1.f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
#ifndef AT_POLY_EVERY
#define AT_POLY_EVERY 0          // 4: one score in four takes ex2_poly; 0: all MUFU
#endif
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// -DAT_TRACE (developer builds only, tools/attn_trace.py): clock64 stamps of one softmax warp pair for the second item of CTA 0,
// written past the end of the lse buffer (the tool allocates the extra room)
#ifdef AT_TRACE
#define AT_TR(slot) do { if (blockIdx.x == 0 && lane == 0 && it == 1 && j < 16) tr[j * 8 + (slot)] = (uint32_t)clock64(); } while (0)
#else
#define AT_TR(slot) do { } while (0)
#endif
struct AttnArgs {
    int L, heads, n_kt;
    int n_qt, n_items;       // work items = (query tile, head, image), walked persistently: item w -> qt = w % n_qt, bh = w / n_qt
    h16* o;
    float* lse2;
    long trace_off;          // AT_TRACE only: element offset of the trace area behind the lse rows
};

// ---------------------------------------------------------------------------------------------------
struct FwdSmem {
    static constexpr int V_STAGES = 3;
    static constexpr int Q = 0;                                              // two Q tiles (this item / next item)
    static constexpr int KRING = 2 * AT_TILE_BYTES;                          // AT_STAGES K tiles
    static constexpr int VRING = KRING + AT_STAGES * AT_TILE_BYTES;          // V_STAGES V^T tiles
    static constexpr int P = VRING + V_STAGES * AT_TILE_BYTES;               // two P buffers of 32 KB
    static constexpr int BARS = P + 2 * 2 * AT_TILE_BYTES;
    static constexpr int STATS = BARS + 512;                                 // [item parity][max | sum][2 halves][128 rows] floats
    static constexpr int TOTAL = STATS + 2 * 2 * 256 * 4 + 1024;             // + alignment slack
};

// HD = head dim (64: CLIP blocks, 32: decoder blocks).  Q / K tiles are always [128][64] fp16 in shared memory: for HD = 32
// the TMA box runs past the tensor's 32-element rows and the tail is zero-filled, and only HD/16 MMA k-steps are issued.
//
// ONE pass over the key tiles (online softmax).  Each softmax warp owns 32 query rows x one 64-key half of every tile and
// keeps its OWN running maximum m (log2 domain) and row sum l; the two halves accumulate into two separate output
// accumulators in tensor memory (O_A += P_A V_A, O_B += P_B V_B), so no per-tile exchange between warps is needed.  The
// maximum is updated lazily: only when a tile's maximum exceeds the one in use by more than 2^8 does the warp rescale its own
// rows of its accumulator (tcgen05.ld / st) — after the first tile this is rare — so P = exp2(S log2e - m) stays below 256 in
// fp16.  The epilogue merges the halves: O = (c_A O_A + c_B O_B) / (c_A l_A + c_B l_B), c_h = 2^(m_h - max(m_A, m_B)).
template <int HD>
__global__ void __launch_bounds__(FW_THREADS, 1) attn_fwd_kernel(const __grid_constant__ CUtensorMap mapQ,
                                                                  const __grid_constant__ CUtensorMap mapK,
                                                                  const __grid_constant__ CUtensorMap mapV,
                                                                  const __grid_constant__ AttnArgs a) {
    constexpr int V_BLK = HD * 128;                // one K-major [HD dims][64 keys] block of V^T
    constexpr int VST = FwdSmem::V_STAGES;
    constexpr float RESCALE_THR = 8.f;             // log2 units
    extern __shared__ uint8_t at_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + FwdSmem::BARS);
    uint64_t* q_full = bars;                       // [2]  Q of item i in buffer i & 1
    uint64_t* q_empty = q_full + 2;                // [2]  every Q K^T of the item has run
    uint64_t* k_full = q_empty + 2;                // [STAGES]  K ring: a slot is free again once its Q K^T has run
    uint64_t* k_empty = k_full + AT_STAGES;        // [STAGES]
    uint64_t* v_full = k_empty + AT_STAGES;        // [VST]     V ring: a slot is free again once its P V has run
    uint64_t* v_empty = v_full + VST;              // [VST]
    uint64_t* s_full = v_empty + VST;              // [3]  S[b] computed            (MMA -> softmax)
    uint64_t* s_empty = s_full + AT_SBUF;          // [3]  S[b] read into registers (softmax -> MMA)
    uint64_t* p_full = s_empty + AT_SBUF;          // [2]  P[b] written             (softmax -> MMA)
    uint64_t* p_empty = p_full + 2;                // [2]  P[b] consumed by P V     (MMA -> softmax)
    uint64_t* o_full = p_empty + 2;                // [1]  the item's accumulators are complete (MMA -> softmax)
    uint64_t* o_empty = o_full + 1;                // [1]  ... and have been read out             (softmax -> MMA)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_kt = a.n_kt;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&mapQ); prefetch_tmap(&mapK); prefetch_tmap(&mapV);
        for (int b = 0; b < 2; ++b) { mbar_init(&q_full[b], 1); mbar_init(&q_empty[b], 1); }
        for (int s = 0; s < AT_STAGES; ++s) { mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); }
        for (int s = 0; s < VST; ++s) { mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1); }
        for (int b = 0; b < AT_SBUF; ++b) { mbar_init(&s_full[b], 1); mbar_init(&s_empty[b], AT_SM_WARPS); }
        for (int b = 0; b < 2; ++b) { mbar_init(&p_full[b], AT_SM_WARPS); mbar_init(&p_empty[b], 1); }
        mbar_init(o_full, 1);
        mbar_init(o_empty, AT_SM_WARPS);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_o = tmem_base + AT_SBUF * AT_BN;     // O_A at +0, O_B at +HD

    // Persistent CTA: item i of this CTA is work item w = blockIdx.x + i * gridDim.x; consecutive w share (image, head), so the
    // CTAs running at the same time read the same K / V from L2.  All ring / barrier indices run on the GLOBAL tile counter
    // g = i * n_kt + j, so the loads, the score MMAs and the softmax of item i+1 start while item i drains (epilogue, output
    // stores): the Q / K load latency, the TMEM allocation and the launch gap are paid once per CTA, not once per item.
    if (warp == 0) {
        if (lane == 0) {
            int i = 0;
            for (int w = blockIdx.x; w < a.n_items; w += gridDim.x, ++i) {
                const int qt = w % a.n_qt, bh = w / a.n_qt, g0 = i * n_kt, qb = i & 1;
                mbar_wait(&q_empty[qb], ((i >> 1) & 1) ^ 1);
                mbar_expect_tx(&q_full[qb], AT_TILE_BYTES);
                tma_load_4d(smem + FwdSmem::Q + qb * AT_TILE_BYTES, &mapQ, &q_full[qb], 0, qt * AT_BM, bh, 0);
                // K and V^T travel in separate rings: a K slot is released by the score MMA (which runs two tiles ahead), a V
                // slot by the P V MMA, so both kinds of load are requested three to four tiles before their use
                auto load_v = [&](int j) {
                    const int g = g0 + j, s = g % VST;
                    mbar_wait(&v_empty[s], ((g / VST) & 1) ^ 1);
                    uint8_t* sv = smem + FwdSmem::VRING + s * AT_TILE_BYTES;
                    mbar_expect_tx(&v_full[s], 2 * V_BLK);
                    if (HD == 64) {
                        // V [128 keys][64 dims] straight from the head-major QKV output: one 128-byte row per key = the
                        // canonical MN-major (N = dims contiguous) 128B-swizzled B operand, no transposed copy of V anywhere
                        tma_load_4d(sv, &mapV, &v_full[s], 0, j * AT_BN, bh, 0);
                    } else {
                        // V^T [HD dims][128 keys] as two K-major [HD][64] blocks (box = 64 keys x HD dims)
                        tma_load_4d(sv, &mapV, &v_full[s], j * AT_BN, 0, bh, 0);
                        tma_load_4d(sv + V_BLK, &mapV, &v_full[s], j * AT_BN + 64, 0, bh, 0);
                    }
                };
                for (int j = 0; j < n_kt; ++j) {
                    const int g = g0 + j, s = g % AT_STAGES;
                    mbar_wait(&k_empty[s], ((g / AT_STAGES) & 1) ^ 1);
                    mbar_expect_tx(&k_full[s], AT_TILE_BYTES);
                    tma_load_4d(smem + FwdSmem::KRING + s * AT_TILE_BYTES, &mapK, &k_full[s], 0, j * AT_BN, bh, 0);
                    if (j >= 2) load_v(j - 2);
                }
                for (int j = (n_kt >= 2 ? n_kt - 2 : 0); j < n_kt; ++j) load_v(j);
            }
        }
    } else if (warp == 1) {
        // score issuer: S[g % 3] = Q K_j^T, as far ahead as the three S buffers allow.  The P V products have their OWN issuing
        // warp (below): one thread issuing all twelve MMAs of a tile plus its five barrier waits took ~2000 clk per tile
        // (clock64 trace), as long as the softmax itself, and tcgen05.commit only tracks the issuing thread's MMAs anyway.
        if (lane == 0) {
            constexpr uint32_t idesc_s = make_idesc(AT_BM, AT_BN);
            int i = 0;
            for (int w = blockIdx.x; w < a.n_items; w += gridDim.x, ++i) {
                const int g0 = i * n_kt, qb = i & 1;
                const uint64_t dq = make_smem_desc(smem_u32(smem + FwdSmem::Q + qb * AT_TILE_BYTES));
                mbar_wait(&q_full[qb], (i >> 1) & 1);
                for (int j = 0; j < n_kt; ++j) {
                    const int g = g0 + j, s = g % AT_STAGES, b = g % AT_SBUF;
                    mbar_wait(&k_full[s], (g / AT_STAGES) & 1);
                    mbar_wait(&s_empty[b], ((g / AT_SBUF) & 1) ^ 1);
                    tc_fence_after();
                    const uint64_t dk = make_smem_desc(smem_u32(smem + FwdSmem::KRING + s * AT_TILE_BYTES));
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k) umma_f16(tmem_base + b * AT_BN, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                    umma_commit(&s_full[b]);
                    umma_commit(&k_empty[s]);
                    if (j == n_kt - 1) umma_commit(&q_empty[qb]);
                }
            }
        }
    } else if (warp == FW_PV_WARP) {
        // P V issuer: O_A += P_A V_A, O_B += P_B V_B
        if (lane == 0) {
            // HD = 64: B = V is MN-major (instruction-descriptor bit 16); a 16-key MMA step then advances 16 rows of 128 bytes
            constexpr uint32_t idesc_o = make_idesc(AT_BM, HD) | (HD == 64 ? (1u << 16) : 0u);
            constexpr uint32_t V_KSTEP = (HD == 64) ? (16 * 128) >> 4 : 2;
            int i = 0;
            for (int w = blockIdx.x; w < a.n_items; w += gridDim.x, ++i) {
                const int g0 = i * n_kt;
                mbar_wait(o_empty, (i & 1) ^ 1);        // the previous item's accumulators have been read out
                for (int j = 0; j < n_kt; ++j) {
                    const int g = g0 + j, s = g % VST, pb = g & 1;
                    mbar_wait(&v_full[s], (g / VST) & 1);
                    mbar_wait(&p_full[pb], (g >> 1) & 1);
                    tc_fence_after();
                    const uint32_t sp = smem_u32(smem + FwdSmem::P + pb * 2 * AT_TILE_BYTES);
                    const uint32_t sv = smem_u32(smem + FwdSmem::VRING + s * AT_TILE_BYTES);
#pragma unroll
                    for (int kb = 0; kb < 2; ++kb) {   // key half kb -> accumulator kb
                        const uint64_t dp = make_smem_desc(sp + kb * AT_TILE_BYTES), dv = make_smem_desc(sv + kb * V_BLK);
#pragma unroll
                        for (int k = 0; k < 4; ++k) umma_f16(tmem_o + kb * HD, dp + 2 * k, dv + V_KSTEP * k, idesc_o, (j | k) != 0);
                    }
                    umma_commit(&p_empty[pb]);
                    umma_commit(&v_empty[s]);
                }
                umma_commit(o_full);
            }
        }
    } else {
        const int q = warp & 3;                              // TMEM lane quarter (hardware: warp id % 4)
        const int half = (warp - 2) >> 2;                    // which 64 key columns of every tile this warp owns
        const int row = q * 32 + lane;
        const uint32_t lanes = (uint32_t)(q * 32) << 16;
        const uint32_t tlane = tmem_base + lanes + half * 64;
        const uint32_t to = tmem_o + lanes + half * HD;      // this warp's rows of its own accumulator
        uint32_t r0[32], r1[32];
        float m_used, l0, l1, l2, l3;                        // maximum in use (log2 domain), row sum of P relative to it
        auto max_chunk = [&](const uint32_t (&r)[32], int nvc) {   // nvc: columns of this chunk that exist (>= 1)
            float c0 = -INFINITY, c1 = -INFINITY, c2 = -INFINITY, c3 = -INFINITY;
            if (nvc >= 32) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    c0 = fmaxf(c0, __uint_as_float(r[i])); c1 = fmaxf(c1, __uint_as_float(r[i + 1]));
                    c2 = fmaxf(c2, __uint_as_float(r[i + 2])); c3 = fmaxf(c3, __uint_as_float(r[i + 3]));
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) if (i < nvc) c0 = fmaxf(c0, __uint_as_float(r[i]));
            }
            return fmaxf(fmaxf(c0, c1), fmaxf(c2, c3));
        };
        auto prob_chunk = [&](const uint32_t (&r)[32], int nvc, uint8_t* blk, int cc0) {
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                float p0 = ex2_approx(fmaf(__uint_as_float(r[2 * i]), AT_LOG2E, -m_used));
                const float x1 = fmaf(__uint_as_float(r[2 * i + 1]), AT_LOG2E, -m_used);
                float p1 = (AT_POLY_EVERY && (2 * i + 1) % AT_POLY_EVERY == AT_POLY_EVERY - 1) ? ex2_poly(x1) : ex2_approx(x1);
                if (nvc < 32) {
                    if (2 * i >= nvc) p0 = 0.f;
                    if (2 * i + 1 >= nvc) p1 = 0.f;
                }
                if (i & 1) { l2 += p0; l3 += p1; } else { l0 += p0; l1 += p1; }
                const h16x2 h = __floats2half2_rn(p0, p1);
                pk[i] = *reinterpret_cast<const uint32_t*>(&h);
            }
#pragma unroll
            for (int t = 0; t < 4; ++t)
                *reinterpret_cast<uint4*>(blk + (((cc0 + t) ^ (row & 7)) << 4)) = make_uint4(pk[4 * t], pk[4 * t + 1], pk[4 * t + 2], pk[4 * t + 3]);
        };
#ifdef AT_TRACE
        uint32_t* tr = reinterpret_cast<uint32_t*>(a.lse2) + (size_t)a.trace_off + (warp == 2 ? 0 : warp == 6 ? 4096 : 12288);
#endif
        if (half == 1) asm volatile("bar.arrive 2, 256;" ::: "memory");      // the first turn goes to half 0
        int it = 0;
        for (int w = blockIdx.x; w < a.n_items; w += gridDim.x, ++it) {
            const int qt = w % a.n_qt, bh = w / a.n_qt, head = bh % a.heads, img = bh / a.heads;
            const int g0 = it * n_kt, grow = qt * AT_BM + row;
            float* stat_m = reinterpret_cast<float*>(smem + FwdSmem::STATS) + (it & 1) * 512;
            float* stat_l = stat_m + 256;
            m_used = -INFINITY; l0 = l1 = l2 = l3 = 0.f;
            mbar_wait(&s_full[g0 % AT_SBUF], (g0 / AT_SBUF) & 1);
            tc_fence_after();
            tmem_ld32_issue(tlane + (g0 % AT_SBUF) * AT_BN, r0);
            tmem_ld32_issue(tlane + (g0 % AT_SBUF) * AT_BN + 32, r1);
            for (int j = 0; j < n_kt; ++j) {
                const int g = g0 + j, b = g % AT_SBUF, pb = g & 1;
                const int nv = a.L - j * AT_BN - half * 64;      // this warp's columns of the tile that exist (may be <= 0)
                uint8_t* blk = smem + FwdSmem::P + pb * 2 * AT_TILE_BYTES + half * AT_TILE_BYTES + row * 128;
                AT_TR(0);
                tmem_ld_wait();                                  // both halves were requested during the previous tile's exp work
                AT_TR(1);
                // the two barrier round trips this tile needs before its exp phase (~200 clk each even when already complete) are
                // put in flight here, under the running-maximum work
                const bool more = j + 1 < n_kt;
                const int bn = (g + 1) % AT_SBUF;
                const uint32_t ok_p = mbar_try(&p_empty[pb], ((g >> 1) & 1) ^ 1);
                const uint32_t ok_s = more ? mbar_try(&s_full[bn], ((g + 1) / AT_SBUF) & 1) : 1u;
                const bool full = nv >= 64;                      // every tile but the last: no column masking (compile-time paths)
                float mx = full ? max_chunk(r0, 32) : nv > 0 ? max_chunk(r0, nv) : -INFINITY;
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_empty[b]);         // S[b] is in registers: the MMA warp may overwrite it
                if (full) mx = fmaxf(mx, max_chunk(r1, 32));
                else if (nv > 32) mx = fmaxf(mx, max_chunk(r1, nv - 32));
                AT_TR(2);
                const float m_new = mx * AT_LOG2E;
                const bool need = m_new > m_used + RESCALE_THR;   // first tile: m_used = -inf
                if (__any_sync(0xffffffffu, need)) {
                    if (j > 0) {
                        // rescale this warp's rows of its accumulator and its row sums: every P V product issued so far must
                        // have finished (p_empty of tile g-1), and P V of tile g cannot start before this warp arrives on p_full
                        const float f = need ? ex2_approx(m_used - m_new) : 1.f;
                        mbar_wait(&p_empty[(g - 1) & 1], ((g - 1) >> 1) & 1);
                        tc_fence_after();
                        uint32_t t[32];
#pragma unroll
                        for (int c = 0; c < HD / 32; ++c) {
                            tmem_ld32_issue(to + c * 32, t);
                            tmem_ld_wait();
#pragma unroll
                            for (int e = 0; e < 32; ++e) t[e] = __float_as_uint(__uint_as_float(t[e]) * f);
                            tmem_st32(to + c * 32, t);
                        }
                        tmem_st_wait();
                        l0 *= f; l1 *= f; l2 *= f; l3 *= f;
                    }
                    if (need) m_used = m_new;
                }
                // both waits come before the exp work; the next tile's TMEM loads are issued as soon as r0 / r1 have been consumed
                AT_TR(3);
                if (!ok_p) mbar_wait(&p_empty[pb], ((g >> 1) & 1) ^ 1);
                if (!ok_s) mbar_wait(&s_full[bn], ((g + 1) / AT_SBUF) & 1);
                tc_fence_after();
                AT_TR(4);
                asm volatile("bar.sync %0, 256;" ::"r"(2 + half) : "memory");
                AT_TR(5);
                if (full) {
                    prob_chunk(r0, 32, blk, 0);
                    if (more) tmem_ld32_issue(tlane + bn * AT_BN, r0);
                    prob_chunk(r1, 32, blk, 4);
                } else {
                    prob_chunk(r0, nv, blk, 0);
                    if (more) tmem_ld32_issue(tlane + bn * AT_BN, r0);
                    prob_chunk(r1, nv - 32, blk, 4);
                }
                asm volatile("bar.arrive %0, 256;" ::"r"(2 + (half ^ 1)) : "memory");
                AT_TR(6);
                if (more) tmem_ld32_issue(tlane + bn * AT_BN + 32, r1);
                fence_async_smem();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[pb]);
                AT_TR(7);
            }
            // ---- merge the two key halves of each row (named barrier over the 256 softmax threads; the statistics are
            //      double-buffered by item parity, so the next item's writes cannot overtake this item's reads)
            stat_m[half * 128 + row] = m_used;
            stat_l[half * 128 + row] = (l0 + l1) + (l2 + l3);
            asm volatile("bar.sync 1, 256;" ::: "memory");
            const float mA = stat_m[row], mB = stat_m[128 + row];
            const float m = fmaxf(mA, mB);                       // key 0 always exists: finite
            const float cA = ex2_approx(mA - m), cB = ex2_approx(mB - m);
            const float l = stat_l[row] * cA + stat_l[128 + row] * cB;
            const float inv_l = 1.f / l;
            if (half == 0 && a.lse2 && grow < a.L) a.lse2[(size_t)bh * a.L + grow] = m + __log2f(l);
            // ---- epilogue: (cA O_A + cB O_B) / l -> fp16, head-interleaved [N*L, heads*HD]; HD = 64: this warp's 32 of the 64
            //      dims, HD = 32: the half-0 warps store the whole row.  The accumulators are handed back to the MMA warp as soon
            //      as they are in registers.
            mbar_wait(o_full, it & 1);
            tc_fence_after();
            const bool storer = (HD == 64 || half == 0);
            const int doff = (HD == 64) ? half * 32 : 0;
            if (storer) {
                tmem_ld32_issue(tmem_o + lanes + doff, r0);
                tmem_ld32_issue(tmem_o + HD + lanes + doff, r1);
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(o_empty);
            if (storer && grow < a.L) {
                h16* orow = a.o + ((size_t)img * a.L + grow) * (size_t)(a.heads * HD) + head * HD + doff;
                const float fa = cA * inv_l, fb = cB * inv_l;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    uint32_t wv[4];
#pragma unroll
                    for (int e4 = 0; e4 < 4; ++e4) {
                        const int e = 8 * t + 2 * e4;
                        const h16x2 h = __floats2half2_rn(fmaf(__uint_as_float(r0[e]), fa, __uint_as_float(r1[e]) * fb),
                                                          fmaf(__uint_as_float(r0[e + 1]), fa, __uint_as_float(r1[e + 1]) * fb));
                        wv[e4] = *reinterpret_cast<const uint32_t*>(&h);
                    }
                    *reinterpret_cast<uint4*>(orow + t * 8) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
                }
            }
        }
        if (half == 0) asm volatile("bar.sync 2, 256;" ::: "memory");        // consume half 1's last hand-over
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------------
struct MeanSmem {
    static constexpr int RING = 0;                                           // stage s: Q_h at +s*32K, K_h at +s*32K+16K
    static constexpr int STG = AT_STAGES * 2 * AT_TILE_BYTES;                // 8 warps x 4 KB transposition blocks
    static constexpr int BARS = STG + AT_SM_WARPS * 32 * 32 * 4;
    static constexpr int TOTAL = BARS + 256 + 1024;
};
struct MeanArgs {
    int L, heads;
    const float* lse2;
    float* avg;
    double* sums;      // optional [n_img]: += sum of avg[img][1:, 1:] (the RFM layer score, clip_tool.py:158-160)
};

__global__ void __launch_bounds__(AT_THREADS, 1) attn_headmean_kernel(const __grid_constant__ CUtensorMap mapQ,
                                                                       const __grid_constant__ CUtensorMap mapK,
                                                                       const __grid_constant__ MeanArgs a) {
    extern __shared__ uint8_t at_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + MeanSmem::BARS);
    uint64_t* full = bars;                         // [STAGES]
    uint64_t* empty = full + AT_STAGES;            // [STAGES]
    uint64_t* s_full = empty + AT_STAGES;          // [2]
    uint64_t* s_empty = s_full + 2;                // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kt = blockIdx.x, qt = blockIdx.y, img = blockIdx.z;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&mapQ); prefetch_tmap(&mapK);
        for (int s = 0; s < AT_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&s_full[b], 1); mbar_init(&s_empty[b], AT_SM_WARPS); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int h = 0; h < a.heads; ++h) {
                const int s = h % AT_STAGES;
                mbar_wait(&empty[s], ((h / AT_STAGES) & 1) ^ 1);
                uint8_t* sq = smem + MeanSmem::RING + s * 2 * AT_TILE_BYTES;
                mbar_expect_tx(&full[s], 2 * AT_TILE_BYTES);
                tma_load_4d(sq, &mapQ, &full[s], 0, qt * AT_BM, img * a.heads + h, 0);
                tma_load_4d(sq + AT_TILE_BYTES, &mapK, &full[s], 0, kt * AT_BN, img * a.heads + h, 0);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = make_idesc(AT_BM, AT_BN);
            for (int h = 0; h < a.heads; ++h) {
                const int s = h % AT_STAGES, b = h & 1;
                {   // both barrier round trips in flight together
                    const uint32_t ok_f = mbar_try(&full[s], (h / AT_STAGES) & 1);
                    const uint32_t ok_e = mbar_try(&s_empty[b], ((h >> 1) & 1) ^ 1);
                    if (!ok_f) mbar_wait(&full[s], (h / AT_STAGES) & 1);
                    if (!ok_e) mbar_wait(&s_empty[b], ((h >> 1) & 1) ^ 1);
                }
                tc_fence_after();
                const uint32_t sq = smem_u32(smem + MeanSmem::RING + s * 2 * AT_TILE_BYTES);
                const uint64_t dq = make_smem_desc(sq), dk = make_smem_desc(sq + AT_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < AT_HD / 16; ++k) umma_f16(tmem_base + b * AT_BN, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                umma_commit(&s_full[b]);
                umma_commit(&empty[s]);
            }
        }
    } else {
        const int q = warp & 3;                              // TMEM lane quarter
        const int half = (warp - 2) >> 2;                    // which 64 key columns this warp accumulates
        const int row = q * 32 + lane, grow = qt * AT_BM + row;
        const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16) + half * 64;
        float* stg = reinterpret_cast<float*>(smem + MeanSmem::STG) + (warp - 2) * 32 * 32;
        float acc[64];
#pragma unroll
        for (int i = 0; i < 64; ++i) acc[i] = 0.f;
        uint32_t r0[32], r1[32];
        float lse = (grow < a.L) ? __ldg(a.lse2 + (size_t)(img * a.heads) * a.L + grow) : 0.f;
        mbar_wait(&s_full[0], 0);
        tc_fence_after();
        tmem_ld32_issue(tlane, r0);
        for (int h = 0; h < a.heads; ++h) {
            const int b = h & 1;
            const float lse_next = (grow < a.L && h + 1 < a.heads) ? __ldg(a.lse2 + (size_t)(img * a.heads + h + 1) * a.L + grow) : 0.f;
            // the TMEM load of the second 32 columns / of the next head's first 32 runs under the exp2 work
            tmem_ld_wait();
            tmem_ld32_issue(tlane + b * AT_BN + 32, r1);
            // the next head's score barrier is probed here, so its round trip runs under the first 32 exponentials
            const uint32_t ok_s = (h + 1 < a.heads) ? mbar_try(&s_full[b ^ 1], ((h + 1) >> 1) & 1) : 1u;
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[i] += ex2_approx(fmaf(__uint_as_float(r0[i]), AT_LOG2E, -lse));
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[b]);         // S[b] is in registers: the next head's MMA may overwrite it
            if (h + 1 < a.heads) {
                if (!ok_s) mbar_wait(&s_full[b ^ 1], ((h + 1) >> 1) & 1);
                tc_fence_after();
                tmem_ld32_issue(tlane + (b ^ 1) * AT_BN, r0);
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[32 + i] += ex2_approx(fmaf(__uint_as_float(r1[i]), AT_LOG2E, -lse));
            lse = lse_next;
        }
        tmem_ld_wait();
        // mean over heads, transposed through the warp's swizzled 32x32 block so that each store is one row x 32 keys
        const float inv_h = 1.f / (float)a.heads;
        const int row0 = qt * AT_BM + q * 32;
        if (a.sums) {
            // sum of this warp's part of avg[1:, 1:] (class-token row and column excluded), one fp64 atomic per warp: the
            // RFM layer selection needs it and would otherwise re-read every map from HBM
            float part = 0.f;
            if (grow >= 1 && grow < a.L) {
                const int key0 = kt * AT_BN + half * 64;
#pragma unroll
                for (int i = 0; i < 64; ++i) part += (key0 + i >= 1 && key0 + i < a.L) ? acc[i] : 0.f;
            }
            part = warp_sum(part);
            if (lane == 0 && part != 0.f) atomicAdd(a.sums + img, (double)(part * inv_h));
        }
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int key = kt * AT_BN + half * 64 + c * 32 + lane;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) =
                    make_float4(acc[c * 32 + 4 * j] * inv_h, acc[c * 32 + 4 * j + 1] * inv_h, acc[c * 32 + 4 * j + 2] * inv_h,
                                acc[c * 32 + 4 * j + 3] * inv_h);
            __syncwarp();
            if (key < a.L) {
                float* dst = a.avg + ((size_t)img * a.L + row0) * a.L + key;
#pragma unroll 4
                for (int rr = 0; rr < 32; ++rr) {
                    if (row0 + rr >= a.L) break;
                    dst[(size_t)rr * a.L] = stg[rr * 32 + (((lane >> 2) ^ (rr & 7)) << 2) + (lane & 3)];
                }
            }
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 256);
    }
}

// ---------------------------------------------------------------------------------------------------
// attn_headmean64_kernel — the same map from 128 x 64 tiles with TWO CTAs resident per SM.  The 128 x 128 kernel above keeps one
// CTA per SM, so its per-tile prologue (barrier init, TMEM allocation, first TMA round trip) and epilogue (transposed store of the
// 64 KB tile) run with the MUFU idle: 107 of its 286 us at L = 1601 do not depend on the head count (tools/hm_slope.py).  Here a
// CTA holds half a tile (thread = row x 32 keys: 32 accumulators, < 102 registers), so two CTAs share an SM and one's
// prologue / epilogue runs under the other's head loop.  Q is loaded once per 64 keys instead of once per 128 (L2 traffic x1.5).
// ---------------------------------------------------------------------------------------------------
constexpr int HM_BN = 64, HM_STAGES = 3;
constexpr int HM_STAGE_BYTES = AT_TILE_BYTES + AT_TILE_BYTES / 2;            // Q_h [128][64] + K_h [64][64] fp16
struct Mean64Smem {
    static constexpr int RING = 0;
    static constexpr int STG = HM_STAGES * HM_STAGE_BYTES;                   // 8 warps x 4 KB transposition blocks
    static constexpr int BARS = STG + AT_SM_WARPS * 32 * 32 * 4;
    static constexpr int TOTAL = BARS + 256 + 1024;
};

__global__ void __launch_bounds__(AT_THREADS, 2) attn_headmean64_kernel(const __grid_constant__ CUtensorMap mapQ,
                                                                         const __grid_constant__ CUtensorMap mapK,
                                                                         const __grid_constant__ MeanArgs a) {
    extern __shared__ uint8_t at_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Mean64Smem::BARS);
    uint64_t* full = bars;                         // [STAGES]
    uint64_t* empty = full + HM_STAGES;            // [STAGES]
    uint64_t* s_full = empty + HM_STAGES;          // [2]
    uint64_t* s_empty = s_full + 2;                // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kt = blockIdx.x, qt = blockIdx.y, img = blockIdx.z;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&mapQ); prefetch_tmap(&mapK);
        for (int s = 0; s < HM_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&s_full[b], 1); mbar_init(&s_empty[b], AT_SM_WARPS); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 2 * HM_BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int h = 0; h < a.heads; ++h) {
                const int s = h % HM_STAGES;
                mbar_wait(&empty[s], ((h / HM_STAGES) & 1) ^ 1);
                uint8_t* sq = smem + Mean64Smem::RING + s * HM_STAGE_BYTES;
                mbar_expect_tx(&full[s], HM_STAGE_BYTES);
                tma_load_4d(sq, &mapQ, &full[s], 0, qt * AT_BM, img * a.heads + h, 0);
                tma_load_4d(sq + AT_TILE_BYTES, &mapK, &full[s], 0, kt * HM_BN, img * a.heads + h, 0);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = make_idesc(AT_BM, HM_BN);
            for (int h = 0; h < a.heads; ++h) {
                const int s = h % HM_STAGES, b = h & 1;
                {   // both barrier round trips in flight together
                    const uint32_t ok_f = mbar_try(&full[s], (h / HM_STAGES) & 1);
                    const uint32_t ok_e = mbar_try(&s_empty[b], ((h >> 1) & 1) ^ 1);
                    if (!ok_f) mbar_wait(&full[s], (h / HM_STAGES) & 1);
                    if (!ok_e) mbar_wait(&s_empty[b], ((h >> 1) & 1) ^ 1);
                }
                tc_fence_after();
                const uint32_t sq = smem_u32(smem + Mean64Smem::RING + s * HM_STAGE_BYTES);
                const uint64_t dq = make_smem_desc(sq), dk = make_smem_desc(sq + AT_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < AT_HD / 16; ++k) umma_f16(tmem_base + b * HM_BN, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                umma_commit(&s_full[b]);
                umma_commit(&empty[s]);
            }
        }
    } else {
        const int q = warp & 3;                              // TMEM lane quarter
        const int half = (warp - 2) >> 2;                    // which 32 key columns this warp accumulates
        const int row = q * 32 + lane, grow = qt * AT_BM + row;
        const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16) + half * 32;
        float* stg = reinterpret_cast<float*>(smem + Mean64Smem::STG) + (warp - 2) * 32 * 32;
        float acc[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[i] = 0.f;
        uint32_t r0[32];
        float lse = (grow < a.L) ? __ldg(a.lse2 + (size_t)(img * a.heads) * a.L + grow) : 0.f;
        for (int h = 0; h < a.heads; ++h) {
            const int b = h & 1;
            const float lse_next = (grow < a.L && h + 1 < a.heads) ? __ldg(a.lse2 + (size_t)(img * a.heads + h + 1) * a.L + grow) : 0.f;
            mbar_wait(&s_full[b], (h >> 1) & 1);
            tc_fence_after();
            tmem_ld32_issue(tlane + b * HM_BN, r0);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[b]);         // S[b] is in registers: the MMA of head h + 2 may overwrite it
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[i] += ex2_approx(fmaf(__uint_as_float(r0[i]), AT_LOG2E, -lse));
            lse = lse_next;
        }
        const float inv_h = 1.f / (float)a.heads;
        const int row0 = qt * AT_BM + q * 32;
        const int key0 = kt * HM_BN + half * 32;
        if (a.sums) {
            float part = 0.f;
            if (grow >= 1 && grow < a.L) {
#pragma unroll
                for (int i = 0; i < 32; ++i) part += (key0 + i >= 1 && key0 + i < a.L) ? acc[i] : 0.f;
            }
            part = warp_sum(part);
            if (lane == 0 && part != 0.f) atomicAdd(a.sums + img, (double)(part * inv_h));
        }
        // mean over heads, transposed through the warp's swizzled 32x32 block so that each store is one row x 32 keys
        const int key = key0 + lane;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(stg + lane * 32 + ((j ^ (lane & 7)) << 2)) =
                make_float4(acc[4 * j] * inv_h, acc[4 * j + 1] * inv_h, acc[4 * j + 2] * inv_h, acc[4 * j + 3] * inv_h);
        __syncwarp();
        if (key < a.L) {
            float* dst = a.avg + ((size_t)img * a.L + row0) * a.L + key;
#pragma unroll 4
            for (int rr = 0; rr < 32; ++rr) {
                if (row0 + rr >= a.L) break;
                dst[(size_t)rr * a.L] = stg[rr * 32 + (((lane >> 2) ^ (rr & 7)) << 2) + (lane & 3)];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * HM_BN);
    }
}

// ---------------------------------------------------------------------------------------------------
// attn_bwd_colsum_kernel — the only L x L work of the Grad-CAM backward (gradcam.cu header):
//     colsum[j] = sum_{i>=1} dS_ij,   dS = P * (dP - delta),  dP = dO V^T,  delta_i = dO_i . O_i
//               = V'_j . W'_j         with  W' = P~^T dO'   (P~ = P without the class-token query row,
//                                            dO' = [dO | -delta_hi | -delta_lo | 0...],  V' = [V | 1 | 1 | 0...])
// so neither P nor dP is ever materialised.  One CTA per (128-key tile, head, instance) walks the query tiles ONCE:
//   S^T = K Q^T (tcgen05, keys on the TMEM lanes)  ->  P^T = exp2(S^T log2e - lse2_i) (the exact probabilities, from the forward's
//   row log-sum-exps; masked queries carry lse = +inf)  ->  fp16 tile in shared memory  ->  W' += P^T dO' (tcgen05, N = 80)
// and the epilogue contracts W'_j with V'_j.  Same warp roles and barriers as attn_fwd_kernel's second pass.
constexpr int BW_ND = 80;                          // 64 head dims + delta hi / lo + padding to a legal MMA N
constexpr int BW_QST = 4;                          // Q ring (released by the score MMA, which runs two tiles ahead)
constexpr int BW_DST = 3;                          // dO'^T ring (released by the P^T dO' MMA)
constexpr int BW_DBLK = BW_ND * 128;               // one K-major [80][64 queries] block of dO'^T (10 KB)
constexpr int BW_MAX_L = 2048;
struct BwdSmem {
    static constexpr int KT = 0;
    static constexpr int QRING = AT_TILE_BYTES;
    static constexpr int DRING = QRING + BW_QST * AT_TILE_BYTES;
    static constexpr int P = DRING + BW_DST * 2 * BW_DBLK;
    static constexpr int LSE = P + 2 * 2 * AT_TILE_BYTES;
    static constexpr int BARS = LSE + BW_MAX_L * 4;
    static constexpr int TOTAL = BARS + 256 + 1024;
};
struct BwdArgs {
    int L, heads, n_qt;
    const float* lse2;       // [N*heads, L]
    const h16* v;            // [N*heads, L, 64]
    float* colsum;           // [n_inst*heads, L]
    BwdInstTable tab;
};

__global__ void __launch_bounds__(AT_THREADS, 1) attn_bwd_colsum_kernel(const __grid_constant__ CUtensorMap mapK,
                                                                         const __grid_constant__ CUtensorMap mapQ,
                                                                         const __grid_constant__ CUtensorMap mapD,
                                                                         const __grid_constant__ BwdArgs a) {
    extern __shared__ uint8_t at_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(at_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BwdSmem::BARS);
    uint64_t* k_full = bars;                       // [1]
    uint64_t* q_full = bars + 1;                   // [BW_QST]
    uint64_t* q_empty = q_full + BW_QST;           // [BW_QST]
    uint64_t* d_full = q_empty + BW_QST;           // [BW_DST]
    uint64_t* d_empty = d_full + BW_DST;           // [BW_DST]
    uint64_t* s_full = d_empty + BW_DST;           // [3]
    uint64_t* s_empty = s_full + AT_SBUF;          // [3]
    uint64_t* p_full = s_empty + AT_SBUF;          // [2]
    uint64_t* p_empty = p_full + 2;                // [2]
    uint64_t* w_full = p_empty + 2;                // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);
    float* lse_s = reinterpret_cast<float*>(smem + BwdSmem::LSE);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int jt = blockIdx.x, head = blockIdx.y, inst = blockIdx.z;
    const int img = a.tab.img[inst];
    const int bh = img * a.heads + head, ih = inst * a.heads + head;
    const int n_qt = a.n_qt;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&mapK); prefetch_tmap(&mapQ); prefetch_tmap(&mapD);
        mbar_init(k_full, 1);
        for (int s = 0; s < BW_QST; ++s) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
        for (int s = 0; s < BW_DST; ++s) { mbar_init(&d_full[s], 1); mbar_init(&d_empty[s], 1); }
        for (int b = 0; b < AT_SBUF; ++b) { mbar_init(&s_full[b], 1); mbar_init(&s_empty[b], AT_SM_WARPS); }
        for (int b = 0; b < 2; ++b) { mbar_init(&p_full[b], AT_SM_WARPS); mbar_init(&p_empty[b], 1); }
        mbar_init(w_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    // row log-sum-exps of this (image, head); the class-token query (i = 0, excluded from the Grad-CAM mean) and the padding
    // queries get +inf so that their probabilities come out as exactly 0
    for (int i = threadIdx.x; i < n_qt * AT_BN; i += AT_THREADS)
        lse_s[i] = (i >= 1 && i < a.L) ? __ldg(a.lse2 + (size_t)bh * a.L + i) : INFINITY;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_w = tmem_base + AT_SBUF * AT_BN;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(k_full, AT_TILE_BYTES);
            tma_load_4d(smem + BwdSmem::KT, &mapK, k_full, 0, jt * AT_BM, bh, 0);
            // separate rings for the two operands of a query tile: both are requested three to four tiles before their use
            auto load_d = [&](int it) {
                const int s = it % BW_DST;
                mbar_wait(&d_empty[s], ((it / BW_DST) & 1) ^ 1);
                uint8_t* sd = smem + BwdSmem::DRING + s * 2 * BW_DBLK;
                mbar_expect_tx(&d_full[s], 2 * BW_DBLK);
                tma_load_4d(sd, &mapD, &d_full[s], it * AT_BN, 0, ih, 0);
                tma_load_4d(sd + BW_DBLK, &mapD, &d_full[s], it * AT_BN + 64, 0, ih, 0);
            };
            for (int it = 0; it < n_qt; ++it) {
                const int s = it % BW_QST;
                mbar_wait(&q_empty[s], ((it / BW_QST) & 1) ^ 1);
                mbar_expect_tx(&q_full[s], AT_TILE_BYTES);
                tma_load_4d(smem + BwdSmem::QRING + s * AT_TILE_BYTES, &mapQ, &q_full[s], 0, it * AT_BN, bh, 0);
                if (it >= 2) load_d(it - 2);
            }
            for (int it = (n_qt >= 2 ? n_qt - 2 : 0); it < n_qt; ++it) load_d(it);
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = make_idesc(AT_BM, AT_BN), idesc_w = make_idesc(AT_BM, BW_ND);
            const uint64_t dk = make_smem_desc(smem_u32(smem + BwdSmem::KT));
            mbar_wait(k_full, 0);
            uint32_t sc = 0;
            auto issue_s = [&](int it_) {          // S^T[sc % 3] = K Q^T for ring slot it_
                const int s = it_ % BW_QST, b = sc % AT_SBUF;
                mbar_wait(&q_full[s], (it_ / BW_QST) & 1);
                mbar_wait(&s_empty[b], ((sc / AT_SBUF) & 1) ^ 1);
                tc_fence_after();
                const uint64_t dq = make_smem_desc(smem_u32(smem + BwdSmem::QRING + s * AT_TILE_BYTES));
#pragma unroll
                for (int k = 0; k < AT_HD / 16; ++k) umma_f16(tmem_base + b * AT_BN, dk + 2 * k, dq + 2 * k, idesc_s, k != 0);
                umma_commit(&s_full[b]);
                umma_commit(&q_empty[s]);
                ++sc;
            };
            issue_s(0);                             // scores run two tiles ahead of the P^T dO' product
            if (n_qt > 1) issue_s(1);
            for (int it = 0; it < n_qt; ++it) {
                if (it + 2 < n_qt) issue_s(it + 2);
                const int s = it % BW_DST, pb = it & 1;
                mbar_wait(&d_full[s], (it / BW_DST) & 1);
                mbar_wait(&p_full[pb], (it >> 1) & 1);
                tc_fence_after();
                const uint32_t sp = smem_u32(smem + BwdSmem::P + pb * 2 * AT_TILE_BYTES);
                const uint32_t sd = smem_u32(smem + BwdSmem::DRING + s * 2 * BW_DBLK);
#pragma unroll
                for (int kb = 0; kb < 2; ++kb) {
                    const uint64_t dp = make_smem_desc(sp + kb * AT_TILE_BYTES), dd = make_smem_desc(sd + kb * BW_DBLK);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_f16(tmem_w, dp + 2 * k, dd + 2 * k, idesc_w, (it | kb | k) != 0);
                }
                umma_commit(&p_empty[pb]);
                umma_commit(&d_empty[s]);
            }
            umma_commit(w_full);
        }
    } else {
        const int q = warp & 3;                              // TMEM lane quarter
        const int half = (warp - 2) >> 2;                    // which 64 query columns of every tile this warp owns
        const int row = q * 32 + lane, gkey = jt * AT_BM + row;
        const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16) + half * 64;
        uint32_t r0[32], r1[32];
        auto prob_chunk = [&](const uint32_t (&r)[32], const float* lse, uint8_t* blk, int cc0) {
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 ls = *reinterpret_cast<const float4*>(lse + 4 * i);      // broadcast: same queries for every lane
                const float p0 = ex2_approx(fmaf(__uint_as_float(r[4 * i]), AT_LOG2E, -ls.x));
                const float p1 = ex2_approx(fmaf(__uint_as_float(r[4 * i + 1]), AT_LOG2E, -ls.y));
                const float p2 = ex2_approx(fmaf(__uint_as_float(r[4 * i + 2]), AT_LOG2E, -ls.z));
                const float p3 = ex2_approx(fmaf(__uint_as_float(r[4 * i + 3]), AT_LOG2E, -ls.w));
                const h16x2 h0 = __floats2half2_rn(p0, p1), h1 = __floats2half2_rn(p2, p3);
                pk[2 * i] = *reinterpret_cast<const uint32_t*>(&h0);
                pk[2 * i + 1] = *reinterpret_cast<const uint32_t*>(&h1);
            }
#pragma unroll
            for (int t = 0; t < 4; ++t)
                *reinterpret_cast<uint4*>(blk + (((cc0 + t) ^ (row & 7)) << 4)) = make_uint4(pk[4 * t], pk[4 * t + 1], pk[4 * t + 2], pk[4 * t + 3]);
        };
        mbar_wait(&s_full[0], 0);
        tc_fence_after();
        tmem_ld32_issue(tlane, r0);
        tmem_ld32_issue(tlane + 32, r1);
        for (int it = 0; it < n_qt; ++it) {
            const int b = it % AT_SBUF, pb = it & 1;
            const float* lse = lse_s + it * AT_BN + half * 64;
            uint8_t* blk = smem + BwdSmem::P + pb * 2 * AT_TILE_BYTES + half * AT_TILE_BYTES + row * 128;
            tmem_ld_wait();                                  // both halves were requested during the previous tile's exp work
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[b]);
            const bool more = it + 1 < n_qt;
            const int bn = (it + 1) % AT_SBUF;
            mbar_wait(&p_empty[pb], ((it >> 1) & 1) ^ 1);
            if (more) mbar_wait(&s_full[bn], ((it + 1) / AT_SBUF) & 1);
            tc_fence_after();
            prob_chunk(r0, lse, blk, 0);
            if (more) tmem_ld32_issue(tlane + bn * AT_BN, r0);
            prob_chunk(r1, lse + 32, blk, 4);
            if (more) tmem_ld32_issue(tlane + bn * AT_BN + 32, r1);
            fence_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[pb]);
        }
        // ---- epilogue: colsum[key] = V'_key . W'_key (the half-0 warps hold the whole row: 3 x 32 TMEM columns)
        mbar_wait(w_full, 0);
        tc_fence_after();
        if (half == 0) {
            const uint32_t wl = tmem_w + ((uint32_t)(q * 32) << 16);
            const int key = gkey < a.L ? gkey : a.L - 1;            // clamped read; out-of-range rows are not stored
            const uint4* vr = reinterpret_cast<const uint4*>(a.v + ((size_t)bh * a.L + key) * AT_HD);
            float acc = 0.f;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                tmem_ld32_issue(wl + c * 32, r0);
                tmem_ld_wait();
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const uint4 pk = __ldg(vr + c * 4 + t);
                    const uint32_t w4[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 vf = __half22float2(*reinterpret_cast<const h16x2*>(&w4[e]));
                        acc = fmaf(vf.x, __uint_as_float(r0[8 * t + 2 * e]), acc);
                        acc = fmaf(vf.y, __uint_as_float(r0[8 * t + 2 * e + 1]), acc);
                    }
                }
            }
            tmem_ld32_issue(wl + 64, r0);
            tmem_ld_wait();
            acc += __uint_as_float(r0[0]) + __uint_as_float(r0[1]);
            if (gkey < a.L) a.colsum[(size_t)ih * a.L + gkey] = acc;
        }
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------------
static int attn_num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
    }
    return n;
}

static int attn_maps(const h16* q, const h16* k, const h16* vt, int nh, int L, int Lp, int hd, CUtensorMap* mq, CUtensorMap* mk,
                     CUtensorMap* mv) {
    int u0, u1, rc;
    GemmOperand oq{q, L, hd, (long)L * hd, 0}, ok{k, L, hd, (long)L * hd, 0};
    if ((rc = make_tma_map(mq, oq, hd, nh, 1, AT_BM, &u0, &u1))) return rc;
    if ((rc = make_tma_map(mk, ok, hd, nh, 1, AT_BN, &u0, &u1))) return rc;
    if (vt && mv) {
        if (hd == 64) {
            // V itself, [nh, L, 64]: same geometry as the K map (rows beyond L read as zero)
            GemmOperand ov{vt, L, hd, (long)L * hd, 0};
            if ((rc = make_tma_map(mv, ov, hd, nh, 1, AT_BN, &u0, &u1))) return rc;
        } else {
            // V^T rows = head dims, K extent = L keys (NOT Lp): columns L..Lp-1 are never written and must read as zero
            GemmOperand ov{vt, hd, Lp, (long)hd * Lp, 0};
            if ((rc = make_tma_map(mv, ov, L, nh, 1, hd, &u0, &u1))) return rc;
        }
    }
    return WC_OK;
}

template <int HD>
static int launch_fwd_hd(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const AttnArgs& a, dim3 grid,
                         cudaStream_t st) {
    WC_SET_MAX_SMEM(attn_fwd_kernel<HD>, FwdSmem::TOTAL);
    attn_fwd_kernel<HD><<<grid, FW_THREADS, FwdSmem::TOTAL, st>>>(mq, mk, mv, a);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

// second-generation head_dim-64 forward (attention2.cu)
int launch_attention_fwd2(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, h16* o, float* lse2, int n_img,
                          int heads, int L, int sms, cudaStream_t st);
// third variant (attention3.cu): 64-key tiles, two CTAs per SM (WC_ATTN2=2)
int launch_attention_fwd3(const CUtensorMap& mq, const h16* k, const h16* v, h16* o, float* lse2, int n_img, int heads, int L,
                          cudaStream_t st);
// head_dim 64: attn_fwd_kernel<64> unless WC_ATTN2=1 (read per launch, so that the tests can run both): the second kernel is within
// 2% of the first at the encoder's shapes (profiles/r02_attention.txt) and owns two query tiles per item, which is the coarser
// granularity at the evaluation driver's batch of two.

int launch_attention_fwd(const h16* q, const h16* k, const h16* vt, h16* o, float* lse2, int n_img, int heads, int L, int Lp,
                         int hd, cudaStream_t st) {
    WC_CHECK_ARG(q && k && vt && o && n_img > 0 && heads > 0 && L > 0 && Lp >= L && (Lp & 7) == 0);
    WC_CHECK_ARG(heads <= 65535 && n_img <= 65535);
    if (hd != 64 && hd != 32) return fail(WC_ERR_UNSUPPORTED, "%s: head_dim %d (64 and 32 are built)", __func__, hd);
    CUtensorMap mq, mk, mv;
    int rc = attn_maps(q, k, vt, n_img * heads, L, Lp, hd, &mq, &mk, &mv);
    if (rc) return rc;
    // a batch of 1 collapses the map's batch dimension; the kernel's batch coordinate is then 0 anyway
    const int n_qt = ceil_div(L, AT_BM);
    const long n_items = (long)n_qt * heads * n_img;
    WC_CHECK_ARG(n_items < (1L << 30));
    AttnArgs a{L, heads, ceil_div(L, AT_BN), n_qt, (int)n_items, o, lse2, (long)n_img * heads * L};
    const int sms = attn_num_sms();
    dim3 grid((unsigned)(n_items < sms ? n_items : sms));
    ProfScope prof(PK_ATTENTION, 4.0 * L * L * hd * heads * n_img, st);
    if (g_prof_on) prof_tag(L, n_img, heads, hd);
    const char* sel = hd == 64 ? getenv("WC_ATTN2") : nullptr;
    if (sel && sel[0] == '2') return launch_attention_fwd3(mq, k, vt, o, lse2, n_img, heads, L, st);
    if (sel && sel[0] == '1') return launch_attention_fwd2(mq, mk, mv, o, lse2, n_img, heads, L, sms, st);
    return hd == 64 ? launch_fwd_hd<64>(mq, mk, mv, a, grid, st) : launch_fwd_hd<32>(mq, mk, mv, a, grid, st);
}

int launch_attn_headmean(const h16* q, const h16* k, const float* lse2, float* avg, int n_img, int heads, int L,
                         cudaStream_t st, double* sums) {
    WC_CHECK_ARG(q && k && lse2 && avg && n_img > 0 && heads > 0 && L > 0 && n_img <= 65535);
    WC_SET_MAX_SMEM(attn_headmean_kernel, MeanSmem::TOTAL);
    CUtensorMap mq, mk;
    int rc = attn_maps(q, k, nullptr, n_img * heads, L, 0, AT_HD, &mq, &mk, nullptr);
    if (rc) return rc;
    if (sums) WC_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * n_img, st));
    MeanArgs a{L, heads, lse2, avg, sums};
    ProfScope prof(PK_SOFTMAX, 4.0 * n_img * (double)L * L, st);
    if (g_prof_on) prof_tag(L, n_img, heads, 1);
    const char* hm = getenv("WC_HEADMEAN64");            // read per launch, so that the tests can run both kernels
    if (!(hm && hm[0] == '0')) {        // 128 x 64 tiles, two CTAs per SM (WC_HEADMEAN64=0: the 128 x 128 kernel)
        WC_SET_MAX_SMEM(attn_headmean64_kernel, Mean64Smem::TOTAL);
        CUtensorMap mk64;
        int u0, u1;
        GemmOperand ok{k, L, AT_HD, (long)L * AT_HD, 0};
        if ((rc = make_tma_map(&mk64, ok, AT_HD, n_img * heads, 1, HM_BN, &u0, &u1))) return rc;
        dim3 grid64(ceil_div(L, HM_BN), ceil_div(L, AT_BM), n_img);
        attn_headmean64_kernel<<<grid64, AT_THREADS, Mean64Smem::TOTAL, st>>>(mq, mk64, a);
        WC_LAUNCH_CHECK();
        return WC_OK;
    }
    dim3 grid(ceil_div(L, AT_BN), ceil_div(L, AT_BM), n_img);
    attn_headmean_kernel<<<grid, AT_THREADS, MeanSmem::TOTAL, st>>>(mq, mk, a);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

int launch_attn_bwd_colsum(const h16* q, const h16* k, const h16* v, const h16* dOT, const float* lse2, float* colsum,
                           const BwdInstTable& tab, int n_inst, int n_img, int heads, int L, int Lp, cudaStream_t st) {
    WC_CHECK_ARG(q && k && v && dOT && lse2 && colsum && n_inst > 0 && n_inst <= BWD_MAX_INST && n_img > 0 && heads > 0);
    WC_CHECK_ARG(L > 1 && Lp >= L && (Lp & 7) == 0 && ceil_div(L, AT_BN) * AT_BN <= BW_MAX_L && heads <= 65535);
    WC_SET_MAX_SMEM(attn_bwd_colsum_kernel, BwdSmem::TOTAL);
    CUtensorMap mq, mk, md;
    int rc = attn_maps(q, k, nullptr, n_img * heads, L, 0, AT_HD, &mq, &mk, nullptr);
    if (rc) return rc;
    int u0, u1;
    // dO'^T rows = 80 (dims, -delta hi/lo, zero padding), K extent = L queries: columns L..Lp-1 read as zero
    GemmOperand od{dOT, BW_ND, Lp, (long)BW_ND * Lp, 0};
    if ((rc = make_tma_map(&md, od, L, n_inst * heads, 1, BW_ND, &u0, &u1))) return rc;
    BwdArgs a{L, heads, ceil_div(L, AT_BN), lse2, v, colsum, tab};
    dim3 grid(ceil_div(L, AT_BM), heads, n_inst);
    ProfScope prof(PK_ATTENTION, 2.0 * L * L * (AT_HD + BW_ND) * heads * n_inst, st);
    if (g_prof_on) prof_tag(L, n_inst, heads, 2);
    attn_bwd_colsum_kernel<<<grid, AT_THREADS, BwdSmem::TOTAL, st>>>(mk, mq, md, a);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

}  // namespace wc

using namespace wc;

// Test / micro-bench entry for the fused attention pair (see include/weclip_b200.h).
extern "C" int wc_attention_fwd(const void* q, const void* k, const void* vt, void* o, float* lse2, float* attn_avg,
                                int n_img, int heads, int L, int Lp, int head_dim, void* stream) {
    WC_CHECK_ARG(lse2 || !attn_avg);
    WC_CHECK_ARG(head_dim == 64 || !attn_avg);
    int rc = launch_attention_fwd(reinterpret_cast<const h16*>(q), reinterpret_cast<const h16*>(k),
                                  reinterpret_cast<const h16*>(vt), reinterpret_cast<h16*>(o), lse2, n_img, heads, L, Lp,
                                  head_dim, as_stream(stream));
    if (rc || !attn_avg) return rc;
    return launch_attn_headmean(reinterpret_cast<const h16*>(q), reinterpret_cast<const h16*>(k), lse2, attn_avg, n_img,
                                heads, L, as_stream(stream));
}
