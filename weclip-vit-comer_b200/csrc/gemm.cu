// tcgen05 GEMM kernel (sm_100a), persistent and warp-specialised.  See gemm.cuh for the contract.
//
// grid = min(#tiles, #SMs) CTAs of 320 threads; each CTA walks output tiles t = blockIdx.x + i*gridDim.x
// (m-block fastest, so concurrently running CTAs share the same weight tile in L2).  Tile = 128 x BN, BN in {256,128,64,32}.
//   warp 0     TMA producer  — cp.async.bulk.tensor (rank-4 maps, 128B swizzle, OOB zero-fill) into a STAGES-deep smem ring
//                              (192 KB: 4 stages at BN=256 ... 8 stages at BN<=64), running ahead across tile boundaries
//   warp 1     MMA issuer    — one lane issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=BN, K=16) x4 per 64-wide k-block;
//                              tcgen05.commit frees the smem stage; accumulators are DOUBLE-BUFFERED in tensor memory
//                              (2 x BN fp32 columns) so the next tile's mainloop overlaps this tile's epilogue
//   warps 2-9  epilogue      — two warps per TMEM lane quarter, each draining half of the tile's columns:
//                              tcgen05.ld 32x32b.x32 (thread = accumulator row), transposed through a per-warp swizzled smem
//                              block so that global traffic is row-contiguous 128-bit accesses (8 lanes = 128 B of one row),
//                              fused bias / activation / residual / fp16 copy / head-major QKV scatter
// Operands are K-major fp16 (A: [M,K], B: [N,K]); accumulators are fp32.
#include "gemm.cuh"
#include "tc_ptx.cuh"

#include <cudaTypedefs.h>
#include <stdlib.h>
#include <mutex>
#include <unordered_map>

namespace wc {

constexpr int BM = 128;
constexpr int BK = 64;               // 64 fp16 = 128 B = one swizzle row
constexpr int EPI_WARPS = 8;         // two warps per TMEM lane quarter, each taking half of the tile's columns
constexpr int GEMM_THREADS = 64 + 32 * EPI_WARPS;
// Per-warp staging block: [32 rows][32 floats], 128-bit vectors XOR-swizzled by (row & 7) -> conflict-free row-per-thread
// writes and 8-lanes-per-row reads without padding (8 warps x 4 KB fit beside the 192 KB operand ring).
__device__ __forceinline__ int stg_off(int row, int vec) { return row * 32 + ((vec ^ (row & 7)) << 2); }
constexpr int RING_BYTES = 192 * 1024;

__device__ __forceinline__ float tanh_approx(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float apply_act(float x, int act) {
    switch (act) {
        case ACT_RELU: return fmaxf(x, 0.f);
        // x * sigmoid(1.702 x) = 0.5 x (1 + tanh(0.851 x)): ONE MUFU op per element instead of ex2 + rcp (the c_fc epilogue is
        // MUFU-paced: 128 x 256 outputs per tile at 16 MUFU/clk/SM).  tanh.approx.f32 is accurate to 2^-11 relative, the
        // same as the fp16 rounding of the result.
        case ACT_QUICKGELU: { const float h = 0.5f * x; return fmaf(h, tanh_approx(0.851f * x), h); }
        case ACT_SIGMOID: return __fdividef(1.f, 1.f + __expf(-x));
        default: return x;
    }
}

struct KernelArgs {
    int M, N, K, nb0;
    int a_b0, a_b1, b_b0, b_b1;   // 1 when the operand really has that batch dimension (else coordinate 0)
    int tiles_m, tiles_n, total_tiles;
    int vec_ok;                   // every pointer / stride of the epilogue allows 4-element vector access
    int variant;                  // variant_id() of the GM_STORE flags (fast epilogue dispatch)
    int n_fastest;                // tile raster (decode_tile)
    int dbg;                      // developer knob (WC_GEMM_DBG): 1 = epilogue only drains tensor memory, 3 = everything but the global stores
    GemmEpilogue e;
};

// Tile raster.  m fastest (default): the CTAs running together share one weight tile and stream A; the n-tiles of an m-block
// run tiles_m tiles apart, so A is re-read from HBM once per n-tile unless it fits L2.  n fastest: the n-tiles of an m-block
// run together and share the A tile through L2, the (small, L2-resident) weight matrix is re-read per m-block instead —
// chosen by the host when A is large and B small (c_proj: A = M x 3072 fp16 is 100-157 MB, read 3x otherwise).
__device__ __forceinline__ void decode_tile(int t, int tiles_m, int tiles_n, int n_fastest, int& mb, int& nb, int& bz) {
    if (n_fastest) {
        nb = t % tiles_n;
        const int rest = t / tiles_n;
        mb = rest % tiles_m; bz = rest / tiles_m;
    } else {
        mb = t % tiles_m;
        const int rest = t / tiles_m;
        nb = rest % tiles_n; bz = rest / tiles_n;
    }
}

__device__ __forceinline__ void store_h4(h16* p, const float* x) {
    h16x2 lo = __floats2half2_rn(x[0], x[1]), hi = __floats2half2_rn(x[2], x[3]);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(p) = pk;
}

// ---------------------------------------------------------------------------------------------------
// Fast epilogue paths.  The flags are template parameters (no per-row branches) and the 8 row iterations are fully
// unrolled so that the 8 LDS.128, the 8 residual LDG.128 and the 8 STG are each issued back to back.
// Preconditions (checked by the caller): full 32-column chunk, 4-element vector access allowed.
// ---------------------------------------------------------------------------------------------------
constexpr int variant_id(int f16, int act, int bias, int resid, int out2) {
    // resid: 0 none, 1 loaded and added in registers, 2 in place through an L2 reduction; out2: 0 none, 1 after act+resid, 2 before
    return f16 | (act << 1) | (bias << 3) | (resid << 4) | (out2 << 6);
}

template <int F16, int ACT, int BIAS, int RESID, int OUT2>
__device__ __forceinline__ void fast_store(const KernelArgs& a, const float* stg, int lane, int row0, int col0, int b0, int b1) {
    const GemmEpilogue& e = a.e;
    const int cc = (lane & 7) * 4, rsub = lane >> 3, col = col0 + cc;
    const int r_first = row0 + rsub;
    float4 bias = make_float4(0.f, 0.f, 0.f, 0.f);
    if (BIAS) bias = __ldg(reinterpret_cast<const float4*>(e.bias + b0 * e.bias_bs0 + b1 * e.bias_bs1 + col));
    float4 x[8], rs[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = *reinterpret_cast<const float4*>(stg + stg_off(i * 4 + rsub, lane & 7));
    if (RESID == 1) {
        const float* rp = e.resid + b0 * e.resid_bs0 + b1 * e.resid_bs1 + (long)r_first * e.ldr + col;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            rs[i] = (r_first + 4 * i < a.M) ? *reinterpret_cast<const float4*>(rp + (long)(4 * i) * e.ldr) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const long ob = b0 * e.out_bs0 + b1 * e.out_bs1 + (long)r_first * e.ldo + col;
    const long o2b = OUT2 ? b0 * e.out2_bs0 + b1 * e.out2_bs1 + (long)r_first * e.ld2 + col : 0;
    if (RESID == 2) {
        // in-place residual (out == resid, fp32): x += alpha*acc + bias as a vector reduction performed by L2
        // (red.global.add.v4.f32) — one fp32 add per element exactly like the register path, but the epilogue warps never
        // wait for the residual tile to come back from HBM
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (r_first + 4 * i >= a.M || a.dbg == 3) break;
            float* p = reinterpret_cast<float*>(e.out) + ob + (long)(4 * i) * e.ldo;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(e.alpha * x[i].x + bias.x),
                         "f"(e.alpha * x[i].y + bias.y), "f"(e.alpha * x[i].z + bias.z), "f"(e.alpha * x[i].w + bias.w) : "memory");
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (r_first + 4 * i >= a.M) break;
        float v[4] = {e.alpha * x[i].x + bias.x, e.alpha * x[i].y + bias.y, e.alpha * x[i].z + bias.z, e.alpha * x[i].w + bias.w};
        if (OUT2 == 2) store_h4(e.out2 + o2b + (long)(4 * i) * e.ld2, v);
#pragma unroll
        for (int t = 0; t < 4; ++t) v[t] = apply_act(v[t], ACT);
        if (RESID == 1) { v[0] += rs[i].x; v[1] += rs[i].y; v[2] += rs[i].z; v[3] += rs[i].w; }
        if (OUT2 == 1) store_h4(e.out2 + o2b + (long)(4 * i) * e.ld2, v);
        if (F16) store_h4(reinterpret_cast<h16*>(e.out) + ob + (long)(4 * i) * e.ldo, v);
        else *reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out) + ob + (long)(4 * i) * e.ldo) = make_float4(v[0], v[1], v[2], v[3]);
    }
}

// returns false when the variant is not instantiated (caller falls back to the generic path)
__device__ __forceinline__ bool fast_store_dispatch(const KernelArgs& a, const float* stg, int lane, int row0, int col0, int b0, int b1) {
#define WC_V(f, ac, bi, re, o2) \
    case variant_id(f, ac, bi, re, o2): fast_store<f, ac, bi, re, o2>(a, stg, lane, row0, col0, b0, b1); return true;
    switch (a.variant) {
        WC_V(0, ACT_NONE, 0, 1, 0)        // patch embed + positional embedding
        WC_V(0, ACT_NONE, 0, 0, 0)        // plain fp32 (scores, dP, c_fc^T)
        WC_V(1, ACT_NONE, 0, 0, 0)        // plain fp16 (P V, out_proj^T)
        WC_V(0, ACT_NONE, 1, 1, 0)        // out-proj / c_proj + residual
        WC_V(0, ACT_NONE, 1, 2, 0)        // out-proj + residual in place: L2 reduction
        WC_V(0, ACT_NONE, 1, 1, 1)        // c_proj + residual + fp16 feature tap
        WC_V(1, ACT_QUICKGELU, 1, 0, 0)   // c_fc
        WC_V(1, ACT_QUICKGELU, 1, 0, 2)   // c_fc keeping the pre-activation (Grad-CAM block)
        WC_V(1, ACT_RELU, 1, 0, 0)        // decoder fuse MLP layer 1
        WC_V(1, ACT_NONE, 1, 0, 0)        // decoder fuse MLP layer 2
        WC_V(0, ACT_NONE, 1, 0, 1)        // linear_fuse (+ fp16 copy)
        WC_V(0, ACT_SIGMOID, 0, 0, 0)     // attn_pred
        default: return false;
    }
#undef WC_V
}

__device__ __forceinline__ void fast_qkv(const KernelArgs& a, const float* stg, int lane, int row0, int col0, int which,
                                         int head, int d0) {
    const GemmEpilogue& e = a.e;
    const int cc = (lane & 7) * 4, rsub = lane >> 3;
    const float4 bias = __ldg(reinterpret_cast<const float4*>(e.bias + col0 + cc));
    const float sc = (which == 0) ? e.qscale : 1.f;     // q pre-scaled by 1/sqrt(hd) (myAtt.py:54)
    h16* dst = (which == 0) ? e.q : (which == 1) ? e.k : e.v;
    float4 x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = *reinterpret_cast<const float4*>(stg + stg_off(i * 4 + rsub, lane & 7));
    int row = row0 + rsub;
    int img = row / e.L, l = row - img * e.L;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (row >= a.M) break;
        float v[4] = {(x[i].x + bias.x) * sc, (x[i].y + bias.y) * sc, (x[i].z + bias.z) * sc, (x[i].w + bias.w) * sc};
        store_h4(dst + (((long)img * e.heads + head) * e.L + l) * e.hd + d0 + cc, v);
        row += 4; l += 4;
        if (l >= e.L) { l -= e.L; ++img; }          // L >= 4 is guaranteed by the caller
    }
}

// ---------------------------------------------------------------------------------------------------
// fp16-output fast path (c_fc, decoder MLPs, packed in-proj).  Bias / scale / activation are applied in the accumulator
// layout (lane = row, 32 consecutive columns in registers; the bias float4s are warp-uniform broadcast loads) and the chunk
// is staged as PACKED fp16: 2 KB each way through shared memory instead of 4 KB — the staging traffic competes with the
// tensor core's operand reads (64 B/clk of the SM's 128) — and 4 x (STS.128, LDS.128, STG.128) per lane instead of 8.
// Read-back layout: lane -> (row group lane >> 2, 16-byte slot lane & 3): a warp instruction writes 8 rows x 64 B.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ int stg16_off(int row, int slot) { return row * 64 + ((slot ^ ((row >> 1) & 3)) << 4); }   // bytes

template <int ACT, int BIAS, int SCALE_AFTER_BIAS>
__device__ __forceinline__ void pack_stage16(const uint32_t* r, float* stg, int lane, const float* bias, float alpha) {
    uint8_t* sb = reinterpret_cast<uint8_t*>(stg);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint4 pk;
        uint32_t* pw = reinterpret_cast<uint32_t*>(&pk);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (BIAS) b = __ldg(reinterpret_cast<const float4*>(bias) + 2 * j + h);
            const int c = 8 * j + 4 * h;
            float v[4] = {__uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3])};
            const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                v[t] = SCALE_AFTER_BIAS ? (v[t] + bb[t]) * alpha : alpha * v[t] + bb[t];
                v[t] = apply_act(v[t], ACT);
            }
            h16x2 lo = __floats2half2_rn(v[0], v[1]), hi = __floats2half2_rn(v[2], v[3]);
            pw[2 * h] = *reinterpret_cast<uint32_t*>(&lo);
            pw[2 * h + 1] = *reinterpret_cast<uint32_t*>(&hi);
        }
        *reinterpret_cast<uint4*>(sb + stg16_off(lane, j)) = pk;
    }
}

template <int ACT, int BIAS>
__device__ __forceinline__ void fast_store16(const KernelArgs& a, const uint32_t* r, float* stg, int lane, int row0, int col0,
                                             int b0, int b1) {
    const GemmEpilogue& e = a.e;
    pack_stage16<ACT, BIAS, 0>(r, stg, lane, BIAS ? e.bias + b0 * e.bias_bs0 + b1 * e.bias_bs1 + col0 : nullptr, e.alpha);
    __syncwarp();
    const uint8_t* sb = reinterpret_cast<const uint8_t*>(stg);
    const int rg = lane >> 2, slot = lane & 3;
    h16* o = reinterpret_cast<h16*>(e.out) + b0 * e.out_bs0 + b1 * e.out_bs1 + (long)(row0 + rg) * e.ldo + col0 + slot * 8;
    uint4 v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = *reinterpret_cast<const uint4*>(sb + stg16_off(i * 8 + rg, slot));
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (row0 + rg + 8 * i < a.M && a.dbg != 3) *reinterpret_cast<uint4*>(o + (long)(8 * i) * e.ldo) = v[i];
    __syncwarp();
}

// packed in-proj + head split on the same scheme; q is scaled by 1/sqrt(hd) after the bias (myAtt.py:54).  L >= 8.
__device__ __forceinline__ void fast_qkv16(const KernelArgs& a, const uint32_t* r, float* stg, int lane, int row0, int col0,
                                           int which, int head, int d0) {
    const GemmEpilogue& e = a.e;
    pack_stage16<ACT_NONE, 1, 1>(r, stg, lane, e.bias + col0, (which == 0) ? e.qscale : 1.f);
    __syncwarp();
    const uint8_t* sb = reinterpret_cast<const uint8_t*>(stg);
    const int rg = lane >> 2, slot = lane & 3;
    h16* dst = (which == 0) ? e.q : (which == 1) ? e.k : e.v;
    uint4 v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = *reinterpret_cast<const uint4*>(sb + stg16_off(i * 8 + rg, slot));
    int row = row0 + rg;
    int img = row / e.L, l = row - img * e.L;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (row < a.M && a.dbg != 3) *reinterpret_cast<uint4*>(dst + (((long)img * e.heads + head) * e.L + l) * e.hd + d0 + slot * 8) = v[i];
        row += 8; l += 8;
        if (l >= e.L) { l -= e.L; ++img; }
    }
    __syncwarp();
}

// GM_MUL_QGELU_GRAD fast path: out = acc * alpha * quickgelu'(u), u = saved pre-activation (fp16).  The 8 aux loads are issued
// back to back before any use; sigmoid(1.702 u) = 0.5 + 0.5 tanh(0.851 u) costs one MUFU op.
__device__ __forceinline__ void fast_qgelu_grad(const KernelArgs& a, const float* stg, int lane, int row0, int col0, int b0, int b1,
                                                int bz) {
    const GemmEpilogue& e = a.e;
    const int cc = (lane & 7) * 4, rsub = lane >> 3, col = col0 + cc, r_first = row0 + rsub;
    const int ab = e.aux_map ? e.aux_map[bz] : bz;
    const h16* u = e.aux + (long)ab * e.aux_bs + (long)r_first * e.ldaux + col;
    h16* o = reinterpret_cast<h16*>(e.out) + b0 * e.out_bs0 + b1 * e.out_bs1 + (long)r_first * e.ldo + col;
    uint2 uu[8];
    float4 x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        uu[i] = (r_first + 4 * i < a.M) ? *reinterpret_cast<const uint2*>(u + (long)(4 * i) * e.ldaux) : make_uint2(0u, 0u);
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = *reinterpret_cast<const float4*>(stg + stg_off(i * 4 + rsub, lane & 7));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (r_first + 4 * i >= a.M) break;
        const float2 f0 = __half22float2(*reinterpret_cast<const h16x2*>(&uu[i].x));
        const float2 f1 = __half22float2(*reinterpret_cast<const h16x2*>(&uu[i].y));
        const float uv[4] = {f0.x, f0.y, f1.x, f1.y};
        float v[4] = {x[i].x, x[i].y, x[i].z, x[i].w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const float sg = fmaf(0.5f, tanh_approx(0.851f * uv[t]), 0.5f);
            v[t] = v[t] * e.alpha * (sg * fmaf(1.702f * uv[t], 1.f - sg, 1.f));
        }
        store_h4(o + (long)(4 * i) * e.ldo, v);
    }
}

// One 32-row x 32-column block of the accumulator owned by this warp (thread `lane` holds row row0+lane, r[j] = column
// col0+j).  `stg` is the warp's private swizzled 32x32 fp32 staging block (stg_off).
__device__ __forceinline__ void epilogue_chunk(const KernelArgs& a, const uint32_t* r, float* stg, int lane, int row0,
                                               int col0, int b0, int b1, int bz) {
    const GemmEpilogue& e = a.e;
    const int ncol = min(32, a.N - col0);       // warp-uniform
    const int myrow = row0 + lane;

    if (e.mode == GM_STORE && e.out_trans) {
        // transposed output (element (m,n) at n*ldo+m): lanes are consecutive m -> already coalesced, no staging
        if (myrow < a.M) {
            const long ob = b0 * e.out_bs0 + b1 * e.out_bs1;
            const float* bias = e.bias ? e.bias + b0 * e.bias_bs0 + b1 * e.bias_bs1 + col0 : nullptr;
            const float* resid = e.resid ? e.resid + b0 * e.resid_bs0 + b1 * e.resid_bs1 + (long)col0 * e.ldr + myrow : nullptr;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < ncol) {
                    float x = e.alpha * __uint_as_float(r[j]);
                    if (bias) x += __ldg(bias + j);
                    x = apply_act(x, e.act);
                    if (resid) x += resid[(long)j * e.ldr];
                    if (e.out_f16) reinterpret_cast<h16*>(e.out)[ob + (long)(col0 + j) * e.ldo + myrow] = __float2half_rn(x);
                    else reinterpret_cast<float*>(e.out)[ob + (long)(col0 + j) * e.ldo + myrow] = x;
                }
            }
        }
        return;
    }

    int which = 0, head = 0, d0 = 0;
    if (e.mode == GM_QKV) {
        // clip/myAtt.py:199-201 (packed in-proj) + :257-268 (head split); a 32-column block never straddles a head
        const int E = e.heads * e.hd;
        which = col0 / E;
        const int within = col0 - which * E;
        head = within / e.hd;
        d0 = within - head * e.hd;
        if (which == 2 && e.vt && myrow < a.M) {
            // V^T [N*heads, hd, Lp]: lanes are consecutive tokens -> coalesced along l
            const int img = myrow / e.L, l = myrow - img * e.L;
            h16* o = e.vt + (((long)img * e.heads + head) * e.hd + d0) * (long)e.Lp + l;
            const float* bias = e.bias + col0;
#pragma unroll
            for (int j = 0; j < 32; ++j) o[(long)j * e.Lp] = __float2half_rn(__uint_as_float(r[j]) + __ldg(bias + j));
        }
        if ((which == 0 && !e.q) || (which == 1 && !e.k) || (which == 2 && !e.v)) return;
    }

    // ---- transpose through shared memory: row-per-thread -> 8 lanes x float4 per row
    {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(stg + stg_off(lane, j)) = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                  __uint_as_float(r[4 * j + 3]));
    }
    __syncwarp();
    if (ncol == 32) {
        if (e.mode == GM_STORE && a.vec_ok && fast_store_dispatch(a, stg, lane, row0, col0, b0, b1)) { __syncwarp(); return; }
        if (e.mode == GM_QKV && e.L >= 4) { fast_qkv(a, stg, lane, row0, col0, which, head, d0); __syncwarp(); return; }
        if (e.mode == GM_MUL_QGELU_GRAD && a.vec_ok) { fast_qgelu_grad(a, stg, lane, row0, col0, b0, b1, bz); __syncwarp(); return; }
    }
    const int cc = (lane & 7) * 4, rsub = lane >> 3;
    const int nv = min(4, ncol - cc);           // valid columns of this lane's group (<= 0: nothing)
    if (nv > 0) {
        const int col = col0 + cc;
        const bool vec = a.vec_ok && nv == 4;
        float bias4[4] = {0.f, 0.f, 0.f, 0.f};
        if (e.bias) {
            const float* bp = e.bias + (e.mode == GM_STORE ? b0 * e.bias_bs0 + b1 * e.bias_bs1 : 0) + col;
#pragma unroll
            for (int t = 0; t < 4; ++t) if (t < nv) bias4[t] = __ldg(bp + t);
        }
#pragma unroll 2
        for (int i = 0; i < 8; ++i) {
            const int rr = i * 4 + rsub, row = row0 + rr;
            if (row >= a.M) break;
            const float4 xv = *reinterpret_cast<const float4*>(stg + stg_off(rr, lane & 7));
            float x[4] = {xv.x, xv.y, xv.z, xv.w};
            if (e.mode == GM_STORE) {
#pragma unroll
                for (int t = 0; t < 4; ++t) x[t] = e.alpha * x[t] + bias4[t];
                h16* o2 = e.out2 ? e.out2 + b0 * e.out2_bs0 + b1 * e.out2_bs1 + (long)row * e.ld2 + col : nullptr;
                if (o2 && e.out2_preact) {
                    if (vec) store_h4(o2, x);
                    else for (int t = 0; t < nv; ++t) o2[t] = __float2half_rn(x[t]);
                }
#pragma unroll
                for (int t = 0; t < 4; ++t) x[t] = apply_act(x[t], e.act);
                if (e.resid) {
                    const float* rp = e.resid + b0 * e.resid_bs0 + b1 * e.resid_bs1 + (long)row * e.ldr + col;
                    if (vec) {
                        const float4 rv = *reinterpret_cast<const float4*>(rp);
                        x[0] += rv.x; x[1] += rv.y; x[2] += rv.z; x[3] += rv.w;
                    } else {
                        for (int t = 0; t < nv; ++t) x[t] += rp[t];
                    }
                }
                if (o2 && !e.out2_preact) {
                    if (vec) store_h4(o2, x);
                    else for (int t = 0; t < nv; ++t) o2[t] = __float2half_rn(x[t]);
                }
                const long ob = b0 * e.out_bs0 + b1 * e.out_bs1 + (long)row * e.ldo + col;
                if (e.out_f16) {
                    h16* o = reinterpret_cast<h16*>(e.out) + ob;
                    if (vec) store_h4(o, x);
                    else for (int t = 0; t < nv; ++t) o[t] = __float2half_rn(x[t]);
                } else {
                    float* o = reinterpret_cast<float*>(e.out) + ob;
                    if (vec) *reinterpret_cast<float4*>(o) = make_float4(x[0], x[1], x[2], x[3]);
                    else for (int t = 0; t < nv; ++t) o[t] = x[t];
                }
            } else if (e.mode == GM_QKV) {
                const float sc = (which == 0) ? e.qscale : 1.f;     // q pre-scaled by 1/sqrt(hd) (myAtt.py:54)
#pragma unroll
                for (int t = 0; t < 4; ++t) x[t] = (x[t] + bias4[t]) * sc;
                const int img = row / e.L, l = row - img * e.L;
                h16* dst = (which == 0) ? e.q : (which == 1) ? e.k : e.v;
                store_h4(dst + (((long)img * e.heads + head) * e.L + l) * e.hd + d0 + cc, x);
            } else {   // GM_MUL_QGELU_GRAD: out = acc * alpha * quickgelu'(u)
                const int ab = e.aux_map ? e.aux_map[bz] : bz;
                const h16* u = e.aux + (long)ab * e.aux_bs + (long)row * e.ldaux + col;
                h16* o = reinterpret_cast<h16*>(e.out) + b0 * e.out_bs0 + b1 * e.out_bs1 + (long)row * e.ldo + col;
                float uu[4] = {0.f, 0.f, 0.f, 0.f};
                if (vec) {
                    const uint2 pk = *reinterpret_cast<const uint2*>(u);
                    const float2 f0 = __half22float2(*reinterpret_cast<const h16x2*>(&pk.x));
                    const float2 f1 = __half22float2(*reinterpret_cast<const h16x2*>(&pk.y));
                    uu[0] = f0.x; uu[1] = f0.y; uu[2] = f1.x; uu[3] = f1.y;
                } else {
                    for (int t = 0; t < nv; ++t) uu[t] = __half2float(u[t]);
                }
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float s = 1.f / (1.f + __expf(-1.702f * uu[t]));
                    x[t] = x[t] * e.alpha * (s * (1.f + 1.702f * uu[t] * (1.f - s)));
                }
                if (vec) store_h4(o, x);
                else for (int t = 0; t < nv; ++t) o[t] = __float2half_rn(x[t]);
            }
        }
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------------
// Kernel
// ---------------------------------------------------------------------------------------------------
template <int BN>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (RING_BYTES / STAGE_BYTES) < 8 ? (RING_BYTES / STAGE_BYTES) : 8;
    static constexpr int STG_BYTES = EPI_WARPS * 32 * 32 * 4;
    static constexpr int TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;         // double-buffered accumulator (power of two)
    static constexpr int TOTAL = STAGES * STAGE_BYTES + STG_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA,
                                                                  const __grid_constant__ CUtensorMap mapB,
                                                                  const __grid_constant__ KernelArgs args) {
    using C = Cfg<BN>;
    constexpr int STAGES = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float* stg_all = reinterpret_cast<float*>(smem + STAGES * C::STAGE_BYTES);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES + C::STG_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;      // [2] accumulator ready   (MMA -> epilogue)
    uint64_t* tempty_bar = tfull_bar + 2;          // [2] accumulator drained (epilogue -> MMA)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_kb = (args.K + BK - 1) / BK;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&mapA);
        prefetch_tmap(&mapB);
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull_bar[s], 1); mbar_init(&tempty_bar[s], EPI_WARPS); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        uint32_t it = 0;     // running k-block counter across tiles
        for (int t = blockIdx.x; t < args.total_tiles; t += gridDim.x) {
            int mb, nb, bz;
            decode_tile(t, args.tiles_m, args.tiles_n, args.n_fastest, mb, nb, bz);
            const int b0 = bz % args.nb0, b1 = bz / args.nb0;
            if (lane == 0) {
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    mbar_expect_tx(&full_bar[s], C::STAGE_BYTES);
                    uint8_t* sa = smem + s * C::STAGE_BYTES;
                    tma_load_4d(sa, &mapA, &full_bar[s], kb * BK, mb * BM, b0 * args.a_b0, b1 * args.a_b1);
                    tma_load_4d(sa + C::A_BYTES, &mapB, &full_bar[s], kb * BK, nb * BN, b0 * args.b_b0, b1 * args.b_b1);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(BM, BN);
            uint32_t it = 0, tl = 0, full_ok = 0;
            for (int t = blockIdx.x; t < args.total_tiles; t += gridDim.x, ++tl) {
                const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
                mbar_wait(&tempty_bar[as], aph ^ 1);            // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tacc = tmem_base + as * BN;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    if (!full_ok) mbar_wait(&full_bar[s], ph);
                    // the NEXT stage's barrier round trip (~200 clk even when the stage has long landed) runs under this stage's
                    // MMA issue instead of in front of the next one
                    full_ok = mbar_try(&full_bar[(it + 1) % STAGES], ((it + 1) / STAGES) & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + s * C::STAGE_BYTES);
                    const uint64_t da = make_smem_desc(sa), db = make_smem_desc(sa + C::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)   // +32 B per UMMA_K inside the 128 B swizzle row -> +2 in the >>4 field
                        umma_f16(tacc, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&tfull_bar[as]);
            }
        }
    } else {
        const int q = warp & 3;                     // TMEM lane quarter this warp may access (warp id % 4)
        const int half = (warp - 2) >> 2;           // which half of the tile's 32-column chunks this warp drains
        constexpr int CH = BN / 32, NCH = (CH >= 2) ? CH / 2 : 1;
        const int c_begin = half * NCH;
        const bool has_work = (CH >= 2) || half == 0;
        float* stg = stg_all + (warp - 2) * 32 * 32;
        uint32_t tl = 0;
        for (int t = blockIdx.x; t < args.total_tiles; t += gridDim.x, ++tl) {
            int mb, nb, bz;
            decode_tile(t, args.tiles_m, args.tiles_n, args.n_fastest, mb, nb, bz);
            const int b0 = bz % args.nb0, b1 = bz / args.nb0;
            const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
            mbar_wait(&tfull_bar[as], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + as * BN + c_begin * 32 + ((uint32_t)(q * 32) << 16);
            const int row0 = mb * BM + q * 32;
            if (has_work && row0 < args.M && nb * BN + c_begin * 32 < args.N) {
                // software-pipelined TMEM reads: chunk c+1 is in flight while chunk c goes through the epilogue
                uint32_t r[2][32];
                tmem_ld32_issue(taddr, r[0]);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    const int col0 = nb * BN + (c_begin + c) * 32;
                    if (col0 >= args.N) break;          // warp-uniform
                    tmem_ld_wait();
                    if (c + 1 < NCH && col0 + 32 < args.N) tmem_ld32_issue(taddr + (c + 1) * 32, r[(c + 1) & 1]);
                    if (args.dbg != 1) epilogue_chunk(args, r[c & 1], stg, lane, row0, col0, b0, b1, bz);
                }
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[as]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------------
// Kernel-level epilogue specialisation.  The generic epilogue_chunk carries every mode and every fast variant; inlined
// once per 32-column chunk it makes the kernel ~30 k instructions and the epilogue warps run out of the instruction
// cache (measured: the K = 768 GEMMs were epilogue-bound at 950 TF/s with a 1400 TF/s main loop).  The hot shapes get a
// kernel instantiated for exactly one epilogue: FE >= 0 is a variant_id() of fast_store, FE_QKV / FE_QGRAD the two
// special modes; FE_GENERIC keeps the full dispatch.  Host-checked preconditions of a specialised kernel: N % 32 == 0
// (only full chunks), vector access allowed, GM_QKV: L >= 4.
// ---------------------------------------------------------------------------------------------------
constexpr int FE_GENERIC = -1, FE_QKV = -2, FE_QGRAD = -3;

template <int FE>
__device__ __forceinline__ void epilogue_chunk_t(const KernelArgs& a, const uint32_t* r, float* stg, int lane, int row0,
                                                 int col0, int b0, int b1, int bz) {
    if constexpr (FE == FE_GENERIC) {
        epilogue_chunk(a, r, stg, lane, row0, col0, b0, b1, bz);
    } else {
        const GemmEpilogue& e = a.e;
        int which = 0, head = 0, d0 = 0;
        if constexpr (FE == FE_QKV) {
            const int E = e.heads * e.hd;
            which = col0 / E;
            const int within = col0 - which * E;
            head = within / e.hd;
            d0 = within - head * e.hd;
            const int myrow = row0 + lane;
            if (which == 2 && e.vt && myrow < a.M) {
                const int img = myrow / e.L, l = myrow - img * e.L;
                h16* o = e.vt + (((long)img * e.heads + head) * e.hd + d0) * (long)e.Lp + l;
                const float* bias = e.bias + col0;
#pragma unroll
                for (int j = 0; j < 32; ++j) o[(long)j * e.Lp] = __float2half_rn(__uint_as_float(r[j]) + __ldg(bias + j));
            }
            if ((which == 0 && !e.q) || (which == 1 && !e.k) || (which == 2 && !e.v)) return;
        }
        constexpr bool STORE16 = FE >= 0 && (FE & 1) && ((FE >> 4) & 3) == 0 && ((FE >> 6) & 3) == 0;
        if constexpr (FE == FE_QKV) {
            fast_qkv16(a, r, stg, lane, row0, col0, which, head, d0);
        } else if constexpr (STORE16) {
            fast_store16<((FE >> 1) & 3), ((FE >> 3) & 1)>(a, r, stg, lane, row0, col0, b0, b1);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(stg + stg_off(lane, j)) = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                                                                 __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
            __syncwarp();
            if constexpr (FE == FE_QGRAD) fast_qgelu_grad(a, stg, lane, row0, col0, b0, b1, bz);
            else fast_store<(FE & 1), ((FE >> 1) & 3), ((FE >> 3) & 1), ((FE >> 4) & 3), ((FE >> 6) & 3)>(a, stg, lane, row0, col0, b0, b1);
            __syncwarp();
        }
    }
}


// ---------------------------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2): the two CTAs of a (2,1,1) cluster share one 256 x 256 output tile.  Each CTA stages its own
// 128 rows of A and HALF of the B tile (128 of the 256 n-rows); CTA 0 issues tcgen05.mma.cta_group::2 (M = 256), which reads A
// and B from both CTAs' shared memory and writes each CTA's 128 accumulator rows into that CTA's tensor memory.  Per CTA and
// k-block this halves the B bytes written by TMA and read by the tensor core (64 KB -> 32 KB of shared-memory traffic per 128
// MMA cycles), which is what caps the single-CTA 128 x 256 kernel near 57 % of the tensor peak.  Ring: 6 stages of 32 KB.
//   full_bar   (CTA 0's only)  both CTAs' TMA loads count their bytes on it
//   empty_bar  (each CTA's)    multicast tcgen05.commit: the stage may be overwritten in both CTAs
//   tfull_bar  (each CTA's)    multicast commit after a tile's last k-block: each CTA's epilogue drains its own 128 rows
//   tempty_bar (CTA 0's only)  16 arrivals: the 8 epilogue warps of CTA 0 and, remotely, the 8 of CTA 1
// ---------------------------------------------------------------------------------------------------
struct Cfg2 {
    static constexpr int BN = 256;
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = (BN / 2) * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = RING_BYTES / STAGE_BYTES;
    static constexpr int STG_BYTES = EPI_WARPS * 32 * 32 * 4;
    static constexpr int TMEM_COLS = 2 * BN;
    static constexpr int TOTAL = STAGES * STAGE_BYTES + STG_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

template <int FE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                const __grid_constant__ KernelArgs args) {
    using C = Cfg2;
    constexpr int STAGES = C::STAGES, BN = C::BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float* stg_all = reinterpret_cast<float*>(smem + STAGES * C::STAGE_BYTES);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES + C::STG_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int num_kb = (args.K + BK - 1) / BK;
    const int tiles_m2 = (args.tiles_m + 1) >> 1;
    const int total = tiles_m2 * args.tiles_n * (args.total_tiles / (args.tiles_m * args.tiles_n));

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&mapA);
        prefetch_tmap(&mapB);
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull_bar[s], 1); mbar_init(&tempty_bar[s], 2 * EPI_WARPS); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();          // the peer's barriers exist before any remote arrive / TMA byte count can reach them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = pair; t < total; t += npairs) {
                int m2, nb, bz;
                decode_tile(t, tiles_m2, args.tiles_n, args.n_fastest, m2, nb, bz);
                const int b0 = bz % args.nb0, b1 = bz / args.nb0;
                const int mb = m2 * 2 + (int)rank;          // may be one past the last m-block: TMA zero-fills, epilogue skips
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * C::STAGE_BYTES);
                    uint8_t* sa = smem + s * C::STAGE_BYTES;
                    const uint32_t fb = mapa_u32(&full_bar[s], 0);
                    tma_load_4d_pair(sa, &mapA, fb, kb * BK, mb * BM, b0 * args.a_b0, b1 * args.a_b1);
                    tma_load_4d_pair(sa + C::A_BYTES, &mapB, fb, kb * BK, nb * BN + (int)rank * (BN / 2), b0 * args.b_b0,
                                     b1 * args.b_b1);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            constexpr uint32_t idesc = make_idesc(2 * BM, BN);
            uint32_t it = 0, tl = 0, full_ok = 0;
            for (int t = pair; t < total; t += npairs, ++tl) {
                const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
                mbar_wait(&tempty_bar[as], aph ^ 1);            // both CTAs' epilogues have drained this accumulator
                tc_fence_after();
                const uint32_t tacc = tmem_base + as * BN;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    if (!full_ok) mbar_wait(&full_bar[s], ph);
                    // the NEXT stage's barrier round trip (~200 clk even when the stage has long landed) runs under this stage's
                    // MMA issue instead of in front of the next one
                    full_ok = mbar_try(&full_bar[(it + 1) % STAGES], ((it + 1) / STAGES) & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + s * C::STAGE_BYTES);
                    const uint64_t da = make_smem_desc(sa), db = make_smem_desc(sa + C::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) umma_f16_pair(tacc, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    umma_commit_pair(&empty_bar[s]);
                }
                umma_commit_pair(&tfull_bar[as]);
            }
        }
    } else {
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        constexpr int NCH = BN / 64;
        const int c_begin = half * NCH;
        float* stg = stg_all + (warp - 2) * 32 * 32;
        const uint32_t tempty_remote0 = mapa_u32(&tempty_bar[0], 0), tempty_remote1 = mapa_u32(&tempty_bar[1], 0);
        uint32_t tl = 0;
        for (int t = pair; t < total; t += npairs, ++tl) {
            int m2, nb, bz;
            decode_tile(t, tiles_m2, args.tiles_n, args.n_fastest, m2, nb, bz);
            const int b0 = bz % args.nb0, b1 = bz / args.nb0;
            const int mb = m2 * 2 + (int)rank;
            const uint32_t as = tl & 1, aph = (tl >> 1) & 1;
            mbar_wait(&tfull_bar[as], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + as * BN + c_begin * 32 + ((uint32_t)(q * 32) << 16);
            const int row0 = mb * BM + q * 32;
            if (row0 < args.M && nb * BN + c_begin * 32 < args.N) {
                uint32_t r[2][32];
                tmem_ld32_issue(taddr, r[0]);
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    const int col0 = nb * BN + (c_begin + c) * 32;
                    if (col0 >= args.N) break;          // warp-uniform
                    tmem_ld_wait();
                    if (c + 1 < NCH && col0 + 32 < args.N) tmem_ld32_issue(taddr + (c + 1) * 32, r[(c + 1) & 1]);
                    if (args.dbg != 1) epilogue_chunk_t<FE>(args, r[c & 1], stg, lane, row0, col0, b0, b1, bz);
                }
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(as ? tempty_remote1 : tempty_remote0);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();          // neither CTA leaves (or frees tensor memory) while the other may still touch it
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, C::TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------------
// Host side: tensor maps (driver entry point resolved at run time: no link against libcuda) + launch
// ---------------------------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    });
    return fn;
}

struct MapKey {
    const void* ptr; long rows, ld, bs0, bs1; int K, nb0, nb1, box_rows;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && rows == o.rows && ld == o.ld && bs0 == o.bs0 && bs1 == o.bs1 && K == o.K &&
               nb0 == o.nb0 && nb1 == o.nb1 && box_rows == o.box_rows;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = reinterpret_cast<size_t>(k.ptr);
        auto mix = [&](size_t v) { h ^= v + 0x9e3779b97f4a7c15ULL + (h << 6) + (h >> 2); };
        mix(k.rows); mix(k.ld); mix(k.bs0); mix(k.bs1); mix(k.K); mix(k.nb0); mix(k.nb1); mix(k.box_rows);
        return h;
    }
};

int make_tma_map(CUtensorMap* out, const GemmOperand& op, int K, int nb0, int nb1, int box_rows, int* use_b0, int* use_b1) {
    auto enc = get_encode();
    if (!enc) return fail(WC_ERR_CUDA, "%s: cuTensorMapEncodeTiled entry point unavailable", __func__);
    if ((reinterpret_cast<uintptr_t>(op.ptr) & 15) || (op.ld & 7) || (op.bs0 & 7) || (op.bs1 & 7))
        return fail(WC_ERR_ARG, "%s: operand pointer/strides must be 16-byte aligned (ld=%ld)", __func__, op.ld);
    *use_b0 = (nb0 > 1 && op.bs0 != 0);
    *use_b1 = (nb1 > 1 && op.bs1 != 0);
    thread_local std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    MapKey key{op.ptr, op.rows, op.ld, op.bs0, op.bs1, K, nb0, nb1, box_rows};
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return WC_OK; }
    const cuuint64_t d2 = *use_b0 ? nb0 : 1, d3 = *use_b1 ? nb1 : 1;
    cuuint64_t dims[4] = {(cuuint64_t)K, (cuuint64_t)op.rows, d2, d3};
    const cuuint64_t row_bytes = (cuuint64_t)op.ld * 2;
    const cuuint64_t s2 = *use_b0 ? (cuuint64_t)op.bs0 * 2 : row_bytes * (cuuint64_t)op.rows;
    const cuuint64_t s3 = *use_b1 ? (cuuint64_t)op.bs1 * 2 : s2 * d2;
    cuuint64_t strides[3] = {row_bytes, s2, s3};
    cuuint32_t box[4] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<h16*>(op.ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(WC_ERR_CUDA, "%s: cuTensorMapEncodeTiled failed (%ld) rows=%ld", __func__, (long)r, op.rows);
    if (cache.size() > 4096) cache.clear();
    cache.emplace(key, *out);
    return WC_OK;
}

static int num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
    }
    return n;
}

static bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

// 4-element vector access in the epilogue needs every touched pointer 16 B (fp32) / 8 B (fp16) aligned and every stride
// a multiple of 4 elements; otherwise the kernel falls back to scalar accesses (same results).
static int epilogue_vec_ok(const GemmEpilogue& e) {
    auto m4 = [](long v) { return (v & 3) == 0; };
    bool ok = true;
    if (e.mode == GM_STORE) {
        if (e.out_trans) return 0;
        ok = ok && aligned(e.out, e.out_f16 ? 8 : 16) && m4(e.ldo) && m4(e.out_bs0) && m4(e.out_bs1);
        if (e.bias) ok = ok && aligned(e.bias, 16) && m4(e.bias_bs0) && m4(e.bias_bs1);
        if (e.resid) ok = ok && aligned(e.resid, 16) && m4(e.ldr) && m4(e.resid_bs0) && m4(e.resid_bs1);
        if (e.out2) ok = ok && aligned(e.out2, 8) && m4(e.ld2) && m4(e.out2_bs0) && m4(e.out2_bs1);
    } else if (e.mode == GM_MUL_QGELU_GRAD) {
        ok = ok && aligned(e.out, 8) && m4(e.ldo) && m4(e.out_bs0) && m4(e.out_bs1) && aligned(e.aux, 8) && m4(e.ldaux) && m4(e.aux_bs);
    }
    return ok ? 1 : 0;
}

// 2 = the residual is the output buffer itself (fp32, no second output, no activation): add through red.global instead of
// loading it; 1 = load + add; 0 = none
static int resid_mode(const GemmEpilogue& e) {
    if (!e.resid) return 0;
    static const bool red_ok = [] { const char* v = getenv("WC_GEMM_RED"); return !(v && v[0] == '0'); }();
    const bool in_place = e.resid == e.out && !e.out_f16 && !e.out2 && !e.out_trans && e.act == ACT_NONE && e.bias &&
                          e.ldr == e.ldo && e.resid_bs0 == e.out_bs0 && e.resid_bs1 == e.out_bs1 && e.mode == GM_STORE;
    return (in_place && red_ok) ? 2 : 1;
}

static int dbg_knob() {
    static const int v = [] { const char* e = getenv("WC_GEMM_DBG"); return e ? atoi(e) : 0; }();
    return v;
}

// A streamed once instead of once per n-tile: worth it when A does not fit L2 next to everything else and B does
static int raster_n_fastest(const GemmProblem& p) {
    const double a_bytes = 2.0 * p.M * p.K, b_bytes = 2.0 * p.N * p.K;
    return (p.nb0 * p.nb1 == 1 && p.N > 256 && a_bytes > 48e6 && b_bytes < 16e6) ? 1 : 0;
}

template <int BN>
static int launch_cfg(const GemmProblem& p, cudaStream_t st) {
    using C = Cfg<BN>;
    WC_SET_MAX_SMEM(gemm_tc_kernel<BN>, C::TOTAL);
    CUtensorMap mA, mB;
    KernelArgs ka;
    ka.M = p.M; ka.N = p.N; ka.K = p.K; ka.nb0 = p.nb0; ka.e = p.epi;
    int rc = make_tma_map(&mA, p.A, p.K, p.nb0, p.nb1, BM, &ka.a_b0, &ka.a_b1);
    if (rc) return rc;
    rc = make_tma_map(&mB, p.B, p.K, p.nb0, p.nb1, BN, &ka.b_b0, &ka.b_b1);
    if (rc) return rc;
    ka.tiles_m = ceil_div(p.M, BM);
    ka.tiles_n = ceil_div(p.N, BN);
    const long total = (long)ka.tiles_m * ka.tiles_n * p.nb0 * p.nb1;
    if (total > 0x7fffffffL) return fail(WC_ERR_ARG, "%s: too many tiles", __func__);
    ka.total_tiles = (int)total;
    ka.vec_ok = epilogue_vec_ok(p.epi);
    ka.n_fastest = raster_n_fastest(p);
    ka.dbg = dbg_knob();
    ka.variant = variant_id(p.epi.out_f16 ? 1 : 0, p.epi.act, p.epi.bias ? 1 : 0, resid_mode(p.epi),
                            p.epi.out2 ? (p.epi.out2_preact ? 2 : 1) : 0);
    const int grid = (int)(total < num_sms() ? total : num_sms());
    ProfScope prof(PK_GEMM, 2.0 * p.M * p.N * p.K * p.nb0 * p.nb1, st);
    if (g_prof_on) prof_tag(p.M, p.N, p.K, p.nb0 * p.nb1);
    gemm_tc_kernel<BN><<<grid, GEMM_THREADS, C::TOTAL, st>>>(mA, mB, ka);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

template <int FE>
static int launch_pair_fe(const CUtensorMap& mA, const CUtensorMap& mB, const KernelArgs& ka, int grid, cudaStream_t st) {
    WC_SET_MAX_SMEM(gemm_tc2_kernel<FE>, Cfg2::TOTAL);
    gemm_tc2_kernel<FE><<<grid, GEMM_THREADS, Cfg2::TOTAL, st>>>(mA, mB, ka);
    return WC_OK;
}

// epilogue specialisation of the pair kernel for this problem (FE_GENERIC when no instantiation matches)
static int pick_fast_epilogue(const GemmProblem& p, const KernelArgs& ka) {
    static const bool on = [] { const char* v = getenv("WC_GEMM_FE"); return !(v && v[0] == '0'); }();
    if (!on || (p.N & 31) || !ka.vec_ok) return FE_GENERIC;
    auto a16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if (p.epi.mode == GM_QKV) {
        const GemmEpilogue& e = p.epi;
        const bool ok = e.L >= 8 && (e.hd & 7) == 0 && (!e.q || a16(e.q)) && (!e.k || a16(e.k)) && (!e.v || a16(e.v));
        return ok ? FE_QKV : FE_GENERIC;
    }
    if (p.epi.mode == GM_MUL_QGELU_GRAD) return FE_QGRAD;
    if (p.epi.mode != GM_STORE || p.epi.out_trans) return FE_GENERIC;
    // fp16 outputs without residual / second output go through 16-byte vector stores
    if (p.epi.out_f16 && !p.epi.resid && !p.epi.out2 &&
        !(a16(p.epi.out) && (p.epi.ldo & 7) == 0 && (p.epi.out_bs0 & 7) == 0 && (p.epi.out_bs1 & 7) == 0))
        return FE_GENERIC;
    return ka.variant;
}

static int launch_pair(const GemmProblem& p, cudaStream_t st) {
    using C = Cfg2;
    CUtensorMap mA, mB;
    KernelArgs ka;
    ka.M = p.M; ka.N = p.N; ka.K = p.K; ka.nb0 = p.nb0; ka.e = p.epi;
    int rc = make_tma_map(&mA, p.A, p.K, p.nb0, p.nb1, BM, &ka.a_b0, &ka.a_b1);
    if (rc) return rc;
    rc = make_tma_map(&mB, p.B, p.K, p.nb0, p.nb1, C::BN / 2, &ka.b_b0, &ka.b_b1);
    if (rc) return rc;
    ka.tiles_m = ceil_div(p.M, BM);
    ka.tiles_n = ceil_div(p.N, C::BN);
    const long total = (long)ka.tiles_m * ka.tiles_n * p.nb0 * p.nb1;
    if (total > 0x7fffffffL) return fail(WC_ERR_ARG, "%s: too many tiles", __func__);
    ka.total_tiles = (int)total;
    ka.vec_ok = epilogue_vec_ok(p.epi);
    ka.n_fastest = raster_n_fastest(p);
    ka.dbg = dbg_knob();
    ka.variant = variant_id(p.epi.out_f16 ? 1 : 0, p.epi.act, p.epi.bias ? 1 : 0, resid_mode(p.epi),
                            p.epi.out2 ? (p.epi.out2_preact ? 2 : 1) : 0);
    const long pair_tiles = (long)((ka.tiles_m + 1) / 2) * ka.tiles_n * p.nb0 * p.nb1;
    const long pairs = num_sms() / 2;
    const int grid = 2 * (int)(pair_tiles < pairs ? pair_tiles : pairs);
    ProfScope prof(PK_GEMM, 2.0 * p.M * p.N * p.K * p.nb0 * p.nb1, st);
    if (g_prof_on) prof_tag(p.M, p.N, p.K, p.nb0 * p.nb1);
#define WC_FE(id) case id: rc = launch_pair_fe<id>(mA, mB, ka, grid, st); break;
    switch (pick_fast_epilogue(p, ka)) {
        WC_FE(FE_QKV)                                       // packed in-proj + head split
        WC_FE(FE_QGRAD)                                     // Grad-CAM c_proj^T with the QuickGELU' factor
        WC_FE(variant_id(1, ACT_QUICKGELU, 1, 0, 0))        // c_fc
        WC_FE(variant_id(1, ACT_QUICKGELU, 1, 0, 2))        // c_fc keeping the pre-activation (Grad-CAM block)
        WC_FE(variant_id(0, ACT_NONE, 1, 2, 0))             // out-proj / c_proj adding into x in place
        WC_FE(variant_id(0, ACT_NONE, 1, 1, 1))             // c_proj + residual + fp16 tap (last encoder block)
        WC_FE(variant_id(0, ACT_NONE, 1, 1, 0))             // + residual into another buffer
        WC_FE(variant_id(1, ACT_RELU, 1, 0, 0))             // decoder fuse MLP layer 1
        WC_FE(variant_id(1, ACT_NONE, 1, 0, 0))             // decoder fuse MLP layer 2
        WC_FE(variant_id(0, ACT_NONE, 1, 0, 1))             // linear_fuse (+ fp16 copy)
        WC_FE(variant_id(0, ACT_NONE, 0, 0, 0))             // plain fp32
        WC_FE(variant_id(1, ACT_NONE, 0, 0, 0))             // plain fp16
        default: rc = launch_pair_fe<FE_GENERIC>(mA, mB, ka, grid, st); break;
    }
#undef WC_FE
    if (rc) return rc;
    WC_LAUNCH_CHECK();
    return WC_OK;
}

int launch_gemm(const GemmProblem& p, cudaStream_t st) {
    WC_CHECK_ARG(p.M > 0 && p.N > 0 && p.K > 0 && p.nb0 > 0 && p.nb1 > 0 && p.A.ptr && p.B.ptr);
    if (p.epi.mode == GM_QKV) WC_CHECK_ARG(p.N % 32 == 0 && p.epi.hd % 32 == 0 && p.epi.bias && aligned(p.epi.bias, 16));
    {   // developer override (tools/gemm_bench.py): WC_GEMM_FORCE_BN = 256 / 128 / 64 runs every problem on cta_group::1 tiles of that width
        static const int force = [] { const char* e = getenv("WC_GEMM_FORCE_BN"); return e ? atoi(e) : 0; }();
        if (force == 256 && p.N > 128) return launch_cfg<256>(p, st);
        if (force == 128 && p.N > 64) return launch_cfg<128>(p, st);
        if (force == 64 && p.N > 32) return launch_cfg<64>(p, st);
    }
    if (p.N <= 32) return launch_cfg<32>(p, st);
    if (p.N <= 64) return launch_cfg<64>(p, st);
    if (p.N <= 128) return launch_cfg<128>(p, st);
    // widest tile that still gives every SM at least one tile (tile quantisation matters more than tile shape on small M)
    const long per_n = (long)ceil_div(p.M, BM) * p.nb0 * p.nb1;
    if (per_n * ceil_div(p.N, 256) >= num_sms()) {
        // enough 256-wide tiles for every SM: CTA pairs (256 x 256 tiles, cta_group::2) unless switched off for A/B runs
        static const bool pair_ok = [] { const char* e = getenv("WC_GEMM_PAIR"); return !(e && e[0] == '0'); }();
        if (pair_ok && ceil_div(p.M, BM) >= 2) return launch_pair(p, st);
        return launch_cfg<256>(p, st);
    }
    if (per_n * ceil_div(p.N, 128) >= num_sms()) return launch_cfg<128>(p, st);
    return launch_cfg<64>(p, st);
}

}  // namespace wc

using namespace wc;

// Generic test / micro-bench entry: out[b] = act(alpha * A[b] @ B[b]^T + bias) (+ resid), see include/weclip_b200.h.
extern "C" int wc_gemm_f16(const void* A, const void* B, void* out, const float* bias, const float* resid,
                            int M, int N, int K, int batch, long lda, long ldb, long ldo, long a_bs, long b_bs, long o_bs,
                            int out_f16, int out_trans, int act, float alpha, void* stream) {
    WC_CHECK_ARG(A && B && out);
    GemmProblem p;
    p.M = M; p.N = N; p.K = K; p.nb0 = batch; p.nb1 = 1;
    p.A = {reinterpret_cast<const h16*>(A), M, lda, a_bs, 0};
    p.B = {reinterpret_cast<const h16*>(B), N, ldb, b_bs, 0};
    p.epi.mode = GM_STORE; p.epi.act = act; p.epi.alpha = alpha; p.epi.bias = bias;
    p.epi.out = out; p.epi.out_f16 = out_f16; p.epi.out_trans = out_trans; p.epi.ldo = ldo; p.epi.out_bs0 = o_bs;
    p.epi.resid = resid; p.epi.ldr = ldo; p.epi.resid_bs0 = o_bs;
    return launch_gemm(p, as_stream(stream));
}

// Two batch dimensions (e.g. heads x images of an attention product) with independent strides for A, B and out.
extern "C" int wc_gemm_f16_ex(const void* A, const void* B, void* out, const float* bias, const float* resid,
                               int M, int N, int K, int nb0, int nb1, long lda, long ldb, long ldo,
                               long a_bs0, long a_bs1, long b_bs0, long b_bs1, long o_bs0, long o_bs1,
                               long bias_bs0, long bias_bs1,
                               int out_f16, int out_trans, int act, float alpha, void* stream) {
    WC_CHECK_ARG(A && B && out && nb0 > 0 && nb1 > 0);
    GemmProblem p;
    p.M = M; p.N = N; p.K = K; p.nb0 = nb0; p.nb1 = nb1;
    p.A = {reinterpret_cast<const h16*>(A), M, lda, a_bs0, a_bs1};
    p.B = {reinterpret_cast<const h16*>(B), N, ldb, b_bs0, b_bs1};
    p.epi.mode = GM_STORE; p.epi.act = act; p.epi.alpha = alpha; p.epi.bias = bias;
    p.epi.bias_bs0 = bias_bs0; p.epi.bias_bs1 = bias_bs1;
    p.epi.out = out; p.epi.out_f16 = out_f16; p.epi.out_trans = out_trans; p.epi.ldo = ldo;
    p.epi.out_bs0 = o_bs0; p.epi.out_bs1 = o_bs1;
    p.epi.resid = resid; p.epi.ldr = ldo; p.epi.resid_bs0 = o_bs0; p.epi.resid_bs1 = o_bs1;
    return launch_gemm(p, as_stream(stream));
}
