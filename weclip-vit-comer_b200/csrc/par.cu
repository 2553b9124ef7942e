// PAR (pixel-adaptive refinement) for sm_100a.  Reference: WeCLIP_model/PAR.py:26-92.
//
// Memory-bound stencil family.  Data layout in HBM:
//   img   [B,3,H,W]    fp32 planar (resized copy lives in the workspace when Hi,Wi != H,W)
//   aff   [B,H,Wb,NN,64]  blocked: NN = 8*n_dil (48) neighbours x 64 consecutive pixels form one contiguous block
//                      (Wb = ceil(W/64) blocks per image row), so the 48 weights a 4-pixel thread needs sit at IMMEDIATE
//                      offsets n*64 from one base pointer (no address arithmetic per load) and a tile row streams one
//                      contiguous 6 KB (fp16) span instead of touching 48 planes 0.5 MB apart; stored as fp32 (exact path)
//                      or fp16 (pipeline path: halves the bytes every propagation step streams)
//   masks [sumKc,H,W]  fp32 planar, ragged over images; two ping-pong buffers [sumKc,H,Wp] in the workspace
// Kernels (fast path = the module's own dilation set 1/2/4/8/12/24):
//   par_resize_kernel     bilinear align_corners=True (PAR.py:67)
//   par_aff_tile_kernel   64x32-pixel tile + 24-pixel replicate halo of the 3 image planes in shared memory; per pixel
//                         mean / unbiased std / logits over the 48 neighbours from shared memory, softmax +
//                         w2*softmax(pos) (PAR.py:70-86).  Algorithmic bytes/pixel: 12 + s_a*NN.
//   par_iter_tile_kernel  one propagation step (PAR.py:88-90): 64xTH tile + halo of the KC mask planes in shared
//                         memory (filled with 128-bit loads, clamped = replicate padding), each thread owns 4 consecutive
//                         pixels: per neighbour ONE 128/64-bit streaming affinity load (L1 no-allocate, L2 evict-first) and
//                         aligned LDS.128 mask vectors (shifted neighbours are register renames).
//                         Algorithmic bytes/pixel/iteration: s_a*NN + 2*4*Kc.
//   par_aff_kernel / par_iter_kernel   generic dilation sets (any n_dil <= 8), L1-cached gathers.
#include "common.cuh"

#include <math.h>
#include <stdlib.h>
#include <vector>

namespace wc {

constexpr int PAR_MAX_NN = 8 * WC_PAR_MAX_DIL;
constexpr int PAR_TX = 32;   // generic kernels: one warp = 32 consecutive pixels of a row
constexpr int PAR_TY = 16;
constexpr int PAR_KC = 4;    // mask channels per thread / per chunk
constexpr int PAR_MAX_CHUNKS = 240;

constexpr int PB = 64;       // pixels per affinity block (= tile width of the fast path)

struct ParGeom {
    int H, W, nn;
    int wp;                  // row pitch of the ping-pong mask planes
    int wb;                  // affinity blocks per image row
    short dy[PAR_MAX_NN];
    short dx[PAR_MAX_NN];
};
struct ParPos {
    float v[PAR_MAX_NN];   // w2 * softmax(pos_aff), data independent (PAR.py:51-62,83,86)
};
struct ParChunks {
    short b[PAR_MAX_CHUNKS];     // image index (selects the affinity)
    short ch0[PAR_MAX_CHUNKS];   // first packed channel
    short cnt[PAR_MAX_CHUNKS];   // channels in this chunk (1..PAR_KC)
};

inline int par_pitch(int W) { return (W + 7) & ~7; }
inline int par_aff_pitch(int W) { return (W + PB - 1) / PB * PB; }
// element index of affinity (b, n, y, x) in the blocked layout
__host__ __device__ __forceinline__ size_t aff_index(int H, int wb, int nn, int b, int n, int y, int x) {
    return ((((size_t)b * H + y) * wb + (x / PB)) * nn + n) * PB + (x % PB);
}

static int make_geom(ParGeom& g, ParPos* pos, int H, int W, const int* dil, int n_dil) {
    if (n_dil < 1 || n_dil > WC_PAR_MAX_DIL) return -1;
    static const int by[8] = {-1, -1, -1, 0, 0, 1, 1, 1};   // PAR.py:10-24 raster order without the centre
    static const int bx[8] = {-1, 0, 1, -1, 1, -1, 0, 1};
    g.H = H; g.W = W; g.nn = 8 * n_dil; g.wp = par_pitch(W); g.wb = (W + PB - 1) / PB;
    for (int d = 0; d < n_dil; ++d)
        for (int j = 0; j < 8; ++j) {
            g.dy[d * 8 + j] = (short)(by[j] * dil[d]);
            g.dx[d * 8 + j] = (short)(bx[j] * dil[d]);
        }
    if (pos) {
        // PAR.py:51-62: pos = d * {sqrt2,1,sqrt2,1,1,sqrt2,1,sqrt2}; :78 unbiased std over all nn; :83,:86.
        // Evaluated in fp32 with the same operation order as torch (the values are constants of the module).
        const int nn = g.nn;
        float p[PAR_MAX_NN];
        const float s2 = (float)sqrt(2.0);
        for (int d = 0; d < n_dil; ++d)
            for (int j = 0; j < 8; ++j) p[d * 8 + j] = ((bx[j] != 0 && by[j] != 0) ? s2 : 1.f) * (float)dil[d];
        double mean = 0; for (int i = 0; i < nn; ++i) mean += p[i]; mean /= nn;
        double ss = 0; for (int i = 0; i < nn; ++i) ss += (p[i] - mean) * (p[i] - mean);
        const float sd = (float)sqrt(ss / (nn - 1));
        float l[PAR_MAX_NN], mx = -INFINITY;
        for (int i = 0; i < nn; ++i) { float t = p[i] / (sd + 1e-8f) / 0.3f; l[i] = -(t * t); mx = fmaxf(mx, l[i]); }
        double sum = 0; for (int i = 0; i < nn; ++i) sum += exp((double)(l[i] - mx));
        for (int i = 0; i < nn; ++i) pos->v[i] = 0.01f * (float)(exp((double)(l[i] - mx)) / sum);
    }
    return 0;
}

static bool is_standard_dilations(const int* dil, int n_dil) {
    static const int std_d[6] = {1, 2, 4, 8, 12, 24};
    if (n_dil != 6) return false;
    for (int i = 0; i < 6; ++i) if (dil[i] != std_d[i]) return false;
    return true;
}

// ---- affinity storage: fp32 (exact path) or fp16 (pipeline path) ---------------------------------------
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_evict_normal_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ float ex2_approx(float x) {      // one MUFU; x <= 0 here
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
template <typename T> struct AffIO;
template <> struct AffIO<float> {
    static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
    static __device__ __forceinline__ float ld(const float* p) {
        float v;
        asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
        return v;
    }
    using Raw = float4;               // 4 consecutive pixels' weights as loaded
    static constexpr int LOOKAHEAD = 1;   // dilations prefetched ahead of the one being accumulated
    static __device__ __forceinline__ Raw ld4(const float* p, uint64_t pol) {
        float4 v;
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol));
        return v;
    }
    static __device__ __forceinline__ float4 unpack(const Raw& r) { return r; }
};
template <> struct AffIO<__half> {
    static __device__ __forceinline__ void st(__half* p, float v) { *p = __float2half_rn(v); }
    static __device__ __forceinline__ float ld(const __half* p) {
        unsigned short u;
        asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(u) : "l"(p));
        return __half2float(__ushort_as_half(u));
    }
    using Raw = uint2;                // 4 x fp16 stay packed until they are used (2 registers per neighbour)
#ifndef WC_PAR_LOOKAHEAD16
#define WC_PAR_LOOKAHEAD16 1     // weight sets prefetched beyond the one in use (1: two sets of 16 registers; 2: three)
#endif
    static constexpr int LOOKAHEAD = WC_PAR_LOOKAHEAD16;
    static __device__ __forceinline__ Raw ld4(const __half* p, uint64_t pol) {
        uint2 v;
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0, %1}, [%2], %3;"
                     : "=r"(v.x), "=r"(v.y) : "l"(p), "l"(pol));
        return v;
    }
    static __device__ __forceinline__ float4 unpack(const Raw& r) {
        const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&r.x));
        const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&r.y));
        return make_float4(lo.x, lo.y, hi.x, hi.y);
    }
};

// ---------------------------------------------------------------------------------------------------
__global__ void par_resize_kernel(const float* __restrict__ src, float* __restrict__ dst, int planes,
                                  int Hi, int Wi, int H, int W) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int p = blockIdx.z;
    if (x >= W || y >= H || p >= planes) return;
    // torch upsample_bilinear2d, align_corners=True: src = dst * (in-1)/(out-1)
    const float sy = (H > 1) ? (float)(Hi - 1) / (float)(H - 1) : 0.f;
    const float sx = (W > 1) ? (float)(Wi - 1) / (float)(W - 1) : 0.f;
    const float fy = sy * y, fx = sx * x;
    const int y0 = (int)fy, x0 = (int)fx;
    const int y1 = y0 + (y0 < Hi - 1 ? 1 : 0), x1 = x0 + (x0 < Wi - 1 ? 1 : 0);
    const float ly = fy - y0, lx = fx - x0;
    const float* s = src + (size_t)p * Hi * Wi;
    const float v = (1.f - ly) * ((1.f - lx) * s[(size_t)y0 * Wi + x0] + lx * s[(size_t)y0 * Wi + x1]) +
                    ly * ((1.f - lx) * s[(size_t)y1 * Wi + x0] + lx * s[(size_t)y1 * Wi + x1]);
    dst[(size_t)p * H * W + (size_t)y * W + x] = v;
}

// ---------------------------------------------------------------------------------------------------
// Generic dilation sets (runtime neighbour table).  Neighbour gathers go through L1.
// ---------------------------------------------------------------------------------------------------
template <typename AffT>
__global__ void __launch_bounds__(256) par_aff_kernel(const float* __restrict__ img, AffT* __restrict__ aff,
                                                      const __grid_constant__ ParGeom g,
                                                      const __grid_constant__ ParPos pos) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int b = blockIdx.z;
    const int H = g.H, W = g.W;
    if (x >= W || y >= H) return;
    const int nn = g.nn;
    const size_t hw = (size_t)H * W;
    const float* im = img + (size_t)b * 3 * hw;
    const float c0 = im[(size_t)y * W + x], c1 = im[hw + (size_t)y * W + x], c2 = im[2 * hw + (size_t)y * W + x];

    // pass 1: per-channel mean of the neighbour values (re-reads in later passes hit L1)
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll 8
    for (int n = 0; n < nn; ++n) {
        const int yy = min(max(y + g.dy[n], 0), H - 1), xx = min(max(x + g.dx[n], 0), W - 1);
        const size_t o = (size_t)yy * W + xx;
        s0 += __ldg(im + o); s1 += __ldg(im + hw + o); s2 += __ldg(im + 2 * hw + o);
    }
    const float inv_n = 1.f / (float)nn;
    const float m0 = s0 * inv_n, m1 = s1 * inv_n, m2 = s2 * inv_n;
    // pass 2: centred sum of squares -> unbiased std (torch.std, PAR.py:77)
    float q0 = 0.f, q1 = 0.f, q2 = 0.f;
#pragma unroll 8
    for (int n = 0; n < nn; ++n) {
        const int yy = min(max(y + g.dy[n], 0), H - 1), xx = min(max(x + g.dx[n], 0), W - 1);
        const size_t o = (size_t)yy * W + xx;
        const float d0 = __ldg(im + o) - m0, d1 = __ldg(im + hw + o) - m1, d2 = __ldg(im + 2 * hw + o) - m2;
        q0 += d0 * d0; q1 += d1 * d1; q2 += d2 * d2;
    }
    const float inv_nm1 = 1.f / (float)(nn - 1);
    const float e0 = sqrtf(q0 * inv_nm1) + 1e-8f, e1 = sqrtf(q1 * inv_nm1) + 1e-8f, e2 = sqrtf(q2 * inv_nm1) + 1e-8f;
    // pass 3: logits = -mean_c((|v-c|/(std+1e-8)/w1)^2) (PAR.py:80-81); row max for the softmax
    auto logit = [&](int n) {
        const int yy = min(max(y + g.dy[n], 0), H - 1), xx = min(max(x + g.dx[n], 0), W - 1);
        const size_t o = (size_t)yy * W + xx;
        const float t0 = fabsf(__ldg(im + o) - c0) / e0 / 0.3f;
        const float t1 = fabsf(__ldg(im + hw + o) - c1) / e1 / 0.3f;
        const float t2 = fabsf(__ldg(im + 2 * hw + o) - c2) / e2 / 0.3f;
        return -(t0 * t0 + t1 * t1 + t2 * t2) / 3.f;
    };
    float mx = -INFINITY;
#pragma unroll 8
    for (int n = 0; n < nn; ++n) mx = fmaxf(mx, logit(n));
    // pass 4: softmax denominator; pass 5: write softmax + w2*softmax(pos) (PAR.py:86)
    float sum = 0.f;
#pragma unroll 8
    for (int n = 0; n < nn; ++n) sum += expf(logit(n) - mx);
    const float inv = 1.f / sum;
    AffT* a = aff + aff_index(H, g.wb, nn, b, 0, y, x);
#pragma unroll 8
    for (int n = 0; n < nn; ++n) AffIO<AffT>::st(a + (size_t)n * PB, expf(logit(n) - mx) * inv + pos.v[n]);
}

template <typename AffT>
__global__ void __launch_bounds__(PAR_TX* PAR_TY) par_iter_kernel(const AffT* __restrict__ aff,
                                                                   const float* __restrict__ in, int in_pitch,
                                                                   float* __restrict__ out, int out_pitch,
                                                                   const __grid_constant__ ParGeom g,
                                                                   const __grid_constant__ ParChunks ch) {
    const int x = blockIdx.x * PAR_TX + threadIdx.x;
    const int y = blockIdx.y * PAR_TY + threadIdx.y;
    const int c = blockIdx.z;
    const int H = g.H, W = g.W;
    if (x >= W || y >= H) return;
    const int nn = g.nn;
    const int cnt = ch.cnt[c];
    const size_t iplane = (size_t)H * in_pitch, oplane = (size_t)H * out_pitch;
    const AffT* a = aff + aff_index(H, g.wb, nn, ch.b[c], 0, y, x);
    const float* m = in + (size_t)ch.ch0[c] * iplane;
    float acc[PAR_KC];
#pragma unroll
    for (int k = 0; k < PAR_KC; ++k) acc[k] = 0.f;
    for (int n = 0; n < nn; ++n) {
        const int yy = min(max(y + g.dy[n], 0), H - 1), xx = min(max(x + g.dx[n], 0), W - 1);
        const float w = AffIO<AffT>::ld(a + (size_t)n * PB);
        const float* p = m + (size_t)yy * in_pitch + xx;
#pragma unroll
        for (int k = 0; k < PAR_KC; ++k)
            if (k < cnt) acc[k] = fmaf(w, __ldg(p + (size_t)k * iplane), acc[k]);
    }
    float* o = out + (size_t)ch.ch0[c] * oplane + (size_t)y * out_pitch + x;
#pragma unroll
    for (int k = 0; k < PAR_KC; ++k)
        if (k < cnt) o[(size_t)k * oplane] = acc[k];
}

// ---------------------------------------------------------------------------------------------------
// Fast path: the module's dilation set {1,2,4,8,12,24} (PAR.py / model_attn_aff_voc.py:92), shared-memory tiles.
// ---------------------------------------------------------------------------------------------------
constexpr int PT_W = 64;                 // tile width in pixels (16 threads x 4 pixels)
constexpr int PT_HALO = 24;              // largest dilation
constexpr int PT_SW = PT_W + 2 * PT_HALO;   // 112 floats per shared-memory row (448 B: 16 B aligned)
constexpr int PT_THREADS = 256;

struct TileArgs {
    int H, W, wb;            // wb: affinity blocks per image row
    int evict_first;         // the affinity does not fit in L2: stream it with an evict-first policy
    int in_pitch, out_pitch; // mask plane pitches of this step's source / destination
    int in_vec, out_vec;     // 128-bit global access allowed (pitch % 4 == 0 and base 16 B aligned)
};

__host__ __device__ constexpr int std_dil(int d) { return d == 0 ? 1 : d == 1 ? 2 : d == 2 ? 4 : d == 3 ? 8 : d == 4 ? 12 : 24; }

// Fill `planes` shared-memory planes [SH][PT_SW] with the (x0-24.., y0-24..) window of src planes, coordinates clamped
// to the image (= F.pad(mode='replicate'), PAR.py:44).  One warp per tile row: lanes 0..27 own the row's 28 16-byte
// vectors, so the only per-iteration index work is the row clamp (the old flat index cost ~25 instructions per vector).
template <int SH, int NT>
__device__ __forceinline__ void fill_tile(float* tile, const float* __restrict__ src, size_t plane, int pitch, int planes,
                                          int x0, int y0, int H, int W, int vec) {
    constexpr int NW = NT / 32;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    if (lane < PT_SW / 4) {
        const int gx = x0 - PT_HALO + 4 * lane;
        const bool fast = vec && gx >= 0 && gx + 3 < W;
        const int c0 = min(max(gx, 0), W - 1), c1 = min(max(gx + 1, 0), W - 1), c2 = min(max(gx + 2, 0), W - 1),
                  c3 = min(max(gx + 3, 0), W - 1);
        const int rows = planes * SH;
        for (int rk = wrp; rk < rows; rk += NW) {
            const int k = rk / SH, r = rk - k * SH;
            const int gy = min(max(y0 - PT_HALO + r, 0), H - 1);
            const float* rowp = src + (size_t)k * plane + (size_t)gy * pitch;
            float* dst = tile + (size_t)rk * PT_SW + 4 * lane;
            if (fast) {
                // interior vector: asynchronous 16-byte global->shared copy (no register staging, many in flight)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)),
                             "l"(rowp + gx) : "memory");
            } else {
                // edge vector (clamped columns, or an unaligned source): four asynchronous 4-byte copies — a register-staged
                // load + store here made every row iteration of an edge tile wait a full L2 round trip (14 % of the
                // kernel's stall samples, ncu r02)
                const uint32_t d32 = (uint32_t)__cvta_generic_to_shared(dst);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32), "l"(rowp + c0) : "memory");
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32 + 4), "l"(rowp + c1) : "memory");
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32 + 8), "l"(rowp + c2) : "memory");
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32 + 12), "l"(rowp + c3) : "memory");
            }
        }
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// Storage type of the mask tile in shared memory.  With fp16 AFFINITIES (the pipeline's approximate mode) the masks of one step
// are also held as fp16 in the tile: values in [0, 1] rounded once per step at the fill (<= 2^-12 absolute), fp32 accumulation,
// fp32 in global memory.  That halves the tile (53.8 KB at Kc = 3) so that THREE CTAs fit per SM, compiled for 80 registers with
// two prefetched weight sets instead of three: the step is bound by the latency of its LDS -> FFMA2 chains (ncu r02 stall
// samples: MIO throttle 24 %, long scoreboard 21 %, short scoreboard 15 %; time linear in Kc), so resident warps are what helps.
// Measured on B200 (16 x 512^2, one step, us at Kc = 2 / 3 / 6):
//     fp32 tiles, 2 CTAs / SM, 3 weight sets (round-2 kernel until this change)   93 / 129 / 254
//     fp16 tiles, 2 CTAs / SM, 3 weight sets                                       88 / 129 / 242   (bytes alone do not help)
//     fp32 tiles, 3 CTAs / SM where they fit (Kc <= 2), 2 weight sets              99 / 140 / 261
//     fp16 tiles, 3 CTAs / SM, 2 weight sets (default)                             87 / 118 / 222
//     fp16 tiles, 4 CTAs / SM (64 registers: 444 B of spills)                      94 / 163 / 298
// fp32 affinities keep fp32 tiles, 2 CTAs / SM and exact arithmetic.  -DWC_PAR_MASK16=0 -DWC_PAR_CTAS16=2 -DWC_PAR_LOOKAHEAD16=2
// restores the previous fp16-affinity kernel.
#ifndef WC_PAR_MASK16
#define WC_PAR_MASK16 1
#endif
template <typename AffT> struct MaskTile { using T = float; };
#if WC_PAR_MASK16
template <> struct MaskTile<__half> { using T = __half; };
#endif
__device__ __forceinline__ float4 ldm4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ldm4(const __half* p) {
    const uint2 u = *reinterpret_cast<const uint2*>(p);
    const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
    const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
    return make_float4(lo.x, lo.y, hi.x, hi.y);
}

// fp16 tile: the same window, through registers (cp.async cannot convert): four rows per warp in flight
template <int SH, int NT>
__device__ __forceinline__ void fill_tile(__half* tile, const float* __restrict__ src, size_t plane, int pitch, int planes,
                                          int x0, int y0, int H, int W, int vec) {
    constexpr int NW = NT / 32;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    if (lane < PT_SW / 4) {
        const int gx = x0 - PT_HALO + 4 * lane;
        const bool fast = vec && gx >= 0 && gx + 3 < W;
        const int c0 = min(max(gx, 0), W - 1), c1 = min(max(gx + 1, 0), W - 1), c2 = min(max(gx + 2, 0), W - 1),
                  c3 = min(max(gx + 3, 0), W - 1);
        const int rows = planes * SH;
        for (int rk0 = wrp; rk0 < rows; rk0 += 4 * NW) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int rk = rk0 + u * NW;
                if (rk < rows) {
                    const int k = rk / SH, r = rk - k * SH;
                    const int gy = min(max(y0 - PT_HALO + r, 0), H - 1);
                    const float* rowp = src + (size_t)k * plane + (size_t)gy * pitch;
                    if (fast) v[u] = __ldg(reinterpret_cast<const float4*>(rowp + gx));
                    else v[u] = make_float4(__ldg(rowp + c0), __ldg(rowp + c1), __ldg(rowp + c2), __ldg(rowp + c3));
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int rk = rk0 + u * NW;
                if (rk < rows) {
                    const __half2 lo = __floats2half2_rn(v[u].x, v[u].y), hi = __floats2half2_rn(v[u].z, v[u].w);
                    uint2 pk;
                    pk.x = *reinterpret_cast<const uint32_t*>(&lo);
                    pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                    *reinterpret_cast<uint2*>(tile + (size_t)rk * PT_SW + 4 * lane) = pk;
                }
            }
        }
    }
}

// Two fp32 FMAs in one instruction (FFMA2, sm_100): halves the issue slots and fma-pipe cycles of the 48 x Kc x 4 products
// a thread accumulates per pass.  The operand pairs are adjacent registers (LDS.128 results, accumulator pairs), so the
// mov.b64 packs below are register renames.
__device__ __forceinline__ void ffma2(float& a0, float& a1, float w0, float w1, float m0, float m1) {
    asm("{\n\t.reg .b64 ra, rw, rm;\n\tmov.b64 ra, {%0, %1};\n\tmov.b64 rw, {%2, %3};\n\tmov.b64 rm, {%4, %5};\n\t"
        "fma.rn.f32x2 ra, rw, rm, ra;\n\tmov.b64 {%0, %1}, ra;\n\t}"
        : "+f"(a0), "+f"(a1) : "f"(w0), "f"(w1), "f"(m0), "f"(m1));
}

// acc[k][i] += w * mask(k, row, x+i+shift): the four pixels of this thread against one neighbour
template <int KC>
__device__ __forceinline__ void fma4(float (&acc)[KC][4], int k, const float4& w, float a, float b, float c, float d) {
    acc[k][0] = fmaf(w.x, a, acc[k][0]);
    acc[k][1] = fmaf(w.y, b, acc[k][1]);
    acc[k][2] = fmaf(w.z, c, acc[k][2]);
    acc[k][3] = fmaf(w.w, d, acc[k][3]);
}
template <int KC>
__device__ __forceinline__ void fma4p(float (&acc)[KC][4], int k, const float4& w, float a, float b, float c, float d) {
    ffma2(acc[k][0], acc[k][1], w.x, w.y, a, b);
    ffma2(acc[k][2], acc[k][3], w.z, w.w, c, d);
}

// One dilation DD: 8 neighbours x KC channels for 4 consecutive pixels.  sc -> centre pixel 0 of channel 0 in the tile.
template <typename AffT, int KC, int SH, int DD, typename MT = float>
__device__ __forceinline__ void accumulate_dilation(float (&acc)[KC][4], const MT* sc,
                                                    const typename AffIO<AffT>::Raw (&wr)[8]) {
    float4 w[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = AffIO<AffT>::unpack(wr[j]);
#pragma unroll
    for (int k = 0; k < KC; ++k) {
#pragma unroll
        for (int r = -1; r <= 1; ++r) {
            const MT* rowp = sc + k * (SH * PT_SW) + r * DD * PT_SW;
            const int jl = (r == -1) ? 0 : (r == 0) ? 3 : 5;      // neighbour index of (r, -1)
            const int jr = (r == -1) ? 2 : (r == 0) ? 4 : 7;      // neighbour index of (r, +1)
            if (DD % 4 == 0) {
                const float4 L = ldm4(rowp - DD);
                const float4 R = ldm4(rowp + DD);
                fma4p<KC>(acc, k, w[jl], L.x, L.y, L.z, L.w);
                fma4p<KC>(acc, k, w[jr], R.x, R.y, R.z, R.w);
                if (r != 0) {
                    const float4 C = ldm4(rowp);
                    fma4p<KC>(acc, k, w[jl + 1], C.x, C.y, C.z, C.w);
                }
            } else {
                const float4 L = ldm4(rowp - 4);
                const float4 C = ldm4(rowp);
                const float4 R = ldm4(rowp + 4);
                if (DD == 1) {      // shifted by one: the value pairs straddle register pairs -> scalar FMAs
                    fma4<KC>(acc, k, w[jl], L.w, C.x, C.y, C.z);
                    fma4<KC>(acc, k, w[jr], C.y, C.z, C.w, R.x);
                } else {   // DD == 2
                    fma4p<KC>(acc, k, w[jl], L.z, L.w, C.x, C.y);
                    fma4p<KC>(acc, k, w[jr], C.z, C.w, R.x, R.y);
                }
                if (r != 0) fma4p<KC>(acc, k, w[jl + 1], C.x, C.y, C.z, C.w);
            }
        }
    }
}

// the 8 weights of dilation d for this thread's 4 pixels: immediate offsets (d*8+j)*64 elements from the block base
template <typename AffT>
__device__ __forceinline__ void load_weights(typename AffIO<AffT>::Raw (&w)[8], const AffT* ap, int d, uint64_t pol) {
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = AffIO<AffT>::ld4(ap + (d * 8 + j) * PB, pol);
}

constexpr int PT_KC = 3;                 // channels per chunk on the tile path: 3 x 35 KB planes -> two CTAs per SM
constexpr int PT_TH = 32;

template <typename AffT, int KC, int TH>
__device__ __forceinline__ void par_iter_tile_body(typename MaskTile<AffT>::T* par_tile, const AffT* __restrict__ aff,
                                                   const float* __restrict__ in,
                                                   float* __restrict__ out, const TileArgs& a, const ParChunks& ch) {
    using MT = typename MaskTile<AffT>::T;
    constexpr int SH = TH + 2 * PT_HALO;
    const int c = blockIdx.z;
    const int x0 = blockIdx.x * PT_W, y0 = blockIdx.y * TH;
    const int H = a.H, W = a.W;
    const size_t iplane = (size_t)H * a.in_pitch, oplane = (size_t)H * a.out_pitch;
    const size_t arow = (size_t)a.wb * (48 * PB);          // affinity elements per image row
    const int ch0 = ch.ch0[c];
    const uint64_t pol = a.evict_first ? l2_evict_first_policy() : l2_evict_normal_policy();
    const int tx4 = threadIdx.x & 15, tyb = threadIdx.x >> 4;
    const int x = x0 + 4 * tx4;
    const bool active = x < W;             // (the affinity blocks are padded to 64 pixels: loads beyond W stay in bounds)
    const AffT* abase = aff + ((size_t)ch.b[c] * H * a.wb + blockIdx.x) * (48 * PB) + 4 * tx4;
    // Software pipeline over (row pass, dilation): two (LOOKAHEAD 1) or three (LOOKAHEAD 2) weight sets rotate, each refilled as
    // soon as its dilation has been accumulated — the refills at the end of a pass already fetch the NEXT row's first
    // dilations, and the very first ones are requested before the tile fill, so no pass starts with an exposed HBM round trip.
    // (fp16 weights: sets of 16 registers, two by default so that three CTAs fit per SM; fp32: two sets of 32)
    constexpr bool THREE = AffIO<AffT>::LOOKAHEAD == 2;
    typename AffIO<AffT>::Raw wa[8], wb[8], wc[THREE ? 8 : 1];
    if (active && y0 + tyb < H) {
        const AffT* ap = abase + (size_t)(y0 + tyb) * arow;
        load_weights<AffT>(wa, ap, 0, pol);
        if constexpr (THREE) { load_weights<AffT>(wb, ap, 1, pol); load_weights<AffT>(wc, ap, 2, pol); }
    }
    fill_tile<SH, PT_THREADS>(par_tile, in + (size_t)ch0 * iplane, iplane, a.in_pitch, KC, x0, y0, H, W, a.in_vec);
    __syncthreads();
    if (!active) return;
    if constexpr (!THREE) {      // fp32 weights: the second set would spill across the fill loop
        if (y0 + tyb < H) load_weights<AffT>(wb, abase + (size_t)(y0 + tyb) * arow, 1, pol);
    }
#pragma unroll 1
    for (int pass = 0; pass < TH / 16; ++pass) {
        const int ty = tyb + 16 * pass, y = y0 + ty;
        if (y >= H) break;
        const AffT* ap = abase + (size_t)y * arow;
        const bool more = pass + 1 < TH / 16 && y + 16 < H;
        const AffT* apn = ap + 16 * arow;
        const MT* sc = par_tile + (ty + PT_HALO) * PT_SW + PT_HALO + 4 * tx4;
        float acc[KC][4];
#pragma unroll
        for (int k = 0; k < KC; ++k) acc[k][0] = acc[k][1] = acc[k][2] = acc[k][3] = 0.f;
        if constexpr (THREE) {
            accumulate_dilation<AffT, KC, SH, 1, MT>(acc, sc, wa);
            load_weights<AffT>(wa, ap, 3, pol);
            accumulate_dilation<AffT, KC, SH, 2, MT>(acc, sc, wb);
            load_weights<AffT>(wb, ap, 4, pol);
            accumulate_dilation<AffT, KC, SH, 4, MT>(acc, sc, wc);
            load_weights<AffT>(wc, ap, 5, pol);
            accumulate_dilation<AffT, KC, SH, 8, MT>(acc, sc, wa);
            if (more) load_weights<AffT>(wa, apn, 0, pol);
            accumulate_dilation<AffT, KC, SH, 12, MT>(acc, sc, wb);
            if (more) load_weights<AffT>(wb, apn, 1, pol);
            accumulate_dilation<AffT, KC, SH, 24, MT>(acc, sc, wc);
            if (more) load_weights<AffT>(wc, apn, 2, pol);
        } else {
            accumulate_dilation<AffT, KC, SH, 1, MT>(acc, sc, wa);
            load_weights<AffT>(wa, ap, 2, pol);
            accumulate_dilation<AffT, KC, SH, 2, MT>(acc, sc, wb);
            load_weights<AffT>(wb, ap, 3, pol);
            accumulate_dilation<AffT, KC, SH, 4, MT>(acc, sc, wa);
            load_weights<AffT>(wa, ap, 4, pol);
            accumulate_dilation<AffT, KC, SH, 8, MT>(acc, sc, wb);
            load_weights<AffT>(wb, ap, 5, pol);
            accumulate_dilation<AffT, KC, SH, 12, MT>(acc, sc, wa);
            if (more) load_weights<AffT>(wa, apn, 0, pol);
            accumulate_dilation<AffT, KC, SH, 24, MT>(acc, sc, wb);
            if (more) load_weights<AffT>(wb, apn, 1, pol);
        }
        float* op = out + (size_t)ch0 * oplane + (size_t)y * a.out_pitch + x;
#pragma unroll
        for (int k = 0; k < KC; ++k) {
            if (a.out_vec && x + 3 < W) {
                *reinterpret_cast<float4*>(op + (size_t)k * oplane) = make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (x + i < W) op[(size_t)k * oplane + i] = acc[k][i];
            }
        }
    }
}

// One launch per propagation step: blockIdx.z = chunk (image, first channel, 1..PT_KC channels); the channel count is
// CTA-uniform, each count has its own fully unrolled body.
template <typename AffT>
#ifndef WC_PAR_CTAS16
#define WC_PAR_CTAS16 3          // resident CTAs per SM the fp16-affinity step is compiled for (3: <= 85 registers)
#endif
__global__ void __launch_bounds__(PT_THREADS, (sizeof(AffT) == 2 ? WC_PAR_CTAS16 : 2)) par_iter_tile_kernel(const AffT* __restrict__ aff,
                                                                       const float* __restrict__ in,
                                                                       float* __restrict__ out,
                                                                       const __grid_constant__ TileArgs a,
                                                                       const __grid_constant__ ParChunks ch) {
    extern __shared__ __align__(16) unsigned char par_tile_raw[];
    typename MaskTile<AffT>::T* par_tile = reinterpret_cast<typename MaskTile<AffT>::T*>(par_tile_raw);
    switch (ch.cnt[blockIdx.z]) {
        case 1: par_iter_tile_body<AffT, 1, PT_TH>(par_tile, aff, in, out, a, ch); break;
        case 2: par_iter_tile_body<AffT, 2, PT_TH>(par_tile, aff, in, out, a, ch); break;
        default: par_iter_tile_body<AffT, 3, PT_TH>(par_tile, aff, in, out, a, ch); break;
    }
}

// ---------------------------------------------------------------------------------------------------
// Pipelined propagation step (fp16 affinity): persistent CTAs, one per SM, walk 64x16-pixel tiles.
//   warp 8       weight producer: the 8 weights x 64 pixels of one (tile row, dilation) are ONE contiguous KB in the blocked
//                layout; 16 lanes each issue a cp.async.bulk (TMA unit, no registers, no LSU issue slots) into a 3-stage
//                ring of [16 rows][8][64] blocks, completion counted on the stage's mbarrier
//   warps 9-15   tile producers: the NEXT tile's mask planes + 24-pixel halo (double-buffered) with cp.async, completion
//                signalled through cp.async.mbarrier.arrive
//   warps 0-7    consumers: thread = 4 pixels of one row; per dilation 8 LDS.64 weights + the aligned LDS.128 mask vectors,
//                FFMA2 accumulation; they never wait on HBM latency themselves (the register-prefetch kernel above spent
//                35 % of its stall samples on exactly that, ncu r02) and the tile fill of tile i+1 runs under tile i.
// Shared memory: 2 x 3 planes x 28 KB + 3 x 16 KB ring + barriers = 216 KB.
// ---------------------------------------------------------------------------------------------------
constexpr int PP_TH = 16;
constexpr int PP_SH = PP_TH + 2 * PT_HALO;                       // 64
constexpr int PP_PLANE = PP_SH * PT_SW;                          // floats per plane
constexpr int PP_NST = 3;
constexpr int PP_STAGE_HALVES = PP_TH * 8 * PB;                  // fp16 elements per ring stage (16 KB)
constexpr int PP_CONSUMERS = 256;
constexpr int PP_FILL_WARPS = 7;
constexpr int PP_THREADS = PP_CONSUMERS + 32 + 32 * PP_FILL_WARPS;
constexpr size_t PP_SMEM = (size_t)2 * PT_KC * PP_PLANE * sizeof(float) + (size_t)PP_NST * PP_STAGE_HALVES * sizeof(__half) + 256 + 128;

__device__ __forceinline__ uint32_t pp_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pp_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pp_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void pp_mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(pp_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pp_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pp_smem_u32(bar)), "r"(bytes) : "memory");
}
// A protocol bug must fault, never hang the GPU: trap after ~2 s of spinning.
__device__ __forceinline__ void pp_mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = pp_smem_u32(bar);
    uint32_t done = 0;
    long long t0 = 0;
    int spins = 0;
    while (true) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) break;
        if (++spins == 64) t0 = clock64();
        if (spins > 64 && clock64() - t0 > 4000000000LL) {
            printf("wc: PAR pipeline mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x);
            __trap();
        }
    }
}
__device__ __forceinline__ void pp_bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(pp_smem_u32(dst)), "l"(src), "r"(bytes), "r"(pp_smem_u32(bar)), "l"(pol) : "memory");
}

struct PipeArgs {
    int tiles_x, tiles_y, n_tiles;
    int dbg;       // timing experiments only (WC_PAR_DBG): bit 0 = no weight copies, bit 1 = no tile fill (results are garbage)
};

template <int KC, int DD>
__device__ __forceinline__ void pp_dilation(float (&acc)[KC][4], const float* sc, const __half* wrow) {
    AffIO<__half>::Raw wr[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) wr[j] = *reinterpret_cast<const uint2*>(wrow + j * PB);
    accumulate_dilation<__half, KC, PP_SH, DD>(acc, sc, wr);
}

template <int KC>
__device__ __forceinline__ void pp_consume_tile(const float* planes, const __half* ring, uint64_t* w_full, uint64_t* w_empty,
                                                int& g, int tx4, int ty, int lane, float (&acc)[KC][4]) {
    const float* sc = planes + (ty + PT_HALO) * PT_SW + PT_HALO + 4 * tx4;
#pragma unroll
    for (int k = 0; k < KC; ++k) acc[k][0] = acc[k][1] = acc[k][2] = acc[k][3] = 0.f;
#pragma unroll
    for (int d = 0; d < 6; ++d) {
        const int s = g % PP_NST;
        pp_mbar_wait(&w_full[s], (g / PP_NST) & 1);
        const __half* wrow = ring + (size_t)s * PP_STAGE_HALVES + ty * (8 * PB) + 4 * tx4;
        if (d == 0) pp_dilation<KC, 1>(acc, sc, wrow);
        else if (d == 1) pp_dilation<KC, 2>(acc, sc, wrow);
        else if (d == 2) pp_dilation<KC, 4>(acc, sc, wrow);
        else if (d == 3) pp_dilation<KC, 8>(acc, sc, wrow);
        else if (d == 4) pp_dilation<KC, 12>(acc, sc, wrow);
        else pp_dilation<KC, 24>(acc, sc, wrow);
        __syncwarp();
        if (lane == 0) pp_mbar_arrive(&w_empty[s]);
        ++g;
    }
}

__global__ void __launch_bounds__(PP_THREADS, 1) par_iter_pipe_kernel(const __half* __restrict__ aff, const float* __restrict__ in,
                                                                       float* __restrict__ out, const __grid_constant__ TileArgs a,
                                                                       const __grid_constant__ ParChunks ch,
                                                                       const __grid_constant__ PipeArgs pa) {
    extern __shared__ __align__(128) uint8_t pp_smem[];
    float* planes = reinterpret_cast<float*>(pp_smem);                                                   // [2][PT_KC][PP_PLANE]
    __half* ring = reinterpret_cast<__half*>(pp_smem + (size_t)2 * PT_KC * PP_PLANE * sizeof(float));      // [PP_NST][16][8][64]
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(ring) + (size_t)PP_NST * PP_STAGE_HALVES * sizeof(__half));
    uint64_t* w_full = bars;                    // [PP_NST]  weights of a (tile, dilation) have landed
    uint64_t* w_empty = w_full + PP_NST;        // [PP_NST]  ... have been read by all 8 consumer warps
    uint64_t* t_full = w_empty + PP_NST;        // [2]       mask planes of a tile are in shared memory
    uint64_t* t_empty = t_full + 2;             // [2]       ... have been read by all 8 consumer warps
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int H = a.H, W = a.W;
    if (threadIdx.x == 0) {
        for (int s = 0; s < PP_NST; ++s) { pp_mbar_init(&w_full[s], 1); pp_mbar_init(&w_empty[s], 8); }
        for (int b = 0; b < 2; ++b) { pp_mbar_init(&t_full[b], 32 * PP_FILL_WARPS); pp_mbar_init(&t_empty[b], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int per_chunk = pa.tiles_x * pa.tiles_y;
    const size_t iplane = (size_t)H * a.in_pitch, oplane = (size_t)H * a.out_pitch;
    const size_t arow = (size_t)a.wb * (48 * PB);

    if (warp == 8) {
        // ---- weight producer
        const uint64_t pol = a.evict_first ? l2_evict_first_policy() : l2_evict_normal_policy();
        int g = 0;
        for (int w = blockIdx.x; w < pa.n_tiles; w += gridDim.x) {
            const int c = w / per_chunk, r = w - c * per_chunk, tyi = r / pa.tiles_x, txi = r - tyi * pa.tiles_x;
            const int y0 = tyi * PP_TH, rows = min(PP_TH, H - y0);
            const __half* base = aff + ((size_t)ch.b[c] * H * a.wb + txi) * (48 * PB);
            for (int d = 0; d < 6; ++d, ++g) {
                const int s = g % PP_NST;
                if (lane == 0) {
                    pp_mbar_wait(&w_empty[s], ((g / PP_NST) & 1) ^ 1);
                    if (pa.dbg & 1) pp_mbar_arrive(&w_full[s]);
                    else pp_mbar_expect_tx(&w_full[s], (uint32_t)rows * 8 * PB * sizeof(__half));
                }
                __syncwarp();
                if (lane < rows && !(pa.dbg & 1))
                    pp_bulk_load(ring + (size_t)s * PP_STAGE_HALVES + lane * (8 * PB), base + (size_t)(y0 + lane) * arow + d * (8 * PB),
                                 8 * PB * sizeof(__half), &w_full[s], pol);
            }
        }
    } else if (warp > 8) {
        // ---- tile producers (PP_FILL_WARPS warps: two could keep only ~10 B/clk of cp.async in flight): planes of tile i into
        //      buffer i & 1
        int i = 0;
        const int pw = warp - 9;
        for (int w = blockIdx.x; w < pa.n_tiles; w += gridDim.x, ++i) {
            const int c = w / per_chunk, r = w - c * per_chunk, tyi = r / pa.tiles_x, txi = r - tyi * pa.tiles_x;
            const int x0 = txi * PT_W, y0 = tyi * PP_TH, buf = i & 1, kc = ch.cnt[c];
            pp_mbar_wait(&t_empty[buf], ((i >> 1) & 1) ^ 1);
            float* tile = planes + (size_t)buf * PT_KC * PP_PLANE;
            const float* src = in + (size_t)ch.ch0[c] * iplane;
            if (lane < PT_SW / 4 && !(pa.dbg & 2)) {
                const int gx = x0 - PT_HALO + 4 * lane;
                const bool fast = a.in_vec && gx >= 0 && gx + 3 < W;
                const int c0 = min(max(gx, 0), W - 1), c1 = min(max(gx + 1, 0), W - 1), c2 = min(max(gx + 2, 0), W - 1),
                          c3 = min(max(gx + 3, 0), W - 1);
                const int rows = kc * PP_SH;
                for (int rk = pw; rk < rows; rk += PP_FILL_WARPS) {
                    const int k = rk / PP_SH, rr = rk - k * PP_SH;
                    const int gy = min(max(y0 - PT_HALO + rr, 0), H - 1);
                    const float* rowp = src + (size_t)k * iplane + (size_t)gy * a.in_pitch;
                    const uint32_t d32 = pp_smem_u32(tile + (size_t)k * PP_PLANE + (size_t)rr * PT_SW + 4 * lane);
                    if (fast) {
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d32), "l"(rowp + gx) : "memory");
                    } else {
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32), "l"(rowp + c0) : "memory");
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32 + 4), "l"(rowp + c1) : "memory");
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32 + 8), "l"(rowp + c2) : "memory");
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32 + 12), "l"(rowp + c3) : "memory");
                    }
                }
            }
            // arrives on t_full[buf] once every cp.async this thread has issued so far has completed (every lane arrives)
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(pp_smem_u32(&t_full[buf])) : "memory");
        }
    } else {
        // ---- consumers
        const int tx4 = threadIdx.x & 15, ty = threadIdx.x >> 4;
        int g = 0, i = 0;
        for (int w = blockIdx.x; w < pa.n_tiles; w += gridDim.x, ++i) {
            const int c = w / per_chunk, r = w - c * per_chunk, tyi = r / pa.tiles_x, txi = r - tyi * pa.tiles_x;
            const int x = txi * PT_W + 4 * tx4, y = tyi * PP_TH + ty, buf = i & 1, kc = ch.cnt[c];
            const float* tile = planes + (size_t)buf * PT_KC * PP_PLANE;
            pp_mbar_wait(&t_full[buf], (i >> 1) & 1);
            float acc[PT_KC][4];
            if (kc == 1) {
                float a1[1][4];
                pp_consume_tile<1>(tile, ring, w_full, w_empty, g, tx4, ty, lane, a1);
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[0][e] = a1[0][e];
            } else if (kc == 2) {
                float a2[2][4];
                pp_consume_tile<2>(tile, ring, w_full, w_empty, g, tx4, ty, lane, a2);
#pragma unroll
                for (int e = 0; e < 4; ++e) { acc[0][e] = a2[0][e]; acc[1][e] = a2[1][e]; }
            } else {
                pp_consume_tile<3>(tile, ring, w_full, w_empty, g, tx4, ty, lane, acc);
            }
            __syncwarp();
            if (lane == 0) pp_mbar_arrive(&t_empty[buf]);             // the planes may be overwritten by tile i+2
            if (x < W && y < H) {
                float* op = out + (size_t)ch.ch0[c] * oplane + (size_t)y * a.out_pitch + x;
#pragma unroll
                for (int k = 0; k < PT_KC; ++k) {
                    if (k < kc) {
                        if (a.out_vec && x + 3 < W) {
                            *reinterpret_cast<float4*>(op + (size_t)k * oplane) = make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (x + e < W) op[(size_t)k * oplane + e] = acc[k][e];
                        }
                    }
                }
            }
        }
    }
}

// Affinity build on a 64x24 tile: the 3 image planes + halo live in shared memory (96.8 KB, 2 CTAs per SM); 192 threads,
// each walks 8 pixels (lanes = consecutive x: conflict-free scalar LDS), TWO passes over the 48 compile-time neighbour
// offsets: (1) sum and sum of squares of d = v - centre per channel -> unbiased variance (shifted two-moment form: exact
// zero on flat regions, no cancellation against the pixel's own level), (2) the 48 logits, kept in registers, with the
// softmax scale folded into one ex2 per neighbour.
constexpr int PA_TH = 24;
constexpr int PA_THREADS = 192;
constexpr int PA_SH = PA_TH + 2 * PT_HALO;

template <typename AffT>
__global__ void __launch_bounds__(PA_THREADS, 2) par_aff_tile_kernel(const float* __restrict__ img, AffT* __restrict__ aff,
                                                                      int H, int W, int wb, int vec,
                                                                      const __grid_constant__ ParPos pos) {
    extern __shared__ __align__(16) float par_tile[];
    constexpr int PLANE = PA_SH * PT_SW;
    const int b = blockIdx.z;
    const int x0 = blockIdx.x * PT_W, y0 = blockIdx.y * PA_TH;
    const size_t hw = (size_t)H * W;
    fill_tile<PA_SH, PA_THREADS>(par_tile, img + (size_t)b * 3 * hw, hw, W, 3, x0, y0, H, W, vec);
    __syncthreads();
    const int tx = threadIdx.x & 63, ty0 = threadIdx.x >> 6;
    const int x = x0 + tx;
    if (x >= W) return;
    constexpr int by[8] = {-1, -1, -1, 0, 0, 1, 1, 1};
    constexpr int bx[8] = {-1, 0, 1, -1, 1, -1, 0, 1};
    constexpr float LOG2E = 1.4426950408889634f;
#pragma unroll 1
    for (int i = 0; i < PA_TH / 3; ++i) {
        const int ty = ty0 + 3 * i, y = y0 + ty;
        if (y >= H) break;
        const float* sc = par_tile + (ty + PT_HALO) * PT_SW + PT_HALO + tx;
        const float c0 = sc[0], c1 = sc[PLANE], c2 = sc[2 * PLANE];
        // pass 1: moments of d = v - centre over the 48 neighbours; var = (sum d^2 - (sum d)^2 / 48) / 47 (torch.std, PAR.py:77)
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f;
#pragma unroll
        for (int n = 0; n < 48; ++n) {
            const int o = by[n & 7] * std_dil(n >> 3) * PT_SW + bx[n & 7] * std_dil(n >> 3);
            const float d0 = sc[o] - c0, d1 = sc[PLANE + o] - c1, d2 = sc[2 * PLANE + o] - c2;
            s0 += d0; s1 += d1; s2 += d2;
            q0 = fmaf(d0, d0, q0); q1 = fmaf(d1, d1, q1); q2 = fmaf(d2, d2, q2);
        }
        const float v0 = fmaxf(fmaf(-s0 * (1.f / 48.f), s0, q0), 0.f), v1 = fmaxf(fmaf(-s1 * (1.f / 48.f), s1, q1), 0.f),
                    v2 = fmaxf(fmaf(-s2 * (1.f / 48.f), s2, q2), 0.f);
        const float e0 = sqrtf(v0 * (1.f / 47.f)) + 1e-8f, e1 = sqrtf(v1 * (1.f / 47.f)) + 1e-8f,
                    e2 = sqrtf(v2 * (1.f / 47.f)) + 1e-8f;
        // pass 2: logits = -mean_c((|v-c|/(std+1e-8)/w1)^2) (PAR.py:80-81) in the log2 domain, kept in registers; the two
        // divisions per term are one multiplication by r = 1/(std+1e-8)/w1 (1-2 ulp from the reference's order)
        const float r0 = 1.f / e0 / 0.3f, r1 = 1.f / e1 / 0.3f, r2 = 1.f / e2 / 0.3f;
        float l[48];
        float mx = -INFINITY;
#pragma unroll
        for (int n = 0; n < 48; ++n) {
            const int o = by[n & 7] * std_dil(n >> 3) * PT_SW + bx[n & 7] * std_dil(n >> 3);
            const float t0 = (sc[o] - c0) * r0, t1 = (sc[PLANE + o] - c1) * r1, t2 = (sc[2 * PLANE + o] - c2) * r2;
            l[n] = (t0 * t0 + t1 * t1 + t2 * t2) * (-LOG2E / 3.f);
            mx = fmaxf(mx, l[n]);
        }
        float sum = 0.f;
#pragma unroll
        for (int n = 0; n < 48; ++n) { l[n] = ex2_approx(l[n] - mx); sum += l[n]; }
        const float inv = 1.f / sum;
        AffT* ap = aff + (((size_t)b * H + y) * wb + blockIdx.x) * (48 * PB) + tx;
#pragma unroll
        for (int n = 0; n < 48; ++n) AffIO<AffT>::st(ap + n * PB, l[n] * inv + pos.v[n]);   // PAR.py:86
    }
}


// ---------------------------------------------------------------------------------------------------
constexpr int AL_MAX_IMG = 64;               // images per launch (channel offsets travel by value)
struct ArgmaxTab { int ch0[AL_MAX_IMG + 1]; };
// blockIdx.y = image of the launch's group; 4 pixels per thread when the plane size allows 16-byte accesses
__global__ void argmax_lut_kernel(const float* __restrict__ masks, const int64_t* __restrict__ lut,
                                  int64_t* __restrict__ labels, const __grid_constant__ ArgmaxTab tab, int b0, size_t hw) {
    const int bi = blockIdx.y;
    const int ch0 = tab.ch0[bi], kc = tab.ch0[bi + 1] - ch0;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= hw) return;
    const float* m = masks + (size_t)ch0 * hw + i;
    float best = m[0];
    int arg = 0;
    for (int k = 1; k < kc; ++k) {
        const float v = m[(size_t)k * hw];
        if (v > best || (v != v && best == best)) { best = v; arg = k; }   // first max wins; NaN propagates like torch
    }
    labels[(size_t)(b0 + bi) * hw + i] = lut[ch0 + arg];
}

// fp16 propagation kernel: 0 = register-prefetch tile kernel (default), 1 = the TMA-pipelined persistent kernel (WC_PAR_PIPE=1).
// Measured on B200 (16 x 512^2, Kc = 3 / 2, one step): register-prefetch 129 / 93 us; pipelined 175 / 136 us — its tile fill
// (86 KB of 16-byte cp.async per 64x16 tile from 7 warps) is not hidden behind the 8 consumer warps' 96 us of LDS + FFMA2
// work, and with the fill removed altogether it only ties (129 us).  Kept selectable as the measured alternative.
static int g_par_pipe = -1;

static int launch_iter_pipe(const __half* aff, const float* src, float* dst, const TileArgs& ta, const ParChunks& chunks, int n,
                            cudaStream_t st) {
    WC_SET_MAX_SMEM(par_iter_pipe_kernel, PP_SMEM);
    PipeArgs pa;
    pa.tiles_x = ceil_div(ta.W, PT_W); pa.tiles_y = ceil_div(ta.H, PP_TH); pa.n_tiles = pa.tiles_x * pa.tiles_y * n;
    { const char* e = getenv("WC_PAR_DBG"); pa.dbg = e ? atoi(e) : 0; }
    int dev = 0, sms = 148;
    WC_CUDA(cudaGetDevice(&dev));
    WC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = pa.n_tiles < sms ? pa.n_tiles : sms;
    par_iter_pipe_kernel<<<grid, PP_THREADS, PP_SMEM, st>>>(aff, src, dst, ta, chunks, pa);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

template <typename AffT>
static int launch_iter_tile(const AffT* aff, const float* src, float* dst, const TileArgs& ta, const ParChunks& chunks,
                            int n, int max_cnt, cudaStream_t st) {
    if (sizeof(AffT) == 2) {
        if (g_par_pipe < 0) { const char* e = getenv("WC_PAR_PIPE"); g_par_pipe = (e && e[0] == '1') ? 1 : 0; }
        if (g_par_pipe) return launch_iter_pipe(reinterpret_cast<const __half*>(aff), src, dst, ta, chunks, n, st);
    }
    constexpr int PLANE = (PT_TH + 2 * PT_HALO) * PT_SW * (int)sizeof(typename MaskTile<AffT>::T);
    WC_SET_MAX_SMEM(par_iter_tile_kernel<AffT>, PT_KC * PLANE);
    dim3 grid(ceil_div(ta.W, PT_W), ceil_div(ta.H, PT_TH), n);
    par_iter_tile_kernel<AffT><<<grid, PT_THREADS, max_cnt * PLANE, st>>>(aff, src, dst, ta, chunks);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// num_iter propagation steps.  masks / out have row pitch W; the ping-pong planes in `scratch` have pitch g.wp.
template <typename AffT>
static int propagate(const AffT* aff, const float* masks, float* out, float* scratch, const int* ch_off, int B,
                     const ParGeom& g, bool fast, int num_iter, cudaStream_t st) {
    const int sum_kc = ch_off[B];
    const size_t hw = (size_t)g.H * g.W, hwp = (size_t)g.H * g.wp;
    if (num_iter == 0) {
        if (out != masks)
            WC_CUDA(cudaMemcpyAsync(out, masks, sizeof(float) * sum_kc * hw, cudaMemcpyDeviceToDevice, st));
        return WC_OK;
    }
    // chunk list: the channels of one image in groups of <= kc_max (tile path 3, generic path 4); an image with more
    // channels streams its affinity once per group
    struct Item { short b, ch0, cnt; };
    std::vector<Item> items;
    const int kc_max = fast ? PT_KC : PAR_KC;
    for (int b = 0; b < B; ++b) {
        const int kc = ch_off[b + 1] - ch_off[b], groups = ceil_div(kc, kc_max);
        int k = ch_off[b];
        for (int gi = 0; gi < groups; ++gi) {          // balanced groups: 4 -> 2+2, 5 -> 3+2, 7 -> 3+2+2
            const int cnt = kc / groups + (gi < kc % groups ? 1 : 0);
            items.push_back(Item{(short)b, (short)k, (short)cnt});
            k += cnt;
        }
    }
    float* ping = scratch;
    float* pong = scratch + (size_t)sum_kc * hwp;
    for (int it = 0; it < num_iter; ++it) {
        const bool first = it == 0, last = it == num_iter - 1;
        const float* src = first ? masks : ((it & 1) ? ping : pong);
        float* dst = last ? out : ((it & 1) ? pong : ping);
        TileArgs ta;
        ta.H = g.H; ta.W = g.W; ta.wb = g.wb;
        // 126 MB of L2: an affinity tensor that fits stays resident across the iterations (no evict-first hint then)
        ta.evict_first = (double)B * g.H * g.wb * PB * g.nn * sizeof(AffT) > 80e6 ? 1 : 0;
        ta.in_pitch = first ? g.W : g.wp;
        ta.out_pitch = last ? g.W : g.wp;
        ta.in_vec = (ta.in_pitch % 4 == 0) && aligned16(src);
        ta.out_vec = (ta.out_pitch % 4 == 0) && aligned16(dst);
        const std::vector<Item>& v = items;
        for (size_t off = 0; off < v.size(); off += PAR_MAX_CHUNKS) {
            const int n = (int)min((size_t)PAR_MAX_CHUNKS, v.size() - off);
            ParChunks chunks;
            int chs = 0, max_cnt = 1, imgs = 0;
            for (int i = 0; i < n; ++i) {
                chunks.b[i] = v[off + i].b; chunks.ch0[i] = v[off + i].ch0; chunks.cnt[i] = v[off + i].cnt;
                chs += v[off + i].cnt;
                max_cnt = max(max_cnt, (int)v[off + i].cnt);
                if (i == 0 || v[off + i].b != v[off + i - 1].b) ++imgs;
            }
            // algorithmic bytes: every image's affinity ONCE per step (an image with more than 3 channels is
            // walked by several chunks; their re-reads are not credited) + read and write of every mask channel
            ProfScope prof(PK_PAR_ITER, (double)hw * ((double)sizeof(AffT) * g.nn * imgs + 8.0 * chs), st);
            if (fast) {
                int rc = launch_iter_tile<AffT>(aff, src, dst, ta, chunks, n, max_cnt, st);
                if (rc) return rc;
            } else {
                dim3 block(PAR_TX, PAR_TY), grid(ceil_div(g.W, PAR_TX), ceil_div(g.H, PAR_TY), n);
                par_iter_kernel<AffT><<<grid, block, 0, st>>>(aff, src, ta.in_pitch, dst, ta.out_pitch, g, chunks);
                WC_LAUNCH_CHECK();
            }
        }
    }
    return WC_OK;
}

template <typename AffT>
static int affinity(const float* img, AffT* aff, int B, const ParGeom& g, const ParPos& pos, bool fast, cudaStream_t st) {
    ProfScope prof(PK_PAR_AFF, (double)B * g.H * g.W * (12.0 + (double)sizeof(AffT) * g.nn), st);
    if (fast) {
        constexpr int SMEM = 3 * PA_SH * PT_SW * (int)sizeof(float);
        WC_SET_MAX_SMEM(par_aff_tile_kernel<AffT>, SMEM);
        dim3 grid(ceil_div(g.W, PT_W), ceil_div(g.H, PA_TH), B);
        const int vec = (g.W % 4 == 0) && aligned16(img);
        par_aff_tile_kernel<AffT><<<grid, PA_THREADS, SMEM, st>>>(img, aff, g.H, g.W, g.wb, vec, pos);
    } else {
        dim3 block(32, 8), grid(ceil_div(g.W, 32), ceil_div(g.H, 8), B);
        par_aff_kernel<AffT><<<grid, block, 0, st>>>(img, aff, g, pos);
    }
    WC_LAUNCH_CHECK();
    return WC_OK;
}

static int check_ragged(const int* ch_off, int B) {
    if (!ch_off || B < 1 || B > 32767 || ch_off[0] != 0) return -1;
    for (int b = 0; b < B; ++b) if (ch_off[b + 1] <= ch_off[b]) return -1;
    if (ch_off[B] > 32767) return -1;
    return 0;
}

}  // namespace wc

using namespace wc;

extern "C" int wc_par_aff_pitch(int W) { return par_aff_pitch(W); }

extern "C" size_t wc_par_workspace_bytes(int B, int sum_kc, int H, int W, int n_dil) {
    const size_t hw = (size_t)H * W, hwp = (size_t)H * par_pitch(W), hwa = (size_t)H * par_aff_pitch(W);
    return sizeof(float) * (hw * (size_t)B * 3 + hwa * (size_t)B * 8 * n_dil + hwp * 2 * (size_t)sum_kc) + 1024;
}

extern "C" int wc_par_affinity(const float* img, void* aff, int aff_f16, int B, int H, int W, const int* dil, int n_dil,
                               void* stream) {
    WC_CHECK_ARG(img && aff && B > 0 && B <= 65535 && H > 0 && W > 0 && dil);
    ParGeom g; ParPos pos;
    WC_CHECK_ARG(make_geom(g, &pos, H, W, dil, n_dil) == 0);
    const bool fast = is_standard_dilations(dil, n_dil) && aligned16(aff);
    if (aff_f16) return affinity<__half>(img, reinterpret_cast<__half*>(aff), B, g, pos, fast, as_stream(stream));
    return affinity<float>(img, reinterpret_cast<float*>(aff), B, g, pos, fast, as_stream(stream));
}

extern "C" int wc_par_propagate(const void* aff, int aff_f16, const float* masks, float* out, float* scratch,
                                const int* ch_off, int B, int H, int W, const int* dil, int n_dil, int num_iter,
                                void* stream) {
    WC_CHECK_ARG(aff && masks && out && scratch && H > 0 && W > 0 && num_iter >= 0);
    WC_CHECK_ARG(check_ragged(ch_off, B) == 0);
    ParGeom g;
    WC_CHECK_ARG(make_geom(g, nullptr, H, W, dil, n_dil) == 0);
    const bool fast = is_standard_dilations(dil, n_dil) && aligned16(aff) && aligned16(scratch);
    if (aff_f16)
        return propagate<__half>(reinterpret_cast<const __half*>(aff), masks, out, scratch, ch_off, B, g, fast, num_iter,
                                 as_stream(stream));
    return propagate<float>(reinterpret_cast<const float*>(aff), masks, out, scratch, ch_off, B, g, fast, num_iter,
                            as_stream(stream));
}

extern "C" int wc_par_fwd(const float* imgs, int Hi, int Wi, const float* masks, float* out, const int* ch_off,
                          int B, int H, int W, const int* dil, int n_dil, int num_iter, int aff_f16, void* workspace,
                          size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(imgs && masks && out && workspace && Hi > 0 && Wi > 0 && H > 0 && W > 0 && num_iter >= 0);
    WC_CHECK_ARG(check_ragged(ch_off, B) == 0);
    if (workspace_bytes < wc_par_workspace_bytes(B, ch_off[B], H, W, n_dil))
        return fail(WC_ERR_WORKSPACE, "%s: workspace too small (%ld bytes)", __func__, (long)workspace_bytes);
    cudaStream_t st = as_stream(stream);
    const size_t hw = (size_t)H * W, hwp = (size_t)H * par_pitch(W);
    float* ws = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    float* img_r = ws;
    float* scratch = img_r + (((size_t)B * 3 * hw + 63) & ~(size_t)63);
    void* aff = scratch + 2 * (size_t)ch_off[B] * hwp;
    const float* img = imgs;
    if (Hi != H || Wi != W) {
        dim3 block(32, 8), grid(ceil_div(W, 32), ceil_div(H, 8), B * 3);
        WC_CHECK_ARG(B * 3 <= 65535);
        ProfScope prof(PK_PAR_AFF, 0.0, st);
        par_resize_kernel<<<grid, block, 0, st>>>(imgs, img_r, B * 3, Hi, Wi, H, W);
        WC_LAUNCH_CHECK();
        img = img_r;
    }
    int rc = wc_par_affinity(img, aff, aff_f16, B, H, W, dil, n_dil, stream);
    if (rc) return rc;
    return wc_par_propagate(aff, aff_f16, masks, out, scratch, ch_off, B, H, W, dil, n_dil, num_iter, stream);
}

extern "C" int wc_argmax_lut(const float* masks, const int64_t* lut, int64_t* labels, const int* ch_off, int B,
                             int H, int W, void* stream) {
    WC_CHECK_ARG(masks && lut && labels && H > 0 && W > 0);
    WC_CHECK_ARG(check_ragged(ch_off, B) == 0);
    const size_t hw = (size_t)H * W;
    ProfScope prof(PK_CAM_UPSAMPLE, 0.0, as_stream(stream));
    for (int b0 = 0; b0 < B; b0 += AL_MAX_IMG) {       // one launch per group of up to 64 images (was: one per image)
        const int nb = (B - b0 < AL_MAX_IMG) ? B - b0 : AL_MAX_IMG;
        ArgmaxTab tab;
        for (int i = 0; i <= nb; ++i) tab.ch0[i] = ch_off[b0 + i];
        for (int i = nb + 1; i <= AL_MAX_IMG; ++i) tab.ch0[i] = tab.ch0[nb];
        dim3 grid((unsigned)ceil_div(hw, (size_t)256), (unsigned)nb);
        argmax_lut_kernel<<<grid, 256, 0, as_stream(stream)>>>(masks, lut, labels, tab, b0, hw);
        WC_LAUNCH_CHECK();
    }
    return WC_OK;
}
