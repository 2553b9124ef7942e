// PAR (pixel-adaptive refinement) for sm_100a.  Reference: WeCLIP_model/PAR.py:26-92.
//
// Memory-bound stencil family.  Data layout in HBM:
//   img   [B,3,H,W]   fp32 planar (resized copy lives in the workspace when Hi,Wi != H,W)
//   aff   [B,NN,H,W]  fp32 planar, NN = 8*n_dil (48): one coalesced stream per neighbour
//   masks [sumKc,H,W] fp32 planar, ragged over images; two ping-pong buffers in the workspace
// Kernels:
//   par_resize_kernel   bilinear align_corners=True (PAR.py:67)
//   par_aff_kernel      per pixel: 3-pass over the NN x 3 neighbour values (mean, centred sum of squares,
//                       logits) -> softmax + w2*softmax(pos)   (PAR.py:70-86).  Algorithmic bytes/pixel: 12 + 4*NN.
//   par_iter_kernel     one propagation step (PAR.py:88-90): streams NN affinities per pixel with
//                       L1::no_allocate loads so that L1 keeps the (re-used) mask neighbourhood.
//                       Algorithmic bytes/pixel/iteration: 4*NN + 2*4*Kc.
#include "common.cuh"

#include <math.h>

namespace wc {

constexpr int PAR_MAX_NN = 8 * WC_PAR_MAX_DIL;
constexpr int PAR_TX = 32;   // one warp = 32 consecutive pixels of a row
constexpr int PAR_TY = 16;   // 512 threads; large tiles keep the 24-px neighbourhood L1-resident
constexpr int PAR_KC = 4;    // mask channels per thread
constexpr int PAR_MAX_CHUNKS = 240;

struct ParGeom {
    int H, W, nn;
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

static int make_geom(ParGeom& g, ParPos* pos, int H, int W, const int* dil, int n_dil) {
    if (n_dil < 1 || n_dil > WC_PAR_MAX_DIL) return -1;
    static const int by[8] = {-1, -1, -1, 0, 0, 1, 1, 1};   // PAR.py:10-24 raster order without the centre
    static const int bx[8] = {-1, 0, 1, -1, 1, -1, 0, 1};
    g.H = H; g.W = W; g.nn = 8 * n_dil;
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

__device__ __forceinline__ float ldg_stream(const float* p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

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
template <int NN>   // NN > 0: compile-time neighbour count (48); NN == 0: runtime g.nn (<= PAR_MAX_NN)
__global__ void __launch_bounds__(256) par_aff_kernel(const float* __restrict__ img, float* __restrict__ aff,
                                                      const __grid_constant__ ParGeom g,
                                                      const __grid_constant__ ParPos pos) {
    const int x = blockIdx.x * 32 + threadIdx.x;
    const int y = blockIdx.y * 8 + threadIdx.y;
    const int b = blockIdx.z;
    const int H = g.H, W = g.W;
    if (x >= W || y >= H) return;
    const int nn = NN > 0 ? NN : g.nn;
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
    // pass 4: softmax denominator; pass 5: write softmax + w2*softmax(pos) (PAR.py:86).  Logits are
    // recomputed instead of kept (48 live registers spill); the re-reads are L1 hits.
    float sum = 0.f;
#pragma unroll 8
    for (int n = 0; n < nn; ++n) sum += expf(logit(n) - mx);
    const float inv = 1.f / sum;
    float* a = aff + (size_t)b * nn * hw + (size_t)y * W + x;
#pragma unroll 8
    for (int n = 0; n < nn; ++n) a[(size_t)n * hw] = expf(logit(n) - mx) * inv + pos.v[n];
}

// ---------------------------------------------------------------------------------------------------
template <int NN>
__global__ void __launch_bounds__(PAR_TX* PAR_TY) par_iter_kernel(const float* __restrict__ aff,
                                                                   const float* __restrict__ in,
                                                                   float* __restrict__ out,
                                                                   const __grid_constant__ ParGeom g,
                                                                   const __grid_constant__ ParChunks ch) {
    const int x = blockIdx.x * PAR_TX + threadIdx.x;
    const int y = blockIdx.y * PAR_TY + threadIdx.y;
    const int c = blockIdx.z;
    const int H = g.H, W = g.W;
    if (x >= W || y >= H) return;
    const int nn = NN > 0 ? NN : g.nn;
    const int cnt = ch.cnt[c];
    const size_t hw = (size_t)H * W;
    const float* a = aff + (size_t)ch.b[c] * nn * hw + (size_t)y * W + x;
    const float* m = in + (size_t)ch.ch0[c] * hw;
    float acc[PAR_KC];
#pragma unroll
    for (int k = 0; k < PAR_KC; ++k) acc[k] = 0.f;
#pragma unroll
    for (int n = 0; n < (NN > 0 ? NN : PAR_MAX_NN); ++n) {
        if (NN == 0 && n >= nn) break;
        const int yy = min(max(y + g.dy[n], 0), H - 1), xx = min(max(x + g.dx[n], 0), W - 1);
        const float w = ldg_stream(a + (size_t)n * hw);
        const float* p = m + (size_t)yy * W + xx;
#pragma unroll
        for (int k = 0; k < PAR_KC; ++k)
            if (k < cnt) acc[k] = fmaf(w, __ldg(p + (size_t)k * hw), acc[k]);
    }
    float* o = out + (size_t)ch.ch0[c] * hw + (size_t)y * W + x;
#pragma unroll
    for (int k = 0; k < PAR_KC; ++k)
        if (k < cnt) o[(size_t)k * hw] = acc[k];
}

// ---------------------------------------------------------------------------------------------------
__global__ void argmax_lut_kernel(const float* __restrict__ masks, const int64_t* __restrict__ lut,
                                  int64_t* __restrict__ labels, int ch0, int kc, int b, size_t hw) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= hw) return;
    const float* m = masks + (size_t)ch0 * hw + i;
    float best = m[0];
    int arg = 0;
    for (int k = 1; k < kc; ++k) {
        const float v = m[(size_t)k * hw];
        if (v > best || (v != v && best == best)) { best = v; arg = k; }   // first max wins; NaN propagates like torch
    }
    labels[(size_t)b * hw + i] = lut[ch0 + arg];
}

static int propagate(const float* aff, const float* masks, float* out, float* scratch, const int* ch_off, int B,
                     const ParGeom& g, int num_iter, cudaStream_t st) {
    const int sum_kc = ch_off[B];
    const size_t hw = (size_t)g.H * g.W;
    if (num_iter == 0) {
        if (out != masks)
            WC_CUDA(cudaMemcpyAsync(out, masks, sizeof(float) * sum_kc * hw, cudaMemcpyDeviceToDevice, st));
        return WC_OK;
    }
    // chunk table
    ParChunks chunks;
    int nchunks_total = 0;
    for (int b = 0; b < B; ++b) nchunks_total += ceil_div(ch_off[b + 1] - ch_off[b], PAR_KC);
    float* ping = scratch;
    float* pong = scratch + (size_t)sum_kc * hw;
    dim3 block(PAR_TX, PAR_TY);
    for (int it = 0; it < num_iter; ++it) {
        const float* src = (it == 0) ? masks : ((it & 1) ? ping : pong);
        float* dst = (it == num_iter - 1) ? out : ((it & 1) ? pong : ping);
        int b = 0, k = ch_off[0];
        int done = 0;
        while (done < nchunks_total) {
            int n = 0;
            while (n < PAR_MAX_CHUNKS && b < B) {
                if (k >= ch_off[b + 1]) { ++b; if (b < B) k = ch_off[b]; continue; }
                chunks.b[n] = (short)b; chunks.ch0[n] = (short)k;
                chunks.cnt[n] = (short)min(PAR_KC, ch_off[b + 1] - k);
                k += chunks.cnt[n]; ++n;
            }
            if (n == 0) break;
            dim3 grid(ceil_div(g.W, PAR_TX), ceil_div(g.H, PAR_TY), n);
            {
                int chs = 0;
                for (int i = 0; i < n; ++i) chs += chunks.cnt[i];
                ProfScope prof(PK_PAR_ITER, (double)hw * (4.0 * g.nn * n + 8.0 * chs), st);
                if (g.nn == 48) par_iter_kernel<48><<<grid, block, 0, st>>>(aff, src, dst, g, chunks);
                else par_iter_kernel<0><<<grid, block, 0, st>>>(aff, src, dst, g, chunks);
            }
            WC_LAUNCH_CHECK();
            done += n;
        }
    }
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

extern "C" size_t wc_par_workspace_bytes(int B, int sum_kc, int H, int W, int n_dil) {
    const size_t hw = (size_t)H * W;
    return sizeof(float) * hw * ((size_t)B * 3 + (size_t)B * 8 * n_dil + 2 * (size_t)sum_kc) + 256;
}

extern "C" int wc_par_affinity(const float* img, float* aff, int B, int H, int W, const int* dil, int n_dil,
                               void* stream) {
    WC_CHECK_ARG(img && aff && B > 0 && H > 0 && W > 0 && dil);
    ParGeom g; ParPos pos;
    WC_CHECK_ARG(make_geom(g, &pos, H, W, dil, n_dil) == 0);
    dim3 block(32, 8), grid(ceil_div(W, 32), ceil_div(H, 8), B);
    ProfScope prof(PK_PAR_AFF, (double)B * H * W * (12.0 + 4.0 * g.nn), as_stream(stream));
    if (g.nn == 48) par_aff_kernel<48><<<grid, block, 0, as_stream(stream)>>>(img, aff, g, pos);
    else par_aff_kernel<0><<<grid, block, 0, as_stream(stream)>>>(img, aff, g, pos);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_par_propagate(const float* aff, const float* masks, float* out, float* scratch,
                                const int* ch_off, int B, int H, int W, const int* dil, int n_dil, int num_iter,
                                void* stream) {
    WC_CHECK_ARG(aff && masks && out && scratch && H > 0 && W > 0 && num_iter >= 0);
    WC_CHECK_ARG(check_ragged(ch_off, B) == 0);
    ParGeom g;
    WC_CHECK_ARG(make_geom(g, nullptr, H, W, dil, n_dil) == 0);
    return propagate(aff, masks, out, scratch, ch_off, B, g, num_iter, as_stream(stream));
}

extern "C" int wc_par_fwd(const float* imgs, int Hi, int Wi, const float* masks, float* out, const int* ch_off,
                          int B, int H, int W, const int* dil, int n_dil, int num_iter, void* workspace,
                          size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(imgs && masks && out && workspace && Hi > 0 && Wi > 0 && H > 0 && W > 0 && num_iter >= 0);
    WC_CHECK_ARG(check_ragged(ch_off, B) == 0);
    if (workspace_bytes < wc_par_workspace_bytes(B, ch_off[B], H, W, n_dil))
        return fail(WC_ERR_WORKSPACE, "%s: workspace too small (%ld bytes)", __func__, (long)workspace_bytes);
    cudaStream_t st = as_stream(stream);
    const size_t hw = (size_t)H * W;
    float* ws = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    float* img_r = ws;
    float* aff = img_r + (size_t)B * 3 * hw;
    float* scratch = aff + (size_t)B * 8 * n_dil * hw;
    const float* img = imgs;
    if (Hi != H || Wi != W) {
        dim3 block(32, 8), grid(ceil_div(W, 32), ceil_div(H, 8), B * 3);
        WC_CHECK_ARG(B * 3 <= 65535);
        ProfScope prof(PK_PAR_AFF, 0.0, st);
        par_resize_kernel<<<grid, block, 0, st>>>(imgs, img_r, B * 3, Hi, Wi, H, W);
        WC_LAUNCH_CHECK();
        img = img_r;
    }
    int rc = wc_par_affinity(img, aff, B, H, W, dil, n_dil, stream);
    if (rc) return rc;
    ParGeom g;
    WC_CHECK_ARG(make_geom(g, nullptr, H, W, dil, n_dil) == 0);
    return propagate(aff, masks, out, scratch, ch_off, B, g, num_iter, st);
}

extern "C" int wc_argmax_lut(const float* masks, const int64_t* lut, int64_t* labels, const int* ch_off, int B,
                             int H, int W, void* stream) {
    WC_CHECK_ARG(masks && lut && labels && H > 0 && W > 0);
    WC_CHECK_ARG(check_ragged(ch_off, B) == 0);
    const size_t hw = (size_t)H * W;
    ProfScope prof(PK_CAM_UPSAMPLE, 0.0, as_stream(stream));
    for (int b = 0; b < B; ++b) {
        argmax_lut_kernel<<<(unsigned)ceil_div(hw, (size_t)256), 256, 0, as_stream(stream)>>>(
            masks, lut, labels, ch_off[b], ch_off[b + 1] - ch_off[b], b, hw);
        WC_LAUNCH_CHECK();
    }
    return WC_OK;
}
