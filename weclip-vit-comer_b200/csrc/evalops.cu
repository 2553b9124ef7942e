// msc+flip fusion and scoring kernels for sm_100a (memory-bound, tiny next to the model forward).
//
// Replaces the ATen calls of the evaluation loop, test_msc_flip_voc.py:60-96 (COCO twin :61-96):
//   wc_resize_bilinear        F.interpolate(mode='bilinear', align_corners=False)        (:63, :80)
//   wc_flip_concat            torch.cat([x, x.flip(-1)], dim=0)                          (:68, :82)
//   wc_msc_accumulate         resize to the base seg size + flip-average + scale sum    (:74, :86-88, :90)
//   wc_upsample_argmax_hist   resize to the label size + argmax (:93-96) + utils/evaluate.py:9-15 _fast_hist
// Bilinear arithmetic follows ATen's upsample_bilinear2d: src = max(scale*(dst+0.5)-0.5, 0), i0 = (int)src,
// i1 = i0 + (i0 < in-1), l1 = src - i0, out = l0y*(l0x*v00 + l1x*v01) + l1y*(l0x*v10 + l1x*v11).
#include "common.cuh"

namespace wc {

struct Tap { int i0, i1; float l0, l1; };

__device__ __forceinline__ Tap make_tap(int dst, float scale, int in_size) {
    const float s = fmaxf(scale * ((float)dst + 0.5f) - 0.5f, 0.f);
    Tap t;
    t.i0 = min((int)s, in_size - 1);
    t.i1 = t.i0 + (t.i0 < in_size - 1 ? 1 : 0);
    t.l1 = s - (float)t.i0;
    t.l0 = 1.f - t.l1;
    return t;
}

__device__ __forceinline__ float bilerp(const float* __restrict__ p, int W, const Tap& ty, const Tap& tx) {
    return ty.l0 * (tx.l0 * p[(size_t)ty.i0 * W + tx.i0] + tx.l1 * p[(size_t)ty.i0 * W + tx.i1]) +
           ty.l1 * (tx.l0 * p[(size_t)ty.i1 * W + tx.i0] + tx.l1 * p[(size_t)ty.i1 * W + tx.i1]);
}

__global__ void __launch_bounds__(256) resize_bilinear_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                              int Hi, int Wi, int Ho, int Wo, float sy, float sx) {
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y, p = blockIdx.z;
    if (x >= Wo || y >= Ho) return;
    const Tap ty = make_tap(y, sy, Hi), tx = make_tap(x, sx, Wi);
    out[((size_t)p * Ho + y) * Wo + x] = bilerp(in + (size_t)p * Hi * Wi, Wi, ty, tx);
}

// out[0] = in, out[1] = in mirrored along x; in [planes,H,W] -> out [2,planes,H,W]
__global__ void __launch_bounds__(256) flip_concat_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                          int planes, int H, int W) {
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y, p = blockIdx.z;
    if (x >= W || y >= H) return;
    const float v = in[((size_t)p * H + y) * W + x];
    out[((size_t)p * H + y) * W + x] = v;
    out[((size_t)(planes + p) * H + y) * W + (W - 1 - x)] = v;
}

// acc[c,y,x] (+)= weight * 0.5 * (R(seg[0])[c,y,x] + R(seg[1])[c,y,w-1-x]),  R = bilinear resize hs x ws -> h x w
__global__ void __launch_bounds__(256) msc_accumulate_kernel(const float* __restrict__ seg2, float* __restrict__ acc,
                                                             int BC, int hs, int ws, int h, int w, float sy, float sx,
                                                             float weight, int first) {
    // c runs over B*C planes; the flipped half of seg2 [2,B,C,hs,ws] starts BC planes later
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y, c = blockIdx.z;
    const int C = BC;
    if (x >= w || y >= h) return;
    float a, b;
    if (hs == h && ws == w) {
        a = seg2[((size_t)c * h + y) * w + x];
        b = seg2[((size_t)(C + c) * h + y) * w + (w - 1 - x)];
    } else {
        const Tap ty = make_tap(y, sy, hs), tx = make_tap(x, sx, ws), txf = make_tap(w - 1 - x, sx, ws);
        a = bilerp(seg2 + (size_t)c * hs * ws, ws, ty, tx);
        b = bilerp(seg2 + (size_t)(C + c) * hs * ws, ws, ty, txf);
    }
    const float v = weight * ((a + b) / 2.f);
    float* o = acc + ((size_t)c * h + y) * w + x;
    *o = first ? v : (*o + v);
}

// pred = argmax_c R(logits)[c], first maximum wins; optional confusion-matrix update hist[C*gt+pred] += 1 for
// 0 <= gt < C (utils/evaluate.py:11-14).  Block-private shared histogram when C*C fits, else global atomics.
__global__ void __launch_bounds__(256) upsample_argmax_hist_kernel(const float* __restrict__ logits, int C, int h, int w,
                                                                   int Ho, int Wo, float sy, float sx,
                                                                   int32_t* __restrict__ pred, const int16_t* __restrict__ gt,
                                                                   unsigned long long* __restrict__ hist, int smem_hist) {
    extern __shared__ unsigned int sh[];
    const int tid = threadIdx.y * 32 + threadIdx.x;
    if (smem_hist) for (int i = tid; i < C * C; i += 256) sh[i] = 0u;
    if (smem_hist) __syncthreads();
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    logits += (size_t)blockIdx.z * C * h * w;
    if (pred) pred += (size_t)blockIdx.z * Ho * Wo;
    if (gt) gt += (size_t)blockIdx.z * Ho * Wo;
    if (x < Wo && y < Ho) {
        const Tap ty = make_tap(y, sy, h), tx = make_tap(x, sx, w);
        float best = bilerp(logits, w, ty, tx);
        int arg = 0;
        for (int c = 1; c < C; ++c) {
            const float v = bilerp(logits + (size_t)c * h * w, w, ty, tx);
            if (v > best || (v != v && best == best)) { best = v; arg = c; }
        }
        if (pred) pred[(size_t)y * Wo + x] = arg;
        if (gt && hist) {
            const int g = gt[(size_t)y * Wo + x];
            if (g >= 0 && g < C) {
                if (smem_hist) atomicAdd(&sh[g * C + arg], 1u);
                else atomicAdd(hist + (size_t)g * C + arg, 1ULL);
            }
        }
    }
    if (smem_hist) {
        __syncthreads();
        for (int i = tid; i < C * C; i += 256)
            if (sh[i]) atomicAdd(hist + i, (unsigned long long)sh[i]);
    }
}

// hist[C*gt+pred] += 1 over flat arrays (the cam-label histogram, test_msc_flip_voc.py:107)
__global__ void __launch_bounds__(256) confusion_hist_kernel(const int64_t* __restrict__ pred, const int16_t* __restrict__ gt,
                                                             long n, int C, unsigned long long* __restrict__ hist, int smem_hist) {
    if (!smem_hist) {   // C*C bins do not fit the static shared-memory window: global atomics
        for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long)gridDim.x * 256) {
            const int g = gt[i];
            const long p = pred[i];
            if (g >= 0 && g < C && p >= 0 && p < C) atomicAdd(hist + (size_t)g * C + p, 1ULL);
        }
        return;
    }
    // CTA-private histogram in shared memory (C*C <= 81*81 bins), flushed once: a few bins take almost every pixel, so
    // global atomics would serialise in L2
    extern __shared__ unsigned int bins[];
    const int nb = C * C;
    for (int b = threadIdx.x; b < nb; b += 256) bins[b] = 0u;
    __syncthreads();
    for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long)gridDim.x * 256) {
        const int g = gt[i];
        const long p = pred[i];
        if (g >= 0 && g < C && p >= 0 && p < C) atomicAdd(bins + g * C + (int)p, 1u);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += 256)
        if (bins[b]) atomicAdd(hist + b, (unsigned long long)bins[b]);
}

static inline float torch_scale(int in_size, int out_size, float scale_factor) {
    // ATen area_pixel_compute_scale(align_corners=false): 1/scale_factor when one was passed, else in/out
    return scale_factor > 0.f ? (float)(1.0 / (double)scale_factor) : (float)in_size / (float)out_size;
}

}  // namespace wc

using namespace wc;

extern "C" int wc_resize_bilinear(const float* in, float* out, int planes, int Hi, int Wi, int Ho, int Wo,
                                  float scale_factor_y, float scale_factor_x, void* stream) {
    WC_CHECK_ARG(in && out && planes > 0 && planes <= 65535 && Hi > 0 && Wi > 0 && Ho > 0 && Wo > 0);
    ProfScope prof(PK_EVAL, 0.0, as_stream(stream));
    dim3 block(32, 8), grid(ceil_div(Wo, 32), ceil_div(Ho, 8), planes);
    resize_bilinear_kernel<<<grid, block, 0, as_stream(stream)>>>(in, out, Hi, Wi, Ho, Wo,
                                                                  torch_scale(Hi, Ho, scale_factor_y),
                                                                  torch_scale(Wi, Wo, scale_factor_x));
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_flip_concat(const float* in, float* out, int planes, int H, int W, void* stream) {
    WC_CHECK_ARG(in && out && planes > 0 && planes <= 65535 && H > 0 && W > 0);
    ProfScope prof(PK_EVAL, 0.0, as_stream(stream));
    dim3 block(32, 8), grid(ceil_div(W, 32), ceil_div(H, 8), planes);
    flip_concat_kernel<<<grid, block, 0, as_stream(stream)>>>(in, out, planes, H, W);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_msc_accumulate(const float* seg2, float* acc, int B, int C_, int hs, int ws, int h, int w, float weight,
                                 int first, void* stream) {
    WC_CHECK_ARG(seg2 && acc && B > 0 && C_ > 0 && (long)B * C_ <= 65535 && hs > 0 && ws > 0 && h > 0 && w > 0);
    const int C = B * C_;
    ProfScope prof(PK_EVAL, 0.0, as_stream(stream));
    dim3 block(32, 8), grid(ceil_div(w, 32), ceil_div(h, 8), C);
    msc_accumulate_kernel<<<grid, block, 0, as_stream(stream)>>>(seg2, acc, C, hs, ws, h, w, torch_scale(hs, h, 0.f),
                                                                 torch_scale(ws, w, 0.f), weight, first);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_upsample_argmax_hist(const float* logits, int B, int C, int h, int w, int Ho, int Wo, int32_t* pred,
                                       const int16_t* gt, int64_t* hist, void* stream) {
    WC_CHECK_ARG(logits && B > 0 && B <= 65535 && C > 0 && h > 0 && w > 0 && Ho > 0 && Wo > 0 && (pred || (gt && hist)));
    const int smem_hist = (gt && hist && (size_t)C * C * 4 <= 40 * 1024) ? 1 : 0;
    ProfScope prof(PK_EVAL, 0.0, as_stream(stream));
    dim3 block(32, 8), grid(ceil_div(Wo, 32), ceil_div(Ho, 8), B);
    upsample_argmax_hist_kernel<<<grid, block, smem_hist ? (size_t)C * C * 4 : 0, as_stream(stream)>>>(
        logits, C, h, w, Ho, Wo, torch_scale(h, Ho, 0.f), torch_scale(w, Wo, 0.f), pred, gt,
        reinterpret_cast<unsigned long long*>(hist), smem_hist);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_confusion_hist(const int64_t* pred, const int16_t* gt, long n, int C, int64_t* hist, void* stream) {
    WC_CHECK_ARG(pred && gt && hist && n > 0 && C > 0);
    const int smem_hist = ((size_t)C * C * 4 <= 40 * 1024) ? 1 : 0;
    ProfScope prof(PK_EVAL, 0.0, as_stream(stream));
    const unsigned grid = (unsigned)min((long)148 * 4, ceil_div(n, 1024L));
    confusion_hist_kernel<<<grid, 256, smem_hist ? (size_t)C * C * 4 : 0, as_stream(stream)>>>(
        pred, gt, n, C, reinterpret_cast<unsigned long long*>(hist), smem_hist);
    WC_LAUNCH_CHECK();
    return WC_OK;
}
