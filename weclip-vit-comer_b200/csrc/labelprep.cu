// Training-time label prep and the two losses that consume the hot path's outputs: the step right
// after WeCLIP.forward in scripts/dist_clip_voc.py:246-258.
//   get_mask_by_radius        scripts/dist_clip_voc.py:116-133   -> wc_radius_mask
//   cams_to_affinity_label    utils/camutils.py:226-247          -> wc_affinity_label
//   get_aff_loss              utils/losses.py:11-22              -> wc_aff_loss (targets given) /
//                                                                   wc_aff_loss_fused (targets derived on the fly from the
//                                                                   pseudo-labels: the [B,P,P] int64 tensor never exists)
//   get_seg_loss              utils/losses.py:24-33              -> wc_seg_loss
// All are HBM-bound passes (integer compares, one exp per logit); each loss also emits d loss / d input so the host side
// can hang them on torch autograd.  Sums are accumulated in fp64 (block partials -> one atomicAdd per CTA).
#include "common.cuh"

#include <atomic>

namespace wc {
namespace {

// torch's nearest-neighbour source index (F.interpolate(mode="nearest"), camutils.py:230): floor(dst * in/out) in fp32.
__device__ __forceinline__ int nearest_src(int dst, float scale, int in) { return min((int)floorf(dst * scale), in - 1); }

__global__ void radius_mask_kernel(uint8_t* __restrict__ mask, int h, int w, int radius) {
    const int P = h * w;
    const long n = (long)P * P;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int r = (int)(i / P), c = (int)(i % P);
        const int dy = r / w - c / w, dx = r % w - c % w;
        mask[i] = (abs(dy) <= radius && abs(dx) <= radius) ? 1 : 0;
    }
}

// Down-sampled pseudo-label of patch p of image b (labels are small non-negative ints or ignore_index: the reference's
// float32 round trip is exact on them).
__global__ void label_downsample_kernel(const int64_t* __restrict__ cam_label, int32_t* __restrict__ small, int B, int H,
                                        int W, int h, int w) {
    const int P = h * w;
    const float sy = (float)H / (float)h, sx = (float)W / (float)w;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B * P; i += gridDim.x * blockDim.x) {
        const int b = i / P, p = i % P;
        const int y = nearest_src(p / w, sy, H), x = nearest_src(p % w, sx, W);
        small[i] = (int32_t)cam_label[((size_t)b * H + y) * W + x];
    }
}

// camutils.py:235-245: equal labels -> 1, different -> 0; outside the radius mask, or either label == ignore -> ignore.
__device__ __forceinline__ int aff_target(int lr, int lc, int in_radius, int ignore) {
    if (!in_radius || lr == ignore || lc == ignore) return ignore;
    return lr == lc ? 1 : 0;
}

__global__ void affinity_label_kernel(const int32_t* __restrict__ small, const uint8_t* __restrict__ mask,
                                      int64_t* __restrict__ out, int B, int P, int ignore) {
    const long n = (long)B * P * P;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % P);
        const long br = i / P;
        const int r = (int)(br % P), b = (int)(br / P);
        const int m = mask ? mask[(size_t)r * P + c] : 1;
        out[i] = aff_target(small[b * P + r], small[b * P + c], m, ignore);
    }
}

struct AffAcc { double pos_sum, neg_sum; unsigned long long pos_n, neg_n; };

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ unsigned long long warp_sum_u(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void aff_block_commit(AffAcc* acc, double ps, double ns, unsigned long long pn,
                                                 unsigned long long nn) {
    __shared__ double s_ps[8], s_ns[8];
    __shared__ unsigned long long s_pn[8], s_nn[8];
    ps = warp_sum_d(ps); ns = warp_sum_d(ns); pn = warp_sum_u(pn); nn = warp_sum_u(nn);
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_ps[wid] = ps; s_ns[wid] = ns; s_pn[wid] = pn; s_nn[wid] = nn; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < (int)(blockDim.x >> 5); ++i) { ps += s_ps[i]; ns += s_ns[i]; pn += s_pn[i]; nn += s_nn[i]; }
        atomicAdd(&acc->pos_sum, ps); atomicAdd(&acc->neg_sum, ns);
        atomicAdd(&acc->pos_n, pn); atomicAdd(&acc->neg_n, nn);
    }
}

// Targets of VEC consecutive elements starting at flat index e of [B,P,P] (VEC == 4 needs P % 4 == 0: one row).
template <bool FUSED, int VEC>
__device__ __forceinline__ void aff_targets(int (&tv)[VEC], long e, const int64_t* __restrict__ t,
                                            const int32_t* __restrict__ small, int P, int w, int radius, int ignore) {
    if (FUSED) {
        const int c0 = (int)(e % P);
        const long br = e / P;
        const int r = (int)(br % P), b = (int)(br / P);
        const int lr = small[b * P + r], ry = r / w, rx = r % w;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const int c = c0 + j;
            const int inr = radius < 0 || (abs(ry - c / w) <= radius && abs(rx - c % w) <= radius);
            tv[j] = aff_target(lr, small[b * P + c], inr, ignore);
        }
    } else if (VEC == 4) {
        const longlong2 a = __ldg(reinterpret_cast<const longlong2*>(t + e));
        const longlong2 b = __ldg(reinterpret_cast<const longlong2*>(t + e) + 1);
        tv[0] = (int)a.x; tv[1] = (int)a.y; tv[2] = (int)b.x; tv[3] = (int)b.y;
    } else {
        tv[0] = (int)t[e];
    }
}

// Pass 1 of get_aff_loss: sum_{t==1}(1 - x), sum_{t==0} x and the two counts.  VEC elements per thread per trip (128-bit
// loads of the inputs when VEC == 4).  FUSED: targets come from the down-sampled labels + the radius, nothing else is read.
template <bool FUSED, int VEC>
__global__ void __launch_bounds__(256) aff_loss_reduce_kernel(const float* __restrict__ x, const int64_t* __restrict__ t,
                                                               const int32_t* __restrict__ small, int P, int w,
                                                               int radius, int ignore, long nv, AffAcc* acc) {
    double ps = 0.0, ns = 0.0;
    unsigned long long pn = 0, nn = 0;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
        float xv[VEC];
        if (VEC == 4) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
            xv[0] = v.x; xv[1 % VEC] = v.y; xv[2 % VEC] = v.z; xv[3 % VEC] = v.w;
        } else {
            xv[0] = __ldg(x + i);
        }
        int tv[VEC];
        aff_targets<FUSED, VEC>(tv, i * VEC, t, small, P, w, radius, ignore);
        float fp = 0.f, fn = 0.f;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (tv[j] == 1) { fp += 1.f - xv[j]; ++pn; }
            else if (tv[j] == 0) { fn += xv[j]; ++nn; }
        }
        ps += fp; ns += fn;
    }
    aff_block_commit(acc, ps, ns, pn, nn);
}

// loss = 0.5 * pos_sum / (n_pos + 1) + 0.5 * neg_sum / (n_neg + 1) (losses.py:13-22)
__global__ void aff_loss_final_kernel(const AffAcc* acc, float* loss, int64_t* counts) {
    const double pc = (double)acc->pos_n + 1.0, nc = (double)acc->neg_n + 1.0;
    loss[0] = (float)(0.5 * acc->pos_sum / pc + 0.5 * acc->neg_sum / nc);
    counts[0] = (int64_t)acc->pos_n + 1;
    counts[1] = (int64_t)acc->neg_n + 1;
}

// d loss / d x = -0.5/(n_pos+1) on positives, +0.5/(n_neg+1) on negatives, 0 on ignored pairs.
template <bool FUSED, int VEC>
__global__ void __launch_bounds__(256) aff_loss_grad_kernel(const int64_t* __restrict__ t, const int32_t* __restrict__ small,
                                                             int P, int w, int radius, int ignore, long nv,
                                                             const AffAcc* __restrict__ acc, float* __restrict__ grad) {
    const float gp = (float)(-0.5 / ((double)acc->pos_n + 1.0)), gn = (float)(0.5 / ((double)acc->neg_n + 1.0));
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
        int tv[VEC];
        aff_targets<FUSED, VEC>(tv, i * VEC, t, small, P, w, radius, ignore);
        float g[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) g[j] = tv[j] == 1 ? gp : tv[j] == 0 ? gn : 0.f;
        if (VEC == 4) reinterpret_cast<float4*>(grad)[i] = make_float4(g[0], g[1 % VEC], g[2 % VEC], g[3 % VEC]);
        else grad[i] = g[0];
    }
}

// ---- get_seg_loss ------------------------------------------------------------------------------------------------------
struct SegAcc { double bg_sum, fg_sum; unsigned long long bg_n, fg_n; };

// Pass 0: how many background (label 0) and foreground (label != 0, != ignore) pixels (the two CE means' denominators).
__global__ void __launch_bounds__(256) seg_count_kernel(const int64_t* __restrict__ label, long n, int C, int ignore, SegAcc* acc) {
    unsigned long long bn = 0, fn = 0;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long long l = label[i];
        // only labels the loss kernel sums over: 0 <= l < C (an out-of-range label that is not ignore_index is a device
        // assert in F.cross_entropy; here it is dropped from the numerator AND the denominator)
        if (l == 0) ++bn; else if (l != ignore && l > 0 && l < C) ++fn;
    }
    bn = warp_sum_u(bn); fn = warp_sum_u(fn);
    if ((threadIdx.x & 31) == 0) { if (bn) atomicAdd(&acc->bg_n, bn); if (fn) atomicAdd(&acc->fg_n, fn); }
}

// Pass 1: per pixel log-sum-exp over the C logits (ONE read of pred: running maximum + rescaled sum), nll = lse -
// pred[label]; block-reduced bg / fg sums; and (optionally) grad[b,c,y,x] = coef(label) * (softmax_c - [c == label]),
// coef = 0.5 / n_bg (label 0), 0.5 / n_fg (foreground), 0 (ignored).  A thread owns VEC consecutive pixels of one image
// (VEC == 4: 128-bit loads / stores per channel plane, H*W % 4 == 0); the gradient loop re-reads the CTA's own lines
// out of L1 / L2.
template <int VEC>
__global__ void __launch_bounds__(256) seg_loss_kernel(const float* __restrict__ pred, const int64_t* __restrict__ label,
                                                        int C, long HW, long nv, int ignore, SegAcc* acc,
                                                        float* __restrict__ grad) {
    const float cb = acc->bg_n ? (float)(0.5 / (double)acc->bg_n) : 0.f;
    const float cf = acc->fg_n ? (float)(0.5 / (double)acc->fg_n) : 0.f;
    constexpr float LOG2E = 1.4426950408889634f;
    double bs = 0.0, fs = 0.0;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
        const long e = i * VEC;
        const long b = e / HW, p = e % HW;
        const float* pp = pred + (size_t)b * C * HW + p;
        long long l[VEC];
        bool valid[VEC], any = false;
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            l[j] = label[e + j];
            valid[j] = l[j] != ignore && l[j] >= 0 && l[j] < C;
            any |= valid[j];
        }
        if (!any && !grad) continue;
        float mx[VEC], se[VEC], tl[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) { mx[j] = -INFINITY; se[j] = 0.f; tl[j] = 0.f; }
#pragma unroll 4
        for (int c = 0; c < C; ++c) {
            float v[VEC];
            if (VEC == 4) {
                const float4 q = __ldg(reinterpret_cast<const float4*>(pp + (size_t)c * HW));
                v[0] = q.x; v[1 % VEC] = q.y; v[2 % VEC] = q.z; v[3 % VEC] = q.w;
            } else {
                v[0] = __ldg(pp + (size_t)c * HW);
            }
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const float m = fmaxf(mx[j], v[j]);
                se[j] = se[j] * exp2f((mx[j] - m) * LOG2E) + exp2f((v[j] - m) * LOG2E);
                mx[j] = m;
                if (c == l[j]) tl[j] = v[j];
            }
        }
        float coef[VEC], inv[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (valid[j]) {
                const float nll = mx[j] + logf(se[j]) - tl[j];
                if (l[j] == 0) bs += nll; else fs += nll;
            }
            coef[j] = !valid[j] ? 0.f : (l[j] == 0 ? cb : cf);
            inv[j] = coef[j] / se[j];
        }
        if (grad) {
            float* gp = grad + (size_t)b * C * HW + p;
#pragma unroll 4
            for (int c = 0; c < C; ++c) {
                float v[VEC], g[VEC];
                if (VEC == 4) {
                    const float4 q = *reinterpret_cast<const float4*>(pp + (size_t)c * HW);
                    v[0] = q.x; v[1 % VEC] = q.y; v[2 % VEC] = q.z; v[3 % VEC] = q.w;
                } else {
                    v[0] = pp[(size_t)c * HW];
                }
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    g[j] = exp2f((v[j] - mx[j]) * LOG2E) * inv[j];
                    if (c == l[j]) g[j] -= coef[j];
                }
                if (VEC == 4) __stcs(reinterpret_cast<float4*>(gp + (size_t)c * HW), make_float4(g[0], g[1 % VEC], g[2 % VEC], g[3 % VEC]));
                else gp[(size_t)c * HW] = g[0];
            }
        }
    }
    bs = warp_sum_d(bs); fs = warp_sum_d(fs);
    __shared__ double s_b[8], s_f[8];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_b[wid] = bs; s_f[wid] = fs; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < (int)(blockDim.x >> 5); ++i) { bs += s_b[i]; fs += s_f[i]; }
        atomicAdd(&acc->bg_sum, bs); atomicAdd(&acc->fg_sum, fs);
    }
}

// (CE_bg + CE_fg) / 2, each a mean over its own pixels; an empty class of pixels gives nan like F.cross_entropy does.
__global__ void seg_loss_final_kernel(const SegAcc* acc, float* loss) {
    const float bg = (float)acc->bg_sum / (float)acc->bg_n, fg = (float)acc->fg_sum / (float)acc->fg_n;
    loss[0] = (bg + fg) * 0.5f;
}

inline int grid_for(long n, int threads, int per_sm = 8) {
    long g = (n + threads - 1) / threads;
    const long cap = 148L * per_sm;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace
}  // namespace wc

using namespace wc;

extern "C" int wc_radius_mask(int h, int w, int radius, uint8_t* mask, void* stream) {
    WC_CHECK_ARG(h > 0 && w > 0 && radius >= 0 && mask);
    const long n = (long)h * w * h * w;
    radius_mask_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(mask, h, w, radius);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" size_t wc_labelprep_workspace_bytes(int B, int H, int W) {
    const size_t P = (size_t)(H / 16) * (W / 16);
    return 64 + ((B * P * sizeof(int32_t) + 63) & ~(size_t)63);
}

extern "C" int wc_affinity_label(const int64_t* cam_label, int B, int H, int W, const uint8_t* radius_mask,
                                 int ignore_index, int64_t* aff_label, void* workspace, size_t workspace_bytes,
                                 void* stream) {
    WC_CHECK_ARG(cam_label && aff_label && workspace && B > 0 && H >= 16 && W >= 16);
    WC_CHECK_ARG(workspace_bytes >= wc_labelprep_workspace_bytes(B, H, W));
    const int h = H / 16, w = W / 16, P = h * w;
    int32_t* small = reinterpret_cast<int32_t*>(static_cast<char*>(workspace) + 64);
    cudaStream_t st = as_stream(stream);
    label_downsample_kernel<<<grid_for((long)B * P, 256), 256, 0, st>>>(cam_label, small, B, H, W, h, w);
    WC_LAUNCH_CHECK();
    affinity_label_kernel<<<grid_for((long)B * P * P, 256), 256, 0, st>>>(small, radius_mask, aff_label, B, P, ignore_index);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

template <bool FUSED>
static int aff_loss_launch(const float* x, const int64_t* t, const int32_t* small, int P, int w, int radius, int ignore,
                           long n, AffAcc* acc, float* loss, int64_t* counts, float* grad, cudaStream_t st) {
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool vec = (FUSED ? P % 4 == 0 : (n % 4 == 0 && al16(t))) && al16(x) && (!grad || al16(grad));
    const long nv = vec ? n / 4 : n;
    if (vec) aff_loss_reduce_kernel<FUSED, 4><<<grid_for(nv, 256), 256, 0, st>>>(x, t, small, P, w, radius, ignore, nv, acc);
    else     aff_loss_reduce_kernel<FUSED, 1><<<grid_for(nv, 256), 256, 0, st>>>(x, t, small, P, w, radius, ignore, nv, acc);
    WC_LAUNCH_CHECK();
    aff_loss_final_kernel<<<1, 1, 0, st>>>(acc, loss, counts);
    WC_LAUNCH_CHECK();
    if (grad) {
        if (vec) aff_loss_grad_kernel<FUSED, 4><<<grid_for(nv, 256), 256, 0, st>>>(t, small, P, w, radius, ignore, nv, acc, grad);
        else     aff_loss_grad_kernel<FUSED, 1><<<grid_for(nv, 256), 256, 0, st>>>(t, small, P, w, radius, ignore, nv, acc, grad);
        WC_LAUNCH_CHECK();
    }
    return WC_OK;
}

extern "C" int wc_aff_loss(const float* inputs, const int64_t* targets, long n, float* loss, int64_t* counts,
                           float* grad, void* workspace, size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(inputs && targets && loss && counts && workspace && workspace_bytes >= 64 && n > 0);
    cudaStream_t st = as_stream(stream);
    AffAcc* acc = static_cast<AffAcc*>(workspace);
    WC_CUDA(cudaMemsetAsync(acc, 0, sizeof(AffAcc), st));
    return aff_loss_launch<false>(inputs, targets, nullptr, 1, 1, -1, 255, n, acc, loss, counts, grad, st);
}

extern "C" int wc_aff_loss_fused(const float* attn_pred, const int64_t* cam_label, int B, int H, int W, int radius,
                                 int ignore_index, float* loss, int64_t* counts, float* grad, void* workspace,
                                 size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(attn_pred && cam_label && loss && counts && workspace && B > 0 && H >= 16 && W >= 16);
    WC_CHECK_ARG(workspace_bytes >= wc_labelprep_workspace_bytes(B, H, W));
    const int h = H / 16, w = W / 16, P = h * w;
    cudaStream_t st = as_stream(stream);
    AffAcc* acc = static_cast<AffAcc*>(workspace);
    int32_t* small = reinterpret_cast<int32_t*>(static_cast<char*>(workspace) + 64);
    WC_CUDA(cudaMemsetAsync(acc, 0, sizeof(AffAcc), st));
    label_downsample_kernel<<<grid_for((long)B * P, 256), 256, 0, st>>>(cam_label, small, B, H, W, h, w);
    WC_LAUNCH_CHECK();
    return aff_loss_launch<true>(attn_pred, nullptr, small, P, w, radius, ignore_index, (long)B * P * P, acc, loss, counts,
                                 grad, st);
}

// ---------------------------------------------------------------------------------------------------
// wc_aff_loss_fused_static: wc_aff_loss_fused for callers that cannot allocate (CUDA-graph capture, serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.  Its scratch (the 64-byte
// accumulator + the down-sampled patch labels) is this file-scope device array: every GPU that loads the library holds its
// 256 KB + 256 B until exit.  In host code the array's name is only a stand-in for the symbol, NOT a device address: the address
// is looked up once per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
// All calls share the array and it keeps the last call's contents: the accumulator is cleared by every call, the label part is
// fully rewritten for the call's B * P patches; two calls that overlap on the GPU (different streams) race.
// ---------------------------------------------------------------------------------------------------
constexpr long AFF_STATIC_MAX_PATCHES = 65536;                       // B * (H/16) * (W/16), e.g. 64 crops of 512 x 512
constexpr size_t AFF_STATIC_BYTES = 256 + AFF_STATIC_MAX_PATCHES * sizeof(int32_t);
static __device__ __align__(256) unsigned char g_aff_static_scratch[AFF_STATIC_BYTES];
constexpr int AFF_STATIC_MAX_DEVICES = 64;
static std::atomic<void*> g_aff_static_ptr[AFF_STATIC_MAX_DEVICES];    // per device: the array's DEVICE address (0 = not looked up)

extern "C" long wc_aff_loss_static_max_patches(void) { return AFF_STATIC_MAX_PATCHES; }

extern "C" int wc_aff_loss_fused_static(const float* attn_pred, const int64_t* cam_label, int B, int H, int W, int radius,
                                        int ignore_index, float* loss, int64_t* counts, float* grad, void* stream) {
    WC_CHECK_ARG(attn_pred && cam_label && loss && counts && B > 0 && H >= 16 && W >= 16);
    const int h = H / 16, w = W / 16, P = h * w;
    if ((long)B * P > AFF_STATIC_MAX_PATCHES)
        return fail(WC_ERR_WORKSPACE, "%s: B * (H/16) * (W/16) = %ld patches exceed the static scratch of %ld; use wc_aff_loss_fused "
                    "with a caller-provided workspace", __func__, (long)B * P, AFF_STATIC_MAX_PATCHES);
    int dev = 0;
    WC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= AFF_STATIC_MAX_DEVICES) return fail(WC_ERR_UNSUPPORTED, "%s: device index %ld", __func__, dev);
    void* base = g_aff_static_ptr[dev].load(std::memory_order_acquire);
    if (!base) {      // set-up, once per device (racing first calls look up the same address)
        WC_CUDA(cudaGetSymbolAddress(&base, g_aff_static_scratch));
        if (!base) return fail(WC_ERR_CUDA, "%s: cudaGetSymbolAddress returned a null address", __func__);
        g_aff_static_ptr[dev].store(base, std::memory_order_release);
    }
    cudaStream_t st = as_stream(stream);
    AffAcc* acc = static_cast<AffAcc*>(base);
    int32_t* small = reinterpret_cast<int32_t*>(static_cast<char*>(base) + 256);
    WC_CUDA(cudaMemsetAsync(acc, 0, sizeof(AffAcc), st));
    label_downsample_kernel<<<grid_for((long)B * P, 256), 256, 0, st>>>(cam_label, small, B, H, W, h, w);
    WC_LAUNCH_CHECK();
    return aff_loss_launch<true>(attn_pred, nullptr, small, P, w, radius, ignore_index, (long)B * P * P, acc, loss, counts,
                                 grad, st);
}

extern "C" int wc_seg_loss(const float* pred, const int64_t* label, int B, int C, int H, int W, int ignore_index,
                           float* loss, float* grad, void* workspace, size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(pred && label && loss && workspace && workspace_bytes >= 64 && B > 0 && C > 0 && H > 0 && W > 0);
    cudaStream_t st = as_stream(stream);
    SegAcc* acc = static_cast<SegAcc*>(workspace);
    WC_CUDA(cudaMemsetAsync(acc, 0, sizeof(SegAcc), st));
    const long HW = (long)H * W, n = (long)B * HW;
    seg_count_kernel<<<grid_for(n, 256), 256, 0, st>>>(label, n, C, ignore_index, acc);
    WC_LAUNCH_CHECK();
    const bool vec = HW % 4 == 0 && (reinterpret_cast<uintptr_t>(pred) & 15) == 0 &&
                     (!grad || (reinterpret_cast<uintptr_t>(grad) & 15) == 0);
    if (vec) seg_loss_kernel<4><<<grid_for(n / 4, 256, 8), 256, 0, st>>>(pred, label, C, HW, n / 4, ignore_index, acc, grad);
    else     seg_loss_kernel<1><<<grid_for(n, 256, 8), 256, 0, st>>>(pred, label, C, HW, n, ignore_index, acc, grad);
    WC_LAUNCH_CHECK();
    seg_loss_final_kernel<<<1, 1, 0, st>>>(acc, loss);
    WC_LAUNCH_CHECK();
    return WC_OK;
}
