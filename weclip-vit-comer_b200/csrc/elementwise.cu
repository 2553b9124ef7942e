// Memory-bound helper kernels of the encoder / decoder (sm_100a): LayerNorm, softmax + head-mean,
// patch gather (im2col), positional-embedding resize, class-token row.
#include "elementwise.cuh"

namespace wc {

// ---------------------------------------------------------------------------------------------------
// LayerNorm over the last dim (clip/model.py:177-183: fp32 statistics, eps 1e-5).  One warp per row,
// the row lives in registers (E <= 1024), 128-bit loads.  Writes h16 (GEMM operand) and/or fp32.
// ---------------------------------------------------------------------------------------------------
template <int VPL>   // float4 vectors per lane: E = 128 * VPL
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                        const float* __restrict__ b, h16* __restrict__ out_h,
                                                        float* __restrict__ out_f, h16* __restrict__ raw_h, long rows,
                                                        int E) {
    const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float4* xr = reinterpret_cast<const float4*>(x + row * E);
    float4 v[VPL];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        v[i] = xr[lane + 32 * i];
        s += v[i].x + v[i].y + v[i].z + v[i].w;
        if (raw_h) {   // fp16 copy of the un-normalised row (the previous block's output tap)
            h16x2 lo = __floats2half2_rn(v[i].x, v[i].y), hi = __floats2half2_rn(v[i].z, v[i].w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo);
            pk.y = *reinterpret_cast<uint32_t*>(&hi);
            reinterpret_cast<uint2*>(raw_h + row * E)[lane + 32 * i] = pk;
        }
    }
    const float mean = warp_sum(s) / (float)E;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const float a = v[i].x - mean, c = v[i].y - mean, d = v[i].z - mean, e = v[i].w - mean;
        q += a * a + c * c + d * d + e * e;
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)E + 1e-5f);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c4 = lane + 32 * i;
        const float4 ww = __ldg(reinterpret_cast<const float4*>(w) + c4);
        const float4 bb = __ldg(reinterpret_cast<const float4*>(b) + c4);
        float4 y;
        y.x = (v[i].x - mean) * rstd * ww.x + bb.x;
        y.y = (v[i].y - mean) * rstd * ww.y + bb.y;
        y.z = (v[i].z - mean) * rstd * ww.z + bb.z;
        y.w = (v[i].w - mean) * rstd * ww.w + bb.w;
        if (out_f) reinterpret_cast<float4*>(out_f + row * E)[c4] = y;
        if (out_h) {
            h16x2 lo = __floats2half2_rn(y.x, y.y), hi = __floats2half2_rn(y.z, y.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo);
            pk.y = *reinterpret_cast<uint32_t*>(&hi);
            reinterpret_cast<uint2*>(out_h + row * E)[c4] = pk;
        }
    }
}

int launch_layernorm(const float* x, const float* w, const float* b, h16* out_h, float* out_f, long rows,
                     int E, cudaStream_t st, h16* raw_h) {
    WC_CHECK_ARG(x && w && b && rows > 0 && (out_h || out_f));
    const unsigned grid = (unsigned)ceil_div(rows, 8L);
    ProfScope prof(PK_LAYERNORM, (double)rows * E * (4.0 + (out_h ? 2.0 : 0.0) + (out_f ? 4.0 : 0.0) + (raw_h ? 2.0 : 0.0)), st);
    switch (E) {
        case 256: layernorm_kernel<2><<<grid, 256, 0, st>>>(x, w, b, out_h, out_f, raw_h, rows, E); break;
        case 512: layernorm_kernel<4><<<grid, 256, 0, st>>>(x, w, b, out_h, out_f, raw_h, rows, E); break;
        case 768: layernorm_kernel<6><<<grid, 256, 0, st>>>(x, w, b, out_h, out_f, raw_h, rows, E); break;
        case 1024: layernorm_kernel<8><<<grid, 256, 0, st>>>(x, w, b, out_h, out_f, raw_h, rows, E); break;
        default: return fail(WC_ERR_UNSUPPORTED, "%s: LayerNorm width %ld not instantiated", __func__, (long)E);
    }
    WC_LAUNCH_CHECK();
    return WC_OK;
}

// ---------------------------------------------------------------------------------------------------
// Row softmax of the attention scores for all heads of one (image, query row) + mean over heads
// (clip/myAtt.py:57 softmax, :323-326 head average).  S fp32 [N*heads, L, Lp] -> P h16 same layout,
// avg fp32 [N, L, L] (optional).  One CTA per (row, image); the head-mean accumulates in shared memory so
// each output element is written once (4*L^2 bytes per image instead of 12x).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) softmax_headmean_kernel(const float* __restrict__ S, h16* __restrict__ P,
                                                               float* __restrict__ avg, int heads, int L, int Lp) {
    extern __shared__ float sm_avg[];   // [L]
    __shared__ float red[32];
    const int row = blockIdx.x, img = blockIdx.y;
    const int tid = threadIdx.x;
    if (avg) for (int c = tid; c < L; c += 256) sm_avg[c] = 0.f;
    const float inv_h = 1.f / (float)heads;
    for (int h = 0; h < heads; ++h) {
        const size_t base = ((size_t)(img * heads + h) * L + row) * Lp;
        const float* s = S + base;
        float mx = -INFINITY;
        for (int c = tid; c < L; c += 256) mx = fmaxf(mx, s[c]);
        mx = block_max(mx, red);
        float sum = 0.f;
        for (int c = tid; c < L; c += 256) sum += __expf(s[c] - mx);
        sum = block_sum(sum, red);
        const float inv = 1.f / sum;
        h16* p = P + base;
        for (int c = tid; c < L; c += 256) {
            const float v = __expf(s[c] - mx) * inv;
            p[c] = __float2half_rn(v);
            if (avg) sm_avg[c] += v * inv_h;   // each thread owns its columns: no race
        }
    }
    if (avg) {
        float* a = avg + ((size_t)img * L + row) * L;
        for (int c = tid; c < L; c += 256) a[c] = sm_avg[c];
    }
}

int launch_softmax_headmean(const float* S, h16* P, float* avg, int n_img, int heads, int L, int Lp, cudaStream_t st) {
    WC_CHECK_ARG(S && P && n_img > 0 && heads > 0 && L > 0 && Lp >= L && n_img <= 65535);
    dim3 grid(L, n_img);
    ProfScope prof(PK_SOFTMAX, (double)n_img * L * L * (6.0 * heads + (avg ? 4.0 : 0.0)), st);
    softmax_headmean_kernel<<<grid, 256, sizeof(float) * L, st>>>(S, P, avg, heads, L, Lp);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

// ---------------------------------------------------------------------------------------------------
// Patch gather: img fp32 [N,3,H,W] -> patches h16 [N*P, 768], k = c*256 + ky*16 + kx (the flattening of
// conv1.weight [768,3,16,16], clip/model.py:246,267).  The conv floors: only h*16 x w*16 pixels are used.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) im2col16_kernel(const float* __restrict__ img, h16* __restrict__ out,
                                                       int H, int W, int h, int w) {
    // one CTA per (patch row py, image): 16 image rows x (w*16) pixels x 3 channels
    const int py = blockIdx.x, n = blockIdx.y;
    const int cols = w * 16;
    for (int i = threadIdx.x; i < 3 * 16 * cols; i += 256) {
        const int x = i % cols, r = i / cols;          // r = c*16 + ky
        const int c = r >> 4, ky = r & 15;
        const float v = img[((size_t)(n * 3 + c) * H + (py * 16 + ky)) * W + x];
        const int px = x >> 4, kx = x & 15;
        out[((size_t)(n * h + py) * w + px) * 768 + c * 256 + ky * 16 + kx] = __float2half_rn(v);
    }
}

int launch_im2col16(const float* img, h16* out, int N, int H, int W, cudaStream_t st) {
    const int h = H / 16, w = W / 16;
    WC_CHECK_ARG(img && out && N > 0 && h > 0 && w > 0 && N <= 65535);
    ProfScope prof(PK_ENC_MISC, 0.0, st);
    im2col16_kernel<<<dim3(h, N), 256, 0, st>>>(img, out, H, W, h, w);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

// ---------------------------------------------------------------------------------------------------
// Positional embedding resize (clip/model.py:11-27): bilinear, align_corners=False, of the gs x gs grid to
// h x w, result rounded through fp16 (the reference stores the resized embedding as .half()).
// Row 0 (class token) is copied.  Also writes x[n, 0, :] = class_embedding + pos[0] (clip/model.py:270-271).
// ---------------------------------------------------------------------------------------------------
__global__ void posemb_resize_kernel(const float* __restrict__ pos, float* __restrict__ out, int gs, int h, int w, int E) {
    const int t = blockIdx.x;   // output token 0..h*w
    for (int e = threadIdx.x; e < E; e += blockDim.x) {
        float v;
        if (t == 0) {
            v = pos[e];
        } else {
            const int oy = (t - 1) / w, ox = (t - 1) % w;
            // torch area_pixel_compute_source_index, align_corners=False: src = scale*(dst+0.5)-0.5, clamped at 0
            const float sy = (float)gs / (float)h, sx = (float)gs / (float)w;
            float fy = fmaxf(sy * (oy + 0.5f) - 0.5f, 0.f), fx = fmaxf(sx * (ox + 0.5f) - 0.5f, 0.f);
            const int y0 = (int)fy, x0 = (int)fx;
            const int y1 = y0 + (y0 < gs - 1 ? 1 : 0), x1 = x0 + (x0 < gs - 1 ? 1 : 0);
            const float ly = fy - y0, lx = fx - x0;
            const float* g = pos + E;   // grid rows start after the class row
            const float v00 = g[(size_t)(y0 * gs + x0) * E + e], v01 = g[(size_t)(y0 * gs + x1) * E + e];
            const float v10 = g[(size_t)(y1 * gs + x0) * E + e], v11 = g[(size_t)(y1 * gs + x1) * E + e];
            v = (1.f - ly) * ((1.f - lx) * v00 + lx * v01) + ly * ((1.f - lx) * v10 + lx * v11);
        }
        out[(size_t)t * E + e] = __half2float(__float2half_rn(v));
    }
}

__global__ void cls_row_kernel(float* __restrict__ x, const float* __restrict__ cls, const float* __restrict__ pos, int L, int E) {
    const int n = blockIdx.x;
    for (int e = threadIdx.x; e < E; e += blockDim.x) x[(size_t)n * L * E + e] = cls[e] + pos[e];
}

int launch_posemb_resize(const float* pos, float* out, int gs, int h, int w, int E, cudaStream_t st) {
    WC_CHECK_ARG(pos && out && gs > 0 && h > 0 && w > 0);
    ProfScope prof(PK_ENC_MISC, 0.0, st);
    posemb_resize_kernel<<<h * w + 1, 256, 0, st>>>(pos, out, gs, h, w, E);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

int launch_cls_row(float* x, const float* cls, const float* pos, int N, int L, int E, cudaStream_t st) {
    ProfScope prof(PK_ENC_MISC, 0.0, st);
    cls_row_kernel<<<N, 256, 0, st>>>(x, cls, pos, L, E);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

// fp32 -> h16 cast (contiguous)
__global__ void cast_bf16_kernel(const float* __restrict__ x, h16* __restrict__ y, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __float2half_rn(x[i]);
}
int launch_cast_bf16(const float* x, h16* y, size_t n, cudaStream_t st) {
    ProfScope prof(PK_ENC_MISC, 0.0, st);
    cast_bf16_kernel<<<(unsigned)ceil_div(n, (size_t)256), 256, 0, st>>>(x, y, n);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

}  // namespace wc
