// Grad-CAM head for sm_100a: block 12 forward + pooled cosine-softmax head + the ANALYTIC input-gradient
// backward to the ln_1 output, batched over (image, class) instances.
//
// Replaces: CLIP.forward_last_layer (clip/model.py:407-429), GradCAM/BaseCAM.forward
// (pytorch_grad_cam/base_cam.py:62-108,116-154), get_cam_weights (grad_cam.py:16-23), the ln_1 hooks
// (activations_and_gradients.py:19-37), reshape_transform (model_attn_aff_voc.py:23-30), ClipOutputTarget
// (clip_tool.py:20-26) and scale_cam_image (utils/image.py:51-61), as driven per class by clip_tool.py:139-147.
//
// Math.  With A = ln_1(x11) the hook tensor, only the q/k/v branch carries gradient
// to A, and Grad-CAM only needs  w = mean_{patch rows}(dp_k/dA) = (1/P) * [sum_rows dQ | sum_rows dK | sum_rows dV] @ W_in.
// The three row sums follow from the attention backward without materialising dQ/dK/dV:
//     dP = dO V^T,  delta_i = sum_j dP_ij P_ij,  dS = P * (dP - delta)
//     sum_{i>=1} dQs_i = sum_j (sum_{i>=1} dS_ij) K_j          (column sums of dS over patch rows)
//     sum_{j>=1} dK_j  = sum_i (-dS_i0) Qs_i                   (rows of dS sum to zero)
//     sum_{j>=1} dV_j  = sum_i (1 - P_i0) dO_i
// The column sums themselves never touch an L x L tensor in memory (attention.cu, attn_bwd_colsum_kernel):
//     delta_i = dO_i . O_i,   colsum_j = V_j . (sum_{i>=1} P_ij dO_i) - sum_{i>=1} P_ij delta_i
// with P recomputed tile by tile from q, k and the forward's row log-sum-exps; dS_i0 and P_i0 are per-row dot products.
#include "attention.cuh"
#include "encoder.cuh"

#include <math.h>

namespace wc {

// 0 (default): reproduce the reference's fp16 rounding of the out-projection's backward (ref_half_round below);
// 1: exact gradient (no reference rounding) — wc_gradcam_set_exact, for analysis.
static int g_gradcam_exact = 0;

constexpr int GC_MAX_IMG = 64;
constexpr int GC_MAX_INST = 64;      // instances processed per backward chunk
constexpr int GC_MAX_T = 128;        // text rows per image (VOC 20+25, COCO 80+23)
constexpr int GC_E = 768, GC_HEADS = 12, GC_HD = 64, GC_EMB = 512;

struct ImgTable {
    int n;
    int text_off[GC_MAX_IMG + 1];
    int k_off[GC_MAX_IMG + 1];
};
struct InstTable {
    int n;                   // instances in this chunk
    int inst0;               // global index of the first instance of the chunk
    short img[GC_MAX_INST];  // image of each instance
};

// image index of each instance of the chunk, as a device array (batch -> aux index of the GEMM epilogue)
__global__ void inst_map_kernel(const __grid_constant__ InstTable it, int* __restrict__ map) {
    if ((int)threadIdx.x < it.n) map[threadIdx.x] = it.img[threadIdx.x];
}

// ---------------------------------------------------------------------------------------------------
// ln_post + column sums over the patch rows (clip/model.py:411-412).  mpart[img][block][c] = sum over the block's 32 rows of
// LN(x2)[row][c]; head_kernel adds the blocks in order (no floating-point atomics: the same bits on every run)
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) lnpost_mean_kernel(const float* __restrict__ x2, const float* __restrict__ gw,
                                                          const float* __restrict__ gb, float* __restrict__ msum, int L) {
    __shared__ float part[8][GC_E];
    const int img = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float acc[24];
#pragma unroll
    for (int i = 0; i < 24; ++i) acc[i] = 0.f;
    for (int r = 0; r < 4; ++r) {
        const int l = 1 + blockIdx.x * 32 + warp * 4 + r;
        if (l >= L) break;
        const float4* xr = reinterpret_cast<const float4*>(x2 + ((size_t)img * L + l) * GC_E);
        float4 v[6];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 6; ++i) { v[i] = xr[lane + 32 * i]; s += v[i].x + v[i].y + v[i].z + v[i].w; }
        const float mean = warp_sum(s) / GC_E;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
            q += a * a + b * b + c * c + d * d;
        }
        const float rstd = rsqrtf(warp_sum(q) / GC_E + 1e-5f);
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const float4 ww = __ldg(reinterpret_cast<const float4*>(gw) + lane + 32 * i);
            const float4 bb = __ldg(reinterpret_cast<const float4*>(gb) + lane + 32 * i);
            acc[4 * i + 0] += (v[i].x - mean) * rstd * ww.x + bb.x;
            acc[4 * i + 1] += (v[i].y - mean) * rstd * ww.y + bb.y;
            acc[4 * i + 2] += (v[i].z - mean) * rstd * ww.z + bb.z;
            acc[4 * i + 3] += (v[i].w - mean) * rstd * ww.w + bb.w;
        }
    }
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) part[warp][(lane + 32 * i) * 4 + j] = acc[4 * i + j];
    __syncthreads();
    for (int c = threadIdx.x; c < GC_E; c += 256) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += part[w][c];
        msum[((size_t)img * gridDim.x + blockIdx.x) * GC_E + c] = s;
    }
}

// ---------------------------------------------------------------------------------------------------
// Pooled head forward + its backward to d(mean-pooled ln_post output) for every class of the image
// (clip/model.py:411-427 and the softmax / L2-norm / proj Jacobians).  One CTA (512 threads) per image.
//   dm[inst][c] = d p_k / d y[l][c] for every patch row l  ( = (proj @ dg)[c] / P )
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) head_kernel(const float* __restrict__ msum, int n_part, const float* __restrict__ proj,
                                                   const float* __restrict__ text, const __grid_constant__ ImgTable tab,
                                                   float logit_scale, int P, float* __restrict__ dm,
                                                   float* __restrict__ gscale, float* __restrict__ probs_out) {
    __shared__ float mvec[GC_E];
    __shared__ __align__(16) float g[GC_EMB], dgh[GC_EMB];
    __shared__ float logit[GC_MAX_T], tnorm[GC_MAX_T], dl[GC_MAX_T];
    __shared__ float red[32];
    const int img = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int t0 = tab.text_off[img], T = tab.text_off[img + 1] - t0;
    const int K = tab.k_off[img + 1] - tab.k_off[img];
    for (int c = tid; c < GC_E; c += 512) {
        float sum = 0.f;
        for (int b = 0; b < n_part; ++b) sum += msum[((size_t)img * n_part + b) * GC_E + c];
        mvec[c] = sum / (float)P;
    }
    __syncthreads();
    {   // g = mean @ proj  (model.py:414-415)
        // 768 dependent-free loads per thread: four accumulators and a deep unroll keep ~16 L2 requests in flight per thread
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 8
        for (int c = 0; c < GC_E; c += 4) {
            s0 = fmaf(mvec[c], __ldg(proj + (size_t)c * GC_EMB + tid), s0);
            s1 = fmaf(mvec[c + 1], __ldg(proj + (size_t)(c + 1) * GC_EMB + tid), s1);
            s2 = fmaf(mvec[c + 2], __ldg(proj + (size_t)(c + 2) * GC_EMB + tid), s2);
            s3 = fmaf(mvec[c + 3], __ldg(proj + (size_t)(c + 3) * GC_EMB + tid), s3);
        }
        g[tid] = (s0 + s1) + (s2 + s3);
    }
    __syncthreads();
    const float gn = sqrtf(block_sum(g[tid] * g[tid], red));
    __syncthreads();
    g[tid] = g[tid] / gn;          // normalised image feature (model.py:420)
    __syncthreads();
    for (int t = warp; t < T; t += 16) {   // cosine logits (model.py:421-424)
        const float* tr = text + (size_t)(t0 + t) * GC_EMB;
        float dot = 0.f, nn = 0.f;
        for (int j = lane; j < GC_EMB; j += 32) { const float v = __ldg(tr + j); dot = fmaf(v, g[j], dot); nn = fmaf(v, v, nn); }
        dot = warp_sum(dot); nn = warp_sum(nn);
        if (lane == 0) { tnorm[t] = sqrtf(nn); logit[t] = logit_scale * dot / sqrtf(nn); }
    }
    __syncthreads();
    if (warp == 0) {   // softmax over the T texts (model.py:427)
        float mx = -INFINITY;
        for (int t = lane; t < T; t += 32) mx = fmaxf(mx, logit[t]);
        mx = warp_max(mx);
        float s = 0.f;
        for (int t = lane; t < T; t += 32) s += expf(logit[t] - mx);
        s = warp_sum(s);
        for (int t = lane; t < T; t += 32) logit[t] = expf(logit[t] - mx) / s;   // now probabilities
    }
    __syncthreads();
    if (probs_out) for (int t = tid; t < T; t += 512) probs_out[t0 + t] = logit[t];
    for (int k = 0; k < K; ++k) {
        const int inst = tab.k_off[img] + k;
        const float pk = logit[k];
        for (int t = tid; t < T; t += 512) dl[t] = pk * ((t == k ? 1.f : 0.f) - logit[t]);   // d p_k / d logit_t
        __syncthreads();
        float s = 0.f;   // d p_k / d ghat_j
#pragma unroll 8
        for (int t = 0; t < T; ++t) s = fmaf(dl[t] / tnorm[t], __ldg(text + (size_t)(t0 + t) * GC_EMB + tid), s);
        s *= logit_scale;
        const float dot = block_sum(s * g[tid], red);
        dgh[tid] = (s - g[tid] * dot) / gn;          // through the L2 normalisation
        __syncthreads();
        // d / d mean = proj @ dg ; every patch row gets 1/P of it.  Four rows per warp in flight (one CTA per image: the kernel is
        // bound by the latency of these L2 reads, not by their volume)
        {
            float4 d4[GC_EMB / 128];
#pragma unroll
            for (int j = 0; j < GC_EMB / 128; ++j) d4[j] = reinterpret_cast<const float4*>(dgh)[lane + 32 * j];
            for (int c = warp; c < GC_E; c += 64) {
                float4 w4[4][GC_EMB / 128];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float4* pr = reinterpret_cast<const float4*>(proj + (size_t)(c + 16 * u) * GC_EMB);
#pragma unroll
                    for (int j = 0; j < GC_EMB / 128; ++j) w4[u][j] = __ldg(pr + lane + 32 * j);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    float a = 0.f;
#pragma unroll
                    for (int j = 0; j < GC_EMB / 128; ++j)
                        a += w4[u][j].x * d4[j].x + w4[u][j].y * d4[j].y + w4[u][j].z * d4[j].z + w4[u][j].w * d4[j].w;
                    a = warp_sum(a);
                    if (lane == 0) dm[(size_t)inst * GC_E + c + 16 * u] = a / (float)P;
                }
            }
        }
        __syncthreads();
        // power-of-two gradient scale: the backward chain runs through fp16 GEMM operands and dp_k/dy is ~1e-6 for
        // near-uniform p (below fp16's normal range); the chain is linear in the gradient, so scaling by 2^e here and
        // dividing the channel weights by 2^e in cam_weight_kernel is exact in fp32.
        float mx = 0.f;
        for (int c = tid; c < GC_E; c += 512) mx = fmaxf(mx, fabsf(dm[(size_t)inst * GC_E + c]));
        mx = block_max(mx, red);
        if (tid == 0) gscale[inst] = (mx > 0.f && isfinite(mx)) ? exp2f(3.f - ceilf(log2f(mx))) : 1.f;
        __syncthreads();
    }
}

// The reference forces the attention out-projection to fp16 on every device (clip/myAtt.py:321), so autograd runs that
// linear's backward in fp16 too: the incoming gradient is cast to half, multiplied by the half weight (fp32 accumulate)
// and the result is rounded to half again.  With near-uniform class probabilities the gradient there is ~1e-7, i.e.
// inside fp16's SUBNORMAL range (quantum 2^-24 = 6e-8): the reference keeps only 2-3 significant bits of it, and that
// quantisation — not arithmetic error — is what its Grad-CAM weights carry (measured: oracle with vs without it differs by
// 1.8e-2 / 3.6e-2 in the normalised CAM at L = 1025 / 1601, tools/gradcam_debug.py).  To return the reference's result this
// chain reproduces the two roundings: the power-of-two scaled gradient g is divided by its scale, rounded to fp16
// (round-to-nearest-even, subnormals kept) and scaled back — exact, because the scale is a power of two.
__device__ __forceinline__ float ref_half_round(float g_scaled, float scale) {
    return __half2float(__float2half_rn(g_scaled / scale)) * scale;
}

// dO (fp32, scaled) -> fp16 operand with the reference's rounding of the out-projection's input gradient (see above).
__global__ void __launch_bounds__(256) quant_ref_half_kernel(const float* __restrict__ in, h16* __restrict__ out,
                                                             const float* __restrict__ gscale, int inst0, long per_inst,
                                                             long total) {
    const long i = ((long)blockIdx.x * 256 + threadIdx.x) * 4;
    if (i >= total) return;
    const float s = gscale[inst0 + (int)(i / per_inst)];
    const float4 v = *reinterpret_cast<const float4*>(in + i);
    h16x2 lo = __floats2half2_rn(ref_half_round(v.x, s), ref_half_round(v.y, s));
    h16x2 hi = __floats2half2_rn(ref_half_round(v.z, s), ref_half_round(v.w, s));
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(out + i) = pk;
}

// ---------------------------------------------------------------------------------------------------
// LayerNorm input-gradient.  out[r] = add[r] + rstd * (g - mean(g) - xhat * mean(g*xhat)),  g = gamma * dy.
// dy is either a matrix [rows,768] (dy_mat) or one vector per instance broadcast over the patch rows with the
// class-token row receiving zero (dy_vec; ln_post feeds only x[:,1:], model.py:412).  One warp per row.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                     const float* __restrict__ dy_mat, const float* __restrict__ dy_vec,
                                                     const float* __restrict__ add, float* __restrict__ out_f,
                                                     h16* __restrict__ out_h, const __grid_constant__ InstTable it,
                                                     int L, const float* __restrict__ gscale, int quant_ref) {
    const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= (long)it.n * L) return;
    const int inst = (int)(row / L), l = (int)(row % L);
    const float4* xr = reinterpret_cast<const float4*>(x + ((size_t)it.img[inst] * L + l) * GC_E);
    float4 v[6], gg[6];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) { v[i] = xr[lane + 32 * i]; s += v[i].x + v[i].y + v[i].z + v[i].w; }
    const float mean = warp_sum(s) / GC_E;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
        q += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
    }
    const float rstd = rsqrtf(warp_sum(q) / GC_E + 1e-5f);
    const bool zero = (dy_vec != nullptr && l == 0);
    const float gs = (dy_vec != nullptr && gscale != nullptr) ? gscale[it.inst0 + inst] : 1.f;
    // quant_ref: out_h is rounded to the fp16 grid of the UNSCALED gradient (see ref_half_round below)
    const float qs = (quant_ref && gscale != nullptr) ? gscale[it.inst0 + inst] : 0.f;
    const float4* dyr = dy_mat ? reinterpret_cast<const float4*>(dy_mat + (size_t)row * GC_E)
                               : reinterpret_cast<const float4*>(dy_vec + (size_t)(it.inst0 + inst) * GC_E);
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * i);
        float4 d = zero ? make_float4(0.f, 0.f, 0.f, 0.f) : dyr[lane + 32 * i];
        v[i].x *= rstd; v[i].y *= rstd; v[i].z *= rstd; v[i].w *= rstd;      // xhat
        gg[i] = make_float4(w.x * d.x * gs, w.y * d.y * gs, w.z * d.z * gs, w.w * d.w * gs);
        c1 += gg[i].x + gg[i].y + gg[i].z + gg[i].w;
        c2 += gg[i].x * v[i].x + gg[i].y * v[i].y + gg[i].z * v[i].z + gg[i].w * v[i].w;
    }
    c1 = warp_sum(c1) / GC_E;
    c2 = warp_sum(c2) / GC_E;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const int c4 = lane + 32 * i;
        float4 o;
        o.x = rstd * (gg[i].x - c1 - v[i].x * c2);
        o.y = rstd * (gg[i].y - c1 - v[i].y * c2);
        o.z = rstd * (gg[i].z - c1 - v[i].z * c2);
        o.w = rstd * (gg[i].w - c1 - v[i].w * c2);
        if (add) {
            const float4 a = reinterpret_cast<const float4*>(add + (size_t)row * GC_E)[c4];
            o.x += a.x; o.y += a.y; o.z += a.z; o.w += a.w;
        }
        if (out_f) reinterpret_cast<float4*>(out_f + (size_t)row * GC_E)[c4] = o;
        if (out_h) {
            if (qs != 0.f) { o.x = ref_half_round(o.x, qs); o.y = ref_half_round(o.y, qs); o.z = ref_half_round(o.z, qs); o.w = ref_half_round(o.w, qs); }
            h16x2 lo = __floats2half2_rn(o.x, o.y), hi = __floats2half2_rn(o.z, o.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo);
            pk.y = *reinterpret_cast<uint32_t*>(&hi);
            reinterpret_cast<uint2*>(out_h + (size_t)row * GC_E)[c4] = pk;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// Operand preparation for attn_bwd_colsum_kernel + the per-row terms.  For (64-query chunk, head, instance):
//   delta_i = dO_i . O_i ;  dOT[inst,h] = [dO^T ; -delta (hi, lo fp16 split) ; 0]  as [80][Lp] fp16 ;
//   P_i0 = exp2(Qs_i . K_0 log2e - lse2_i) ;  dS_i0 = P_i0 (dO_i . V_0 - delta_i).
// dO fp16 [n_inst*L, 768] (chunk-local instances), O fp16 [N*L, 768] (block-12 attention output, head-interleaved).
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void load16h(const h16* p, float* out) {
    const uint4 a = *reinterpret_cast<const uint4*>(p), b = *reinterpret_cast<const uint4*>(p + 8);
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const float2 f = __half22float2(*reinterpret_cast<const h16x2*>(&w[e]));
        out[2 * e] = f.x; out[2 * e + 1] = f.y;
    }
}

__global__ void __launch_bounds__(256) do_prep_kernel(const h16* __restrict__ dO, const h16* __restrict__ O,
                                                      const h16* __restrict__ q, const h16* __restrict__ k,
                                                      const h16* __restrict__ v, const float* __restrict__ lse2,
                                                      h16* __restrict__ dOT, float* __restrict__ ds0, float* __restrict__ p0,
                                                      const __grid_constant__ InstTable it, int L, int Lp) {
    __shared__ h16x2 tile[64][33];       // [query][dim pair], padded: the transposed read is 2-way conflicted at worst
    __shared__ h16 dl[2][64];
    const int i0 = blockIdx.x * 64, h = blockIdx.y, inst = blockIdx.z, tid = threadIdx.x;
    const int img = it.img[inst];
    const size_t hb = (size_t)img * GC_HEADS + h, ih = (size_t)inst * GC_HEADS + h;
    const int r = tid >> 2, qd = tid & 3, i = i0 + r;
    float dlt = 0.f, dp0 = 0.f, s0 = 0.f;
    float g[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) g[e] = 0.f;
    if (i < L) {
        float o[16], qq[16], k0[16], v0[16];
        load16h(dO + ((size_t)inst * L + i) * GC_E + h * GC_HD + qd * 16, g);
        load16h(O + ((size_t)img * L + i) * GC_E + h * GC_HD + qd * 16, o);
        load16h(q + (hb * L + i) * GC_HD + qd * 16, qq);
        load16h(k + (hb * L) * GC_HD + qd * 16, k0);
        load16h(v + (hb * L) * GC_HD + qd * 16, v0);
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            dlt = fmaf(g[e], o[e], dlt);
            dp0 = fmaf(g[e], v0[e], dp0);
            s0 = fmaf(qq[e], k0[e], s0);
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) tile[r][qd * 8 + e] = __floats2half2_rn(g[2 * e], g[2 * e + 1]);
#pragma unroll
    for (int m = 1; m <= 2; m <<= 1) {
        dlt += __shfl_xor_sync(0xffffffffu, dlt, m);
        dp0 += __shfl_xor_sync(0xffffffffu, dp0, m);
        s0 += __shfl_xor_sync(0xffffffffu, s0, m);
    }
    if (qd == 0) {
        if (i < L) {
            const float p = exp2f(s0 * 1.4426950408889634f - __ldg(lse2 + hb * L + i));
            p0[ih * L + i] = p;
            ds0[ih * L + i] = p * (dp0 - dlt);
        }
        const h16 hi = __float2half_rn(-dlt);
        dl[0][r] = hi;
        dl[1][r] = __float2half_rn(-dlt - __half2float(hi));
    }
    __syncthreads();
    const h16* th = reinterpret_cast<const h16*>(&tile[0][0]);
    for (int idx = tid; idx < 80 * 8; idx += 256) {
        const int d = idx >> 3, c8 = (idx & 7) * 8, col = i0 + c8;
        if (col >= Lp) continue;
        __align__(16) h16 vals[8];
#pragma unroll
        for (int e = 0; e < 8; ++e)
            vals[e] = d < 64 ? th[(c8 + e) * 66 + d] : d == 64 ? dl[0][c8 + e] : d == 65 ? dl[1][c8 + e] : __float2half_rn(0.f);
        *reinterpret_cast<uint4*>(dOT + (ih * 80 + d) * (size_t)Lp + col) = *reinterpret_cast<const uint4*>(vals);
    }
}

// ---------------------------------------------------------------------------------------------------
// Row sums of dQ, dK, dV over the patch tokens -> dqkv[inst][2304] (see file header).  One CTA per (head, inst).
// ---------------------------------------------------------------------------------------------------
// 16 row groups of 64 threads, four rows per group in flight (the first version walked the rows with 4 groups and one row at a
// time: 180 CTAs of latency-bound 2-byte loads, 0.48 TB/s).
constexpr int QS_GROUPS = 16;
__global__ void __launch_bounds__(64 * QS_GROUPS) qkv_grad_sum_kernel(const float* __restrict__ colsum, const float* __restrict__ ds0,
                                                                     const h16* __restrict__ q, const h16* __restrict__ k,
                                                                     const float* __restrict__ p0, const h16* __restrict__ dO,
                                                                     float* __restrict__ dqkv, const __grid_constant__ InstTable it,
                                                                     int L, int Lp, float qscale) {
    __shared__ float part[3][QS_GROUPS][GC_HD];
    const int h = blockIdx.x, inst = blockIdx.y, d = threadIdx.x & 63, grp = threadIdx.x >> 6;
    const int img = it.img[inst];
    const size_t hb = (size_t)(img * GC_HEADS + h);
    const float* cs = colsum + (size_t)(inst * GC_HEADS + h) * L;
    const float* d0 = ds0 + (size_t)(inst * GC_HEADS + h) * L;
    const float* pp = p0 + (size_t)(inst * GC_HEADS + h) * L;
    const h16* kp = k + hb * L * GC_HD + d;
    const h16* qp = q + hb * L * GC_HD + d;
    const h16* op = dO + (size_t)inst * L * GC_E + h * GC_HD + d;
    float gq = 0.f, gk = 0.f, gv = 0.f;
    int i = grp;
    for (; i + 3 * QS_GROUPS < L; i += 4 * QS_GROUPS) {
        float a[4], b[4], c[4], kv[4], qv[4], ov[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int r = i + u * QS_GROUPS;
            a[u] = __ldg(cs + r); b[u] = __ldg(d0 + r); c[u] = __ldg(pp + r);
            kv[u] = __half2float(kp[(size_t)r * GC_HD]);
            qv[u] = __half2float(qp[(size_t)r * GC_HD]);
            ov[u] = __half2float(op[(size_t)r * GC_E]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            gq = fmaf(a[u], kv[u], gq);
            gk = fmaf(-b[u], qv[u], gk);
            gv = fmaf(1.f - c[u], ov[u], gv);
        }
    }
    for (; i < L; i += QS_GROUPS) {
        gq = fmaf(cs[i], __half2float(kp[(size_t)i * GC_HD]), gq);
        gk = fmaf(-d0[i], __half2float(qp[(size_t)i * GC_HD]), gk);
        gv = fmaf(1.f - pp[i], __half2float(op[(size_t)i * GC_E]), gv);
    }
    part[0][grp][d] = gq; part[1][grp][d] = gk; part[2][grp][d] = gv;
    __syncthreads();
    if (grp < 3) {
        float s = 0.f;
#pragma unroll
        for (int g = 0; g < QS_GROUPS; ++g) s += part[grp][g][d];
        dqkv[(size_t)inst * 3 * GC_E + grp * GC_E + h * GC_HD + d] = (grp == 0) ? s * qscale : s;
    }
}

// w[inst][c] = (1/P) * sum_r dqkv[inst][r] * W_in[r][c]   (mean over patch tokens of the hooked gradient, grad_cam.py:23)
// 64 columns x 4 row groups per CTA (the row groups are reduced through shared memory).
__global__ void __launch_bounds__(256) cam_weight_kernel(const float* __restrict__ dqkv, const float* __restrict__ w_in,
                                                         float* __restrict__ wvec, const float* __restrict__ gscale, int P) {
    __shared__ float dq[3 * GC_E];
    __shared__ float part[4][64];
    const int inst = blockIdx.y, cl = threadIdx.x & 63, grp = threadIdx.x >> 6, c = blockIdx.x * 64 + cl;
    for (int r = threadIdx.x; r < 3 * GC_E; r += 256) dq[r] = dqkv[(size_t)inst * 3 * GC_E + r];
    __syncthreads();
    float s0 = 0.f, s1 = 0.f;
    const int r0 = grp * (3 * GC_E / 4);
#pragma unroll 8
    for (int r = r0; r < r0 + 3 * GC_E / 4; r += 2) {
        s0 = fmaf(dq[r], __ldg(w_in + (size_t)r * GC_E + c), s0);
        s1 = fmaf(dq[r + 1], __ldg(w_in + (size_t)(r + 1) * GC_E + c), s1);
    }
    part[grp][cl] = s0 + s1;
    __syncthreads();
    if (grp == 0) wvec[(size_t)inst * GC_E + c] = (part[0][cl] + part[1][cl] + part[2][cl] + part[3][cl]) / (float)P / gscale[inst];
}

// cam = relu(sum_c w_c A_c) over the patch tokens (base_cam.py:55-59,144-145): one warp per (patch, instance), the grid
// covers every patch of every instance so the A32 stream runs at memory speed (the K classes of an image re-read it from L2).
__global__ void __launch_bounds__(256) cam_dot_kernel(const float* __restrict__ A32, const float* __restrict__ wvec,
                                                      float* __restrict__ cams, const __grid_constant__ InstTable it, int L) {
    __shared__ float4 wv[GC_E / 4];
    const int inst = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int P = L - 1, img = it.img[inst];
    if (tid < GC_E / 4) wv[tid] = reinterpret_cast<const float4*>(wvec + (size_t)(it.inst0 + inst) * GC_E)[tid];
    __syncthreads();
    float* o = cams + (size_t)(it.inst0 + inst) * P;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int p = (blockIdx.x * 8 + warp) * 4 + r;
        if (p >= P) break;
        const float4* a = reinterpret_cast<const float4*>(A32 + ((size_t)img * L + 1 + p) * GC_E);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const float4 x = a[lane + 32 * i], w = wv[lane + 32 * i];
            s += x.x * w.x + x.y * w.y + x.z * w.z + x.w * w.w;
        }
        s = warp_sum(s);
        if (lane == 0) o[p] = fmaxf(s, 0.f);
    }
}

// scale_cam_image twice (utils/image.py:51-61, base_cam.py:150-154), in place, one CTA per instance.
__global__ void __launch_bounds__(256) cam_norm_kernel(float* __restrict__ cams, int inst0, int P) {
    extern __shared__ float cam[];   // [P]
    __shared__ float red[32];
    const int tid = threadIdx.x;
    float* o = cams + (size_t)(inst0 + blockIdx.x) * P;
    for (int p = tid; p < P; p += 256) cam[p] = o[p];
    __syncthreads();
    for (int rep = 0; rep < 2; ++rep) {
        float mn = INFINITY;
        for (int p = tid; p < P; p += 256) mn = fminf(mn, cam[p]);
        mn = block_min(mn, red);
        float mx = -INFINITY;
        for (int p = tid; p < P; p += 256) mx = fmaxf(mx, cam[p] - mn);
        mx = block_max(mx, red);
        const float den = 1e-7f + mx;
        for (int p = tid; p < P; p += 256) cam[p] = (cam[p] - mn) / den;
        __syncthreads();
    }
    for (int p = tid; p < P; p += 256) o[p] = cam[p];
}

// ---------------------------------------------------------------------------------------------------
struct GcLayout {
    // forward (per call)
    float *a32, *x1, *x2, *lse, *msum, *dm, *dqkv, *wvec, *gscale;
    h16 *a, *q, *k, *v, *vt, *o, *u, *upre;
    // backward (per chunk of <= chunk instances)
    float *dx2, *da2, *colsum, *ds0, *p0;
    h16 *dx2b, *dh, *dx1b, *dO, *dOT;
    int* inst_map;
};

static void gc_layout(Arena& ar, GcLayout& g, int N, int L, int total_inst, int chunk) {
    const long M = (long)N * L, Lp = round8(L), E = GC_E;
    g.a32 = ar.get<float>(M * E); g.x1 = ar.get<float>(M * E); g.x2 = ar.get<float>(M * E);
    g.lse = ar.get<float>((long)N * GC_HEADS * L);
    g.msum = ar.get<float>((long)N * ceil_div(L - 1, 32) * E); g.dm = ar.get<float>((long)total_inst * E);
    g.dqkv = ar.get<float>((long)total_inst * 3 * E); g.wvec = ar.get<float>((long)total_inst * E);
    g.gscale = ar.get<float>((long)total_inst);
    g.a = ar.get<h16>(M * E); g.q = ar.get<h16>(M * E); g.k = ar.get<h16>(M * E); g.v = ar.get<h16>(M * E);
    g.vt = nullptr;   // head_dim 64: the fused attention reads V itself
    g.o = ar.get<h16>(M * E); g.u = ar.get<h16>(M * 4 * E); g.upre = ar.get<h16>(M * 4 * E);
    const long R = (long)chunk * L;
    g.dx2 = ar.get<float>(R * E); g.da2 = ar.get<float>(R * E);
    g.colsum = ar.get<float>((long)chunk * GC_HEADS * L); g.ds0 = ar.get<float>((long)chunk * GC_HEADS * L);
    g.p0 = ar.get<float>((long)chunk * GC_HEADS * L);
    g.dx2b = ar.get<h16>(R * E); g.dh = ar.get<h16>(R * 4 * E); g.dx1b = ar.get<h16>(R * E); g.dO = ar.get<h16>(R * E);
    g.dOT = ar.get<h16>((long)chunk * GC_HEADS * 80 * Lp);
    g.inst_map = ar.get<int>(GC_MAX_INST);
}

// instances per backward chunk: 16 (bounds the dP scratch), but never less than the largest image's class count
static int chunk_size(int total_inst, int max_k) {
    int c = total_inst < 16 ? total_inst : 16;
    if (c < max_k) c = max_k;
    return c < 1 ? 1 : c;
}

}  // namespace wc

using namespace wc;

extern "C" size_t wc_gradcam_workspace_bytes(int N, int h, int w, int total_inst, int max_k) {
    Arena ar(nullptr, 0);
    GcLayout g;
    gc_layout(ar, g, N, h * w + 1, total_inst, chunk_size(total_inst, max_k));
    return ar.off + 512;
}

extern "C" int wc_gradcam_set_exact(int exact) { const int old = g_gradcam_exact; g_gradcam_exact = exact ? 1 : 0; return old; }

// Test hook: byte offsets (from the workspace pointer passed to wc_gradcam, which must be 256-byte aligned) of the
// intermediate buffers of the LAST backward chunk, in the order
//   0 a32, 1 x1, 2 x2, 3 lse, 4 dm, 5 dqkv, 6 wvec, 7 gscale, 8 o(h16), 9 upre(h16), 10 dx2, 11 da2, 12 colsum, 13 ds0, 14 p0,
//   15 dx2b(h16), 16 dh(h16), 17 dx1b(h16), 18 dO(h16), 19 dOT(h16), 20 q(h16), 21 k(h16), 22 v(h16); returns the chunk size.
extern "C" int wc_gradcam_debug_offsets(int N, int h, int w, int total_inst, int max_k, size_t* offs) {
    WC_CHECK_ARG(offs && N > 0 && h > 0 && w > 0 && total_inst > 0);
    Arena ar(reinterpret_cast<void*>((uintptr_t)1 << 40), (size_t)1 << 50);     // fake aligned base: only offsets are read
    GcLayout g;
    const int chunk = chunk_size(total_inst, max_k);
    gc_layout(ar, g, N, h * w + 1, total_inst, chunk);
    const uint8_t* b = reinterpret_cast<const uint8_t*>((uintptr_t)1 << 40);
    const void* ptrs[23] = {g.a32, g.x1, g.x2, g.lse, g.dm, g.dqkv, g.wvec, g.gscale, g.o, g.upre, g.dx2, g.da2, g.colsum,
                            g.ds0, g.p0, g.dx2b, g.dh, g.dx1b, g.dO, g.dOT, g.q, g.k, g.v};
    for (int i = 0; i < 23; ++i) offs[i] = (size_t)(reinterpret_cast<const uint8_t*>(ptrs[i]) - b);
    return chunk;
}

extern "C" int wc_gradcam(const wc_vit_weights* w, const float* x11, int N, int h, int wd, const float* text,
                          const int* text_off_host, const int* k_off_host, float* cams, float* attn_last,
                          double* attn_last_sum, float* probs, void* workspace, size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(w && x11 && text && text_off_host && k_off_host && cams && attn_last && workspace);
    WC_CHECK_ARG(N > 0 && N <= GC_MAX_IMG && h > 0 && wd > 0);
    cudaStream_t st = as_stream(stream);
    const int P = h * wd, L = P + 1, E = GC_E;
    const long M = (long)N * L, Lp = round8(L);
    ImgTable tab; tab.n = N;
    for (int i = 0; i <= N; ++i) { tab.text_off[i] = text_off_host[i]; tab.k_off[i] = k_off_host[i]; }
    const int total = k_off_host[N];
    int max_k = 0;
    for (int i = 0; i < N; ++i) {
        const int K = k_off_host[i + 1] - k_off_host[i], T = text_off_host[i + 1] - text_off_host[i];
        WC_CHECK_ARG(K >= 1 && K <= T && T <= GC_MAX_T);
        max_k = K > max_k ? K : max_k;
    }
    const int chunk = chunk_size(total, max_k);
    WC_CHECK_ARG(chunk <= GC_MAX_INST);
    WC_CHECK_ARG(L <= 2048 && chunk <= BWD_MAX_INST);
    Arena ar(workspace, workspace_bytes);
    GcLayout g;
    gc_layout(ar, g, N, L, total, chunk);
    if (!ar.ok) return fail(WC_ERR_WORKSPACE, "%s: workspace too small (%ld bytes)", __func__, (long)workspace_bytes);
    int rc;
    // ---- block 12 forward, keeping what the backward needs (clip/model.py:408)
    {
        BlockBuffers b{};
        b.x_in = x11; b.x_mid = g.x1; b.x_out = g.x2;
        b.a = g.a; b.a_f32 = g.a32; b.q = g.q; b.k = g.k; b.v = g.v; b.vt = g.vt; b.lse = g.lse; b.o = g.o;   // fused attention
        b.u = g.u; b.u_pre = g.upre; b.attn_avg = attn_last; b.attn_sum = attn_last_sum;
        if ((rc = run_block(w->blocks[11], b, N, L, E, GC_HEADS, st))) return rc;
    }
    // ---- pooled head + its Jacobian for every (image, class)
    WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, lnpost_mean_kernel<<<dim3(ceil_div(P, 32), N), 256, 0, st>>>(g.x2, w->ln_post_w, w->ln_post_b, g.msum, L));
    WC_LAUNCH_CHECK();
    WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, head_kernel<<<N, 512, 0, st>>>(g.msum, ceil_div(P, 32), w->proj, text, tab, w->logit_scale_exp, P, g.dm, g.gscale, probs));
    WC_LAUNCH_CHECK();
    // ---- backward, in chunks of whole images holding <= chunk instances
    const float qscale = 1.f / sqrtf((float)GC_HD);
    int img0 = 0;
    while (img0 < N) {
        int img1 = img0, n_inst = 0;
        while (img1 < N && n_inst + (k_off_host[img1 + 1] - k_off_host[img1]) <= chunk) {
            n_inst += k_off_host[img1 + 1] - k_off_host[img1];
            ++img1;
        }
        if (img1 == img0) return fail(WC_ERR_UNSUPPORTED, "%s: more than %ld classes in one image", __func__, (long)chunk);
        InstTable it; it.n = n_inst; it.inst0 = k_off_host[img0];
        for (int i = img0, c = 0; i < img1; ++i)
            for (int k = k_off_host[i]; k < k_off_host[i + 1]; ++k) it.img[c++] = (short)i;
        const long R = (long)n_inst * L;
        const unsigned ln_grid = (unsigned)ceil_div(R, 8L);
        // ln_post backward (dy = dm broadcast over patch rows, class row zero)
        WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, ln_bwd_kernel<<<ln_grid, 256, 0, st>>>(g.x2, w->ln_post_w, nullptr, g.dm, nullptr, g.dx2, g.dx2b, it, L, g.gscale, 0));
        WC_LAUNCH_CHECK();
        // c_proj^T and the QuickGELU derivative (model.py:198-203): ONE launch batched over the chunk's instances; the saved
        // pre-activation u_pre belongs to the image, so the epilogue maps instance -> image through inst_map
        inst_map_kernel<<<1, GC_MAX_INST, 0, st>>>(it, g.inst_map);
        WC_LAUNCH_CHECK();
        {
            GemmProblem p; p.M = L; p.N = 4 * E; p.K = E; p.nb0 = n_inst;
            p.A = GemmOperand{g.dx2b, L, E, (long)L * E, 0};
            p.B = GemmOperand{reinterpret_cast<const h16*>(w->w_proj_t), 4 * E, E, 0, 0};
            p.epi.mode = GM_MUL_QGELU_GRAD; p.epi.out = g.dh; p.epi.ldo = 4 * E; p.epi.out_bs0 = (long)L * 4 * E;
            p.epi.aux = g.upre; p.epi.ldaux = 4 * E; p.epi.aux_bs = (long)L * 4 * E; p.epi.aux_map = g.inst_map;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        {   // c_fc^T
            GemmProblem p; p.M = (int)R; p.N = E; p.K = 4 * E;
            p.A = GemmOperand{g.dh, R, 4 * E, 0, 0};
            p.B = GemmOperand{reinterpret_cast<const h16*>(w->w_fc_t), E, 4 * E, 0, 0};
            p.epi.out = g.da2; p.epi.ldo = E;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        // ln_2 backward + the residual branch: dx1 = dx2 + LN2'(da2)
        WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, ln_bwd_kernel<<<ln_grid, 256, 0, st>>>(g.x1, w->blocks[11].ln2_w, g.da2, nullptr, g.dx2, nullptr, g.dx1b, it, L, g.gscale, g_gradcam_exact ? 0 : 1));
        WC_LAUNCH_CHECK();
        {   // out_proj^T (myAtt.py:321)
            GemmProblem p; p.M = (int)R; p.N = E; p.K = E;
            p.A = GemmOperand{g.dx1b, R, E, 0, 0};
            p.B = GemmOperand{reinterpret_cast<const h16*>(w->w_out_t), E, E, 0, 0};
            if (g_gradcam_exact) { p.epi.out = g.dO; p.epi.out_f16 = 1; }
            else p.epi.out = g.da2;            // fp32 (da2 is free again), rounded like the reference's half result below
            p.epi.ldo = E;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        if (!g_gradcam_exact) {
            const long tot = R * E;
            WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, quant_ref_half_kernel<<<(unsigned)ceil_div(tot / 4, 256L), 256, 0, st>>>(
                g.da2, g.dO, g.gscale, it.inst0, (long)L * E, tot));
            WC_LAUNCH_CHECK();
        }
        // delta, dO'^T, dS_i0, P_i0; then the column sums of dS without materialising P or dP (attention.cu)
        WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, do_prep_kernel<<<dim3(ceil_div(L, 64), GC_HEADS, n_inst), 256, 0, st>>>(
            g.dO, g.o, g.q, g.k, g.v, g.lse, g.dOT, g.ds0, g.p0, it, L, (int)Lp));
        WC_LAUNCH_CHECK();
        {
            BwdInstTable bt;
            for (int c = 0; c < n_inst; ++c) bt.img[c] = it.img[c];
            if ((rc = launch_attn_bwd_colsum(g.q, g.k, g.v, g.dOT, g.lse, g.colsum, bt, n_inst, N, GC_HEADS, L, (int)Lp, st))) return rc;
        }
        WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, qkv_grad_sum_kernel<<<dim3(GC_HEADS, n_inst), 64 * QS_GROUPS, 0, st>>>(g.colsum, g.ds0, g.q, g.k, g.p0, g.dO,
                                                                    g.dqkv + (long)it.inst0 * 3 * E, it, L, (int)Lp, qscale));
        WC_LAUNCH_CHECK();
        WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, cam_weight_kernel<<<dim3(E / 64, n_inst), 256, 0, st>>>(g.dqkv + (long)it.inst0 * 3 * E, w->w_in_f32,
                                                                 g.wvec + (long)it.inst0 * E, g.gscale + it.inst0, P));
        WC_LAUNCH_CHECK();
        WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, cam_dot_kernel<<<dim3(ceil_div(P, 32), n_inst), 256, 0, st>>>(g.a32, g.wvec, cams, it, L));
        WC_LAUNCH_CHECK();
        WC_PROF_LAUNCH(PK_GRADCAM_MISC, 0.0, st, cam_norm_kernel<<<n_inst, 256, sizeof(float) * P, st>>>(cams, it.inst0, P));
        WC_LAUNCH_CHECK();
        img0 = img1;
    }
    (void)M;
    return WC_OK;
}
