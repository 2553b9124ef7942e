// Building blocks of the trainable-decoder backward for sm_100a: everything that is not a GEMM.
//
// The reference trains decoder_fts_fuse + decoder + attn_pred with torch autograd (scripts/dist_clip_voc.py:246-267,
// WeCLIP_model/segformer_head.py:49-80, Decoder/TransDecoder.py:70-125, model_attn_aff_voc.py:134-137).  Here every matrix
// product of the forward, the input gradients and the weight gradients runs on the tcgen05 GEMM (wc_gemm_f16_ex) and these
// kernels supply the rest: operand transposes (the GEMM takes K-major operands only), LayerNorm / softmax / activation
// backward, bias gradients (column sums), the sigmoid-Gram backward of attn_pred and Dropout2d.  Host orchestration:
// weclip-vit-comer_b200/train.py (a torch.autograd.Function).
#include "common.cuh"

#include <math.h>

namespace wc {

// out[b][c][r] = alpha * in[b][r][c] (r < R, c < C): turns a row-major activation / gradient / weight into the K-major
// operand of the product that contracts over its rows.  Pad columns R..ld_out-1 are left untouched (callers zero them).
// 32 x 32 tiles through shared memory.
template <typename T>
__global__ void __launch_bounds__(256) transpose_f16_kernel(const T* __restrict__ in, h16* __restrict__ out, int R, int C,
                                                            long ld_in, long ld_out, long in_bs, long out_bs, float alpha) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z, r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const T* ip = in + (size_t)b * in_bs;
    h16* op = out + (size_t)b * out_bs;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = r0 + ty + 8 * i, c = c0 + tx;
        tile[ty + 8 * i][tx] = (r < R && c < C) ? alpha * (float)ip[(size_t)r * ld_in + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = c0 + ty + 8 * i, r = r0 + tx;
        if (c < C && r < R) op[(size_t)c * ld_out + r] = __float2half_rn(tile[tx][ty + 8 * i]);
    }
}

__global__ void __launch_bounds__(256) cast_f16_kernel(const float* __restrict__ in, h16* __restrict__ out, long n, float alpha,
                                                       float ref_scale) {
    const long i = (long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    float v = alpha * in[i];
    if (ref_scale > 0.f) v = __half2float(__float2half_rn(v / ref_scale)) * ref_scale;     // the reference's own fp16 rounding
    out[i] = __float2half_rn(v);
}

// LayerNorm backward (eps 1e-5), one warp per row, E = 32 * EPL:
//   dx = rstd * (g - mean(g) - xhat * mean(g * xhat)) [+ add],  g = gamma * dy ;  dgamma += sum_rows dy * xhat ;  dbeta += sum_rows dy
template <int EPL>
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ dy, const float* __restrict__ add,
                                                            float* __restrict__ dx, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta, long rows) {
    constexpr int E = 32 * EPL;
    __shared__ float sg[E], sb[E];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    for (int c = threadIdx.x; c < E; c += 256) { sg[c] = 0.f; sb[c] = 0.f; }
    __syncthreads();
    float pg[EPL], pb[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) { pg[i] = 0.f; pb[i] = 0.f; }
    for (long row = (long)blockIdx.x * 8 + wrp; row < rows; row += (long)gridDim.x * 8) {
        float v[EPL], d[EPL];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) { v[i] = x[row * E + lane + 32 * i]; s += v[i]; }
        const float mean = warp_sum(s) / E;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) { v[i] -= mean; q += v[i] * v[i]; }
        const float rstd = rsqrtf(warp_sum(q) / E + 1e-5f);
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            v[i] *= rstd;                                         // xhat
            d[i] = dy[row * E + lane + 32 * i];
            pg[i] += d[i] * v[i]; pb[i] += d[i];
            d[i] *= gamma[lane + 32 * i];                         // g
            c1 += d[i]; c2 += d[i] * v[i];
        }
        c1 = warp_sum(c1) / E; c2 = warp_sum(c2) / E;
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            float o = rstd * (d[i] - c1 - v[i] * c2);
            if (add) o += add[row * E + lane + 32 * i];
            dx[row * E + lane + 32 * i] = o;
        }
    }
#pragma unroll
    for (int i = 0; i < EPL; ++i) { atomicAdd(&sg[lane + 32 * i], pg[i]); atomicAdd(&sb[lane + 32 * i], pb[i]); }
    __syncthreads();
    for (int c = threadIdx.x; c < E; c += 256) { atomicAdd(dgamma + c, sg[c]); atomicAdd(dbeta + c, sb[c]); }
}

// Softmax backward per row: dS = P * (dP - sum_j P_j dP_j).  P h16 / dP fp32 / dS h16, rows of pitch Lp (pad columns -> 0).
__global__ void __launch_bounds__(256) softmax_bwd_kernel(const h16* __restrict__ P, const float* __restrict__ dP,
                                                          h16* __restrict__ dS, long rows, int L, int Lp) {
    const long row = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const h16* p = P + row * Lp;
    const float* d = dP + row * Lp;
    float s = 0.f;
    for (int c = lane; c < L; c += 32) s = fmaf(__half2float(p[c]), d[c], s);
    s = warp_sum(s);
    h16* o = dS + row * Lp;
    for (int c = lane; c < Lp; c += 32) o[c] = c < L ? __float2half_rn(__half2float(p[c]) * (d[c] - s)) : __float2half_rn(0.f);
}

// out[c] = sum_r x[r][c]  (bias gradients).  One CTA per 32 columns, 8 row groups.
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, float* __restrict__ out, long M, int N, long ld,
                                                     float alpha) {
    __shared__ float part[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, c = blockIdx.x * 32 + tx;
    float s = 0.f;
    if (c < N)
        for (long r = ty; r < M; r += 8) s += (float)x[r * ld + c];
    part[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < N) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += part[i][tx];
        out[c] = alpha * t;
    }
}

// Activation backward: out = dy * f'(.), kind 1: ReLU with aux = the activation output (mask aux > 0), kind 2: QuickGELU
// with aux = the pre-activation u: f'(u) = s (1 + 1.702 u (1 - s)), s = sigmoid(1.702 u)  (clip/model.py:186-188).
__global__ void __launch_bounds__(256) act_bwd_kernel(int kind, const float* __restrict__ dy, const h16* __restrict__ aux,
                                                      h16* __restrict__ out, long n, float alpha) {
    const long i = (long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float u = __half2float(aux[i]), g = alpha * dy[i];
    float r;
    if (kind == 1) r = u > 0.f ? g : 0.f;
    else { const float s = 1.f / (1.f + __expf(-1.702f * u)); r = g * s * (1.f + 1.702f * u * (1.f - s)); }
    out[i] = __float2half_rn(r);
}

// attn_pred = sigmoid(F F^T) (model_attn_aff_voc.py:134-137): with g = dA * s * (1 - s) the gradient to F is (g + g^T) F.
// G[n][i][j] = alpha * (g_ij + g_ji) as h16 [N, P, Pp] (pad columns zero): 32 x 32 tile pairs through shared memory.
__global__ void __launch_bounds__(256) sigmoid_gram_bwd_kernel(const float* __restrict__ dA, const float* __restrict__ S,
                                                               h16* __restrict__ G, int P, int Pp, float alpha) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z, i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const size_t base = (size_t)n * P * P;
#pragma unroll
    for (int k = 0; k < 4; ++k) {          // tile (j0.., i0..) of g, to be read transposed
        const int r = j0 + ty + 8 * k, c = i0 + tx;
        float v = 0.f;
        if (r < P && c < P) { const float s = S[base + (size_t)r * P + c]; v = dA[base + (size_t)r * P + c] * s * (1.f - s); }
        tile[ty + 8 * k][tx] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int i = i0 + ty + 8 * k, j = j0 + tx;
        if (i < P && j < Pp) {
            float v = 0.f;
            if (j < P) { const float s = S[base + (size_t)i * P + j]; v = dA[base + (size_t)i * P + j] * s * (1.f - s) + tile[tx][ty + 8 * k]; }
            G[((size_t)n * P + i) * Pp + j] = __float2half_rn(alpha * v);
        }
    }
}

// Dropout2d (segformer_head.py:79): y[n][p][c] = x[n][p][c] * mask[n][c] (mask holds 0 or 1/(1-p)); optional h16 copy.
__global__ void __launch_bounds__(256) dropout2d_kernel(const float* __restrict__ x, const float* __restrict__ mask,
                                                        float* __restrict__ y, h16* __restrict__ y_h, int P, int C, long n) {
    const long i = (long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const int c = (int)(i % C);
    const long img = i / ((long)P * C);
    const float v = x[i] * mask[img * C + c];
    if (y) y[i] = v;
    if (y_h) y_h[i] = __float2half_rn(v);
}

}  // namespace wc

using namespace wc;

extern "C" int wc_transpose_f16(const void* in, int in_f32, void* out, int batch, int R, int C, long ld_in, long ld_out,
                                long in_bs, long out_bs, float alpha, void* stream) {
    WC_CHECK_ARG(in && out && batch > 0 && batch <= 65535 && R > 0 && C > 0 && ld_in >= C && ld_out >= R);
    dim3 grid((unsigned)ceil_div(R, 32), (unsigned)ceil_div(C, 32), batch);
    WC_CHECK_ARG(grid.y <= 65535);
    if (in_f32)
        transpose_f16_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float*>(in), reinterpret_cast<h16*>(out),
                                                                         R, C, ld_in, ld_out, in_bs, out_bs, alpha);
    else
        transpose_f16_kernel<h16><<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<const h16*>(in), reinterpret_cast<h16*>(out),
                                                                       R, C, ld_in, ld_out, in_bs, out_bs, alpha);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_cast_f16(const float* in, void* out, long n, float alpha, float ref_scale, void* stream) {
    WC_CHECK_ARG(in && out && n > 0);
    cast_f16_kernel<<<(unsigned)ceil_div(n, 256L), 256, 0, as_stream(stream)>>>(in, reinterpret_cast<h16*>(out), n, alpha, ref_scale);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_layernorm_bwd(const float* x, const float* gamma, const float* dy, const float* add, float* dx,
                                float* dgamma, float* dbeta, long rows, int E, void* stream) {
    WC_CHECK_ARG(x && gamma && dy && dx && dgamma && dbeta && rows > 0);
    const unsigned grid = (unsigned)(ceil_div(rows, 8L) < 592 ? ceil_div(rows, 8L) : 592);
    cudaStream_t st = as_stream(stream);
    switch (E) {
        case 256: layernorm_bwd_kernel<8><<<grid, 256, 0, st>>>(x, gamma, dy, add, dx, dgamma, dbeta, rows); break;
        case 512: layernorm_bwd_kernel<16><<<grid, 256, 0, st>>>(x, gamma, dy, add, dx, dgamma, dbeta, rows); break;
        case 768: layernorm_bwd_kernel<24><<<grid, 256, 0, st>>>(x, gamma, dy, add, dx, dgamma, dbeta, rows); break;
        default: return fail(WC_ERR_UNSUPPORTED, "%s: LayerNorm width %ld not instantiated", __func__, (long)E);
    }
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_softmax_bwd(const void* P, const float* dP, void* dS, long rows, int L, int Lp, void* stream) {
    WC_CHECK_ARG(P && dP && dS && rows > 0 && L > 0 && Lp >= L);
    softmax_bwd_kernel<<<(unsigned)ceil_div(rows, 8L), 256, 0, as_stream(stream)>>>(reinterpret_cast<const h16*>(P), dP,
                                                                                    reinterpret_cast<h16*>(dS), rows, L, Lp);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_colsum(const void* x, int x_f16, float* out, long M, int N, long ld, float alpha, void* stream) {
    WC_CHECK_ARG(x && out && M > 0 && N > 0 && ld >= N);
    const unsigned grid = (unsigned)ceil_div(N, 32);
    if (x_f16) colsum_kernel<h16><<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<const h16*>(x), out, M, N, ld, alpha);
    else colsum_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float*>(x), out, M, N, ld, alpha);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_act_bwd(int kind, const float* dy, const void* aux, void* out, long n, float alpha, void* stream) {
    WC_CHECK_ARG((kind == 1 || kind == 2) && dy && aux && out && n > 0);
    act_bwd_kernel<<<(unsigned)ceil_div(n, 256L), 256, 0, as_stream(stream)>>>(kind, dy, reinterpret_cast<const h16*>(aux),
                                                                              reinterpret_cast<h16*>(out), n, alpha);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_sigmoid_gram_bwd(const float* dA, const float* S, void* G, int N, int P, int Pp, float alpha, void* stream) {
    WC_CHECK_ARG(dA && S && G && N > 0 && N <= 65535 && P > 0 && Pp >= P);
    dim3 grid((unsigned)ceil_div(Pp, 32), (unsigned)ceil_div(P, 32), N);
    sigmoid_gram_bwd_kernel<<<grid, 256, 0, as_stream(stream)>>>(dA, S, reinterpret_cast<h16*>(G), P, Pp, alpha);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_dropout2d(const float* x, const float* mask, float* y, void* y_f16, int N, int P, int C, void* stream) {
    WC_CHECK_ARG(x && mask && (y || y_f16) && N > 0 && P > 0 && C > 0);
    const long n = (long)N * P * C;
    dropout2d_kernel<<<(unsigned)ceil_div(n, 256L), 256, 0, as_stream(stream)>>>(x, mask, y, reinterpret_cast<h16*>(y_f16), P, C, n);
    WC_LAUNCH_CHECK();
    return WC_OK;
}
