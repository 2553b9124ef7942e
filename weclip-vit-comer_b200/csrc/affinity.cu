// Attention-affinity CAM refinement + CAM post-processing for sm_100a (memory-bound kernels).
//
// Replaces clip/clip_tool.py:152-191 (RFM layer selection, compute_trans_mat :64-80, box mask via
// clip/utils.py:115-142 scoremap2bbox, masked transition) and :202-216 generate_cam_label +
// model_attn_aff_voc.py:160-163 (cv2 bilinear upsample, background channel).
//
// compute_trans_mat is evaluated in factored form.  With B the fused attention [P,P], the three
// (column, row) normalisation rounds only rescale rows and columns, T = diag(r) B diag(c), so each round is a
// matrix-vector product with B (r = 1/(B c), c = 1/(B^T r)); the symmetrised square is never materialised:
//     cam' = T_s (T_s v),  T_s x = (r * (B (c*x)) + c * (B^T (r*x))) / 2,  v = mask * cam
// which is algebraically (T_s T_s) v of clip_tool.py:73-76,186-191 with 4 passes over B instead of a P^3 GEMM.
#include "common.cuh"

#include <math.h>

namespace wc {

constexpr int AFF_MAX_IMG = 64;
constexpr int AFF_KMAX = 8;

struct AffTable {
    int n;
    int k_off[AFF_MAX_IMG + 1];
};

__device__ __forceinline__ const float* attn_row(const float* attn, int l, int img, int N, int L, int i) {
    return attn + (((size_t)l * N + img) * L + 1 + i) * L + 1;   // [:, 1:, 1:] (clip_tool.py:155)
}

// sums[img][l] = sum_ij W_l[i][j] (l < nl), sums[img][nl] = sum_ij attn_pred[i][j]
__global__ void __launch_bounds__(256) aff_layer_sums_kernel(const float* __restrict__ attn, const float* __restrict__ ap,
                                                             double* __restrict__ sums, int nl, int N, int L) {
    __shared__ float red[32];
    const int P = L - 1, img = blockIdx.z, l = blockIdx.y, r0 = blockIdx.x * 32, r1 = min(r0 + 32, P);
    float s = 0.f;
    for (int i = r0; i < r1; ++i) {
        const float* row = (l < nl) ? attn_row(attn, l, img, N, L, i) : ap + ((size_t)img * P + i) * P;
        for (int j = threadIdx.x; j < P; j += 256) s += row[j];
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0) atomicAdd(sums + (size_t)img * (nl + 1) + l, (double)s);
}

// The row loop of aff_build_kernel for NK kept layers (NK = 0: run-time count): 4 rows x (NK + 1) independent loads in flight per
// thread (the first version issued one row's loads at a time behind a per-layer branch and ran at 1.3-1.6 TB/s).
template <int NK>
__device__ __forceinline__ void aff_build_rows(const float* const (&base)[16], const float (&wk)[16], const float* __restrict__ ap,
                                               float* __restrict__ Bm, float* __restrict__ colz, int img, int P, int L, int r0,
                                               int r1, int use_rfm, int nk_rt = 0) {
    const int nk = NK ? NK : nk_rt;
    const size_t irow = (size_t)img * P;
    for (int j = threadIdx.x; j < P; j += 256) {
        float cs = 0.f;
        int i = r0;
        for (; i + 4 <= r1; i += 4) {
            float v[4] = {0.f, 0.f, 0.f, 0.f}, a[4] = {1.f, 1.f, 1.f, 1.f};
            if (NK) {
#pragma unroll
                for (int q = 0; q < (NK ? NK : 1); ++q) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) v[u] = fmaf(wk[q], __ldg(base[q] + (size_t)(i + u) * L + j), v[u]);
                }
            } else {
                for (int q = 0; q < nk; ++q) {
#pragma unroll
                    for (int u = 0; u < 4; ++u) v[u] = fmaf(wk[q], __ldg(base[q] + (size_t)(i + u) * L + j), v[u]);
                }
            }
            if (use_rfm) {
#pragma unroll
                for (int u = 0; u < 4; ++u) a[u] = __ldg(ap + (irow + i + u) * P + j);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (use_rfm) v[u] *= a[u];
                Bm[(irow + i + u) * P + j] = v[u];
                cs += v[u];
            }
        }
        for (; i < r1; ++i) {
            float v = 0.f;
            for (int q = 0; q < nk; ++q) v = fmaf(wk[q], __ldg(base[q] + (size_t)i * L + j), v);
            if (use_rfm) v *= ap[(irow + i) * P + j];
            Bm[(irow + i) * P + j] = v;
            cs += v;
        }
        (void)cs; (void)colz;      // (the first column sums come from aff_col_pass_kernel: no floating-point atomics on the path)
    }
}

// B = RFM fusion (clip_tool.py:158-170) or the plain layer mean (:173-174); also the first column sums.
__global__ void __launch_bounds__(256) aff_build_kernel(const float* __restrict__ attn, const float* __restrict__ ap,
                                                        const double* __restrict__ sums, float* __restrict__ Bm,
                                                        float* __restrict__ colz, int nl, int N, int L, int use_rfm,
                                                        float* __restrict__ sel_out, int sums_layer_major) {
    __shared__ float wl[16];
    const int P = L - 1, img = blockIdx.y, r0 = blockIdx.x * 16, r1 = min(r0 + 16, P);
    if (threadIdx.x == 0) {
        if (use_rfm) {
            // attn_diff_l = sum(seg_attn - W_l); keep layers with diff <= mean(diff).  sum(seg_attn) is the same for every
            // layer and cancels in diff_l - mean(diff), so only the layer sums S_l matter: d_l - mean = mean(S) - S_l
            // (sums: [img][nl+1] from aff_layer_sums_kernel, or [nl][N] accumulated by the head-mean kernel)
            float d[16], mean = 0.f;
            double sbar = 0.0;
            for (int l = 0; l < nl; ++l) sbar += sums_layer_major ? sums[(size_t)l * N + img] : sums[(size_t)img * (nl + 1) + l];
            sbar /= (double)nl;
            for (int l = 0; l < nl; ++l)
                d[l] = (float)(sbar - (sums_layer_major ? sums[(size_t)l * N + img] : sums[(size_t)img * (nl + 1) + l]));
            float cnt = 0.f;
            for (int l = 0; l < nl; ++l) { wl[l] = (d[l] <= mean) ? 1.f : 0.f; cnt += wl[l]; }
            if (sel_out && blockIdx.x == 0)        // test hook: which layers were kept, and the margins they were kept by
                for (int l = 0; l < nl; ++l) { sel_out[((size_t)img * 2) * nl + l] = wl[l]; sel_out[((size_t)img * 2 + 1) * nl + l] = d[l] - mean; }
            for (int l = 0; l < nl; ++l) wl[l] /= (cnt + 1e-5f);
        } else {
            for (int l = 0; l < nl; ++l) wl[l] = 1.f / (float)nl;
            if (sel_out && blockIdx.x == 0)
                for (int l = 0; l < nl; ++l) { sel_out[((size_t)img * 2) * nl + l] = 1.f; sel_out[((size_t)img * 2 + 1) * nl + l] = 0.f; }
        }
    }
    __syncthreads();
    // kept layers, in ascending order (the fmaf chain below adds them in the same order as a loop over all layers would)
    int nk = 0;
    const float* base[16];
    float wk[16];
    for (int l = 0; l < nl; ++l)
        if (wl[l] != 0.f) { base[nk] = attn_row(attn, l, img, N, L, 0); wk[nk] = wl[l]; ++nk; }      // dropped layers are not read
    switch (nk) {
        case 1: aff_build_rows<1>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm); break;
        case 2: aff_build_rows<2>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm); break;
        case 3: aff_build_rows<3>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm); break;
        case 4: aff_build_rows<4>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm); break;
        case 5: aff_build_rows<5>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm); break;
        case 6: aff_build_rows<6>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm); break;
        case 8: aff_build_rows<8>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm); break;
        default: aff_build_rows<0>(base, wk, ap, Bm, colz, img, P, L, r0, r1, use_rfm, nk); break;
    }
}

// r_i = 1 / sum_j B_ij / z_j      (row normalisation with c = 1/z)
__global__ void __launch_bounds__(256) aff_row_pass_kernel(const float* __restrict__ Bm, const float* __restrict__ z,
                                                           float* __restrict__ r, int P) {
    const int img = blockIdx.y, i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= P) return;
    const float* row = Bm + ((size_t)img * P + i) * P;
    const float* zz = z + (size_t)img * P;
    float s = 0.f;
    float s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if ((P & 3) == 0) {      // 16-byte loads, two per lane in flight (rows start 16-byte aligned when P % 4 == 0)
        const float4* row4 = reinterpret_cast<const float4*>(row);
        const float4* z4 = reinterpret_cast<const float4*>(zz);
        const int n4 = P >> 2;
        int j = lane;
        for (; j + 32 < n4; j += 64) {
            const float4 b0 = __ldg(row4 + j), b1 = __ldg(row4 + j + 32), q0 = __ldg(z4 + j), q1 = __ldg(z4 + j + 32);
            s = fmaf(b0.x, __frcp_rn(q0.x), s); s1 = fmaf(b0.y, __frcp_rn(q0.y), s1);
            s2 = fmaf(b0.z, __frcp_rn(q0.z), s2); s3 = fmaf(b0.w, __frcp_rn(q0.w), s3);
            s = fmaf(b1.x, __frcp_rn(q1.x), s); s1 = fmaf(b1.y, __frcp_rn(q1.y), s1);
            s2 = fmaf(b1.z, __frcp_rn(q1.z), s2); s3 = fmaf(b1.w, __frcp_rn(q1.w), s3);
        }
        for (; j < n4; j += 32) {
            const float4 b0 = __ldg(row4 + j), q0 = __ldg(z4 + j);
            s = fmaf(b0.x, __frcp_rn(q0.x), s); s1 = fmaf(b0.y, __frcp_rn(q0.y), s1);
            s2 = fmaf(b0.z, __frcp_rn(q0.z), s2); s3 = fmaf(b0.w, __frcp_rn(q0.w), s3);
        }
        s = warp_sum((s + s1) + (s2 + s3));
        if (lane == 0) r[(size_t)img * P + i] = 1.f / s;
        return;
    }
    int j = lane;
    for (; j + 96 < P; j += 128) {          // four independent streams; 1/z by reciprocal + multiply (<= 1 ulp from the division)
        s = fmaf(row[j], __frcp_rn(zz[j]), s);
        s1 = fmaf(row[j + 32], __frcp_rn(zz[j + 32]), s1);
        s2 = fmaf(row[j + 64], __frcp_rn(zz[j + 64]), s2);
        s3 = fmaf(row[j + 96], __frcp_rn(zz[j + 96]), s3);
    }
    for (; j < P; j += 32) s = fmaf(row[j], __frcp_rn(zz[j]), s);
    s = warp_sum((s + s1) + (s2 + s3));
    if (lane == 0) r[(size_t)img * P + i] = 1.f / s;
}

// z_j = sum_i r_i B_ij      (column sums of diag(r) B; r == nullptr: plain column sums)
// One CTA owns a strip of 32 columns over ALL rows: lane = column, the 8 warps interleave the rows, eight 128-byte row segments
// per warp in flight, and the eight partial sums are added in a fixed order — every z_j is the same bits on every run (the first
// version added row-chunk partials with atomicAdd, whose order is not fixed).
__global__ void __launch_bounds__(256) aff_col_pass_kernel(const float* __restrict__ Bm, const float* __restrict__ r,
                                                           float* __restrict__ z, int P) {
    __shared__ float part[8][32];
    const int img = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, j = blockIdx.x * 32 + lane;
    const bool ok = j < P;
    const float* col = Bm + (size_t)img * P * P + (ok ? j : 0);
    const float* rr = r ? r + (size_t)img * P : nullptr;
    float s0 = 0.f, s1 = 0.f;
    int i = warp;
    for (; i + 56 < P; i += 64) {
        float b[8], w[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            b[u] = __ldg(col + (size_t)(i + 8 * u) * P);
            w[u] = rr ? __ldg(rr + i + 8 * u) : 1.f;
        }
#pragma unroll
        for (int u = 0; u < 8; u += 2) { s0 = fmaf(w[u], b[u], s0); s1 = fmaf(w[u + 1], b[u + 1], s1); }
    }
    for (; i < P; i += 8) s0 = fmaf(rr ? __ldg(rr + i) : 1.f, __ldg(col + (size_t)i * P), s0);
    part[warp][lane] = s0 + s1;
    __syncthreads();
    if (warp == 0 && ok) {
        float t = part[0][lane];
#pragma unroll
        for (int w8 = 1; w8 < 8; ++w8) t += part[w8][lane];
        z[(size_t)img * P + j] = t;
    }
}

// ---------------------------------------------------------------------------------------------------
// Box mask (clip/utils.py:115-142 + clip_tool.py:179-185), one CTA per class instance, grid <= 64x64 cells:
// uint8 quantisation, strict '>' threshold against int(thr*max), 8-connected component labelling by iterated
// min-label propagation in shared memory, per-component bounding box, cv2's min(x+w, W-1) clip and the
// exclusive-upper slice fill.  Output v = mask * cam (the masked CAM that multiplies the transition matrix).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) box_mask_kernel(const float* __restrict__ cams, float* __restrict__ v,
                                                       float* __restrict__ mask_out, int h, int w, double thr) {
    extern __shared__ int sm[];
    const int P = h * w, tid = threadIdx.x, inst = blockIdx.x;
    int* label = sm;                 // [P]
    int* x0 = sm + P; int* x1 = x0 + P; int* y0 = x1 + P; int* y1 = y0 + P;
    unsigned char* msk = reinterpret_cast<unsigned char*>(y1 + P);   // [P]
    __shared__ int s_max, s_changed;
    const float* cam = cams + (size_t)inst * P;
    if (tid == 0) s_max = 0;
    __syncthreads();
    int lmax = 0;
    for (int p = tid; p < P; p += 256) lmax = max(lmax, (int)(unsigned char)(int)(cam[p] * 255.f));
    atomicMax(&s_max, lmax);
    __syncthreads();
    const int t = (int)(thr * (double)s_max);
    for (int p = tid; p < P; p += 256) {
        const int u8 = (int)(unsigned char)(int)(cam[p] * 255.f);
        label[p] = (u8 > t) ? p : -1;
        x0[p] = w; x1[p] = -1; y0[p] = h; y1[p] = -1; msk[p] = 0;
    }
    __syncthreads();
    for (int iter = 0; iter < 4 * P; ++iter) {
        if (tid == 0) s_changed = 0;
        __syncthreads();
        for (int p = tid; p < P; p += 256) {
            int l = label[p];
            if (l < 0) continue;
            const int y = p / w, x = p - y * w;
            int m = l;
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    const int yy = y + dy, xx = x + dx;
                    if (yy < 0 || yy >= h || xx < 0 || xx >= w) continue;
                    const int q = label[yy * w + xx];
                    if (q >= 0) m = min(m, q);
                }
            m = min(m, label[m]);           // pointer jumping
            if (m < l) { label[p] = m; s_changed = 1; }
        }
        __syncthreads();
        if (!s_changed) break;
        __syncthreads();
    }
    for (int p = tid; p < P; p += 256) {
        const int l = label[p];
        if (l < 0) continue;
        const int y = p / w, x = p - y * w;
        atomicMin(&x0[l], x); atomicMax(&x1[l], x); atomicMin(&y0[l], y); atomicMax(&y1[l], y);
    }
    __syncthreads();
    for (int p = tid; p < P; p += 256) {
        if (label[p] != p) continue;   // component root
        const int xa = x0[p], ya = y0[p], xb = min(x1[p] + 1, w - 1), yb = min(y1[p] + 1, h - 1);
        for (int y = ya; y < yb; ++y)
            for (int x = xa; x < xb; ++x) msk[y * w + x] = 1;
    }
    __syncthreads();
    for (int p = tid; p < P; p += 256) {
        const float m = msk[p] ? 1.f : 0.f;
        v[(size_t)inst * P + p] = m * cam[p];
        if (mask_out) mask_out[(size_t)inst * P + p] = m;
    }
}

// out_ik = 0.5 * r_i * sum_j B_ij (x_jk / z_j)          (row half of T_s x; writes)
__global__ void __launch_bounds__(256) tsym_row_kernel(const float* __restrict__ Bm, const float* __restrict__ r,
                                                       const float* __restrict__ z, const float* __restrict__ x,
                                                       float* __restrict__ out, const __grid_constant__ AffTable tab,
                                                       int P, int kbase) {
    const int img = blockIdx.y, i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int k0 = tab.k_off[img] + kbase, kn = min(AFF_KMAX, tab.k_off[img + 1] - k0);
    if (i >= P || kn <= 0) return;
    const float* row = Bm + ((size_t)img * P + i) * P;
    const float* zz = z + (size_t)img * P;
    float acc[AFF_KMAX];
#pragma unroll
    for (int k = 0; k < AFF_KMAX; ++k) acc[k] = 0.f;
    if ((P & 3) == 0) {      // 16-byte loads (B row, z, x): two per lane in flight
        const float4* row4 = reinterpret_cast<const float4*>(row);
        const float4* z4 = reinterpret_cast<const float4*>(zz);
        const int n4 = P >> 2;
        auto scaled = [&](int jj) {
            float4 b = __ldg(row4 + jj);
            const float4 q = __ldg(z4 + jj);
            b.x *= __frcp_rn(q.x); b.y *= __frcp_rn(q.y); b.z *= __frcp_rn(q.z); b.w *= __frcp_rn(q.w);
            return b;
        };
        int j = lane;
        for (; j + 32 < n4; j += 64) {
            const float4 b0 = scaled(j), b1 = scaled(j + 32);
#pragma unroll
            for (int k = 0; k < AFF_KMAX; ++k)
                if (k < kn) {
                    const float4* xk = reinterpret_cast<const float4*>(x + (size_t)(k0 + k) * P);
                    const float4 a0 = __ldg(xk + j), a1 = __ldg(xk + j + 32);
                    float t0 = fmaf(b0.x, a0.x, fmaf(b0.y, a0.y, fmaf(b0.z, a0.z, b0.w * a0.w)));
                    float t1 = fmaf(b1.x, a1.x, fmaf(b1.y, a1.y, fmaf(b1.z, a1.z, b1.w * a1.w)));
                    acc[k] += t0 + t1;
                }
        }
        for (; j < n4; j += 32) {
            const float4 b0 = scaled(j);
#pragma unroll
            for (int k = 0; k < AFF_KMAX; ++k)
                if (k < kn) {
                    const float4 a0 = __ldg(reinterpret_cast<const float4*>(x + (size_t)(k0 + k) * P) + j);
                    acc[k] += fmaf(b0.x, a0.x, fmaf(b0.y, a0.y, fmaf(b0.z, a0.z, b0.w * a0.w)));
                }
        }
        const float ri = r[(size_t)img * P + i];
#pragma unroll
        for (int k = 0; k < AFF_KMAX; ++k) {
            if (k < kn) {
                const float sk = warp_sum(acc[k]);
                if (lane == 0) out[(size_t)(k0 + k) * P + i] = 0.5f * ri * sk;
            }
        }
        return;
    }
    // four independent column groups per trip (loads of the next groups in flight under the FMAs of the first)
    int j = lane;
    for (; j + 96 < P; j += 128) {
        float b[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) b[u] = row[j + 32 * u] * __frcp_rn(zz[j + 32 * u]);
#pragma unroll
        for (int k = 0; k < AFF_KMAX; ++k)
            if (k < kn) {
                const float* xk = x + (size_t)(k0 + k) * P + j;
                acc[k] = fmaf(b[0], xk[0], fmaf(b[1], xk[32], fmaf(b[2], xk[64], fmaf(b[3], xk[96], acc[k]))));
            }
    }
    for (; j < P; j += 32) {
        const float b = row[j] * __frcp_rn(zz[j]);
#pragma unroll
        for (int k = 0; k < AFF_KMAX; ++k)
            if (k < kn) acc[k] = fmaf(b, x[(size_t)(k0 + k) * P + j], acc[k]);
    }
    const float ri = r[(size_t)img * P + i];
#pragma unroll
    for (int k = 0; k < AFF_KMAX; ++k) {
        if (k < kn) {
            const float s = warp_sum(acc[k]);
            if (lane == 0) out[(size_t)(k0 + k) * P + i] = 0.5f * ri * s;
        }
    }
}

// out_jk += 0.5 / z_j * sum_i B_ij r_i x_ik   (column half of T_s x; adds to the row half written by tsym_row_kernel)
// Same strip scheme as aff_col_pass_kernel (32 columns x all rows per CTA, fixed summation order, one writer per element).
__global__ void __launch_bounds__(256) tsym_col_kernel(const float* __restrict__ Bm, const float* __restrict__ r,
                                                       const float* __restrict__ z, const float* __restrict__ x,
                                                       float* __restrict__ out, const __grid_constant__ AffTable tab,
                                                       int P, int kbase) {
    __shared__ float part[8][AFF_KMAX][32];
    const int img = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, j = blockIdx.x * 32 + lane;
    const int k0 = tab.k_off[img] + kbase, kn = min(AFF_KMAX, tab.k_off[img + 1] - k0);
    if (kn <= 0) return;
    const bool ok = j < P;
    const float* col = Bm + (size_t)img * P * P + (ok ? j : 0);
    const float* rr = r + (size_t)img * P;
    float acc[AFF_KMAX];
#pragma unroll
    for (int k = 0; k < AFF_KMAX; ++k) acc[k] = 0.f;
    int i = warp;
    for (; i + 24 < P; i += 32) {
        float b[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) b[u] = __ldg(col + (size_t)(i + 8 * u) * P) * __ldg(rr + i + 8 * u);
#pragma unroll
        for (int k = 0; k < AFF_KMAX; ++k)
            if (k < kn) {
                const float* xk = x + (size_t)(k0 + k) * P + i;       // warp-uniform addresses: broadcast loads
                acc[k] = fmaf(b[0], __ldg(xk), fmaf(b[1], __ldg(xk + 8), fmaf(b[2], __ldg(xk + 16), fmaf(b[3], __ldg(xk + 24), acc[k]))));
            }
    }
    for (; i < P; i += 8) {
        const float b = __ldg(col + (size_t)i * P) * __ldg(rr + i);
#pragma unroll
        for (int k = 0; k < AFF_KMAX; ++k)
            if (k < kn) acc[k] = fmaf(b, __ldg(x + (size_t)(k0 + k) * P + i), acc[k]);
    }
#pragma unroll
    for (int k = 0; k < AFF_KMAX; ++k) part[warp][k][lane] = acc[k];
    __syncthreads();
    if (ok) {
        const float cj = 0.5f / z[(size_t)img * P + j];
        for (int k = warp; k < kn; k += 8) {
            float t = part[0][k][lane];
#pragma unroll
            for (int w8 = 1; w8 < 8; ++w8) t += part[w8][k][lane];
            out[(size_t)(k0 + k) * P + j] += cj * t;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// generate_cam_label (clip_tool.py:202-216): per class min-max (scale_cam_image) then cv2.resize INTER_LINEAR
// (half-pixel centres, border taps clamped with zero weight), and the background channel
// bg = (1 - max_k)^cam_bg_thres in front (model_attn_aff_voc.py:162-163).  Output packed [sumK + N, Hc, Wc]:
// image i owns channels [k_off[i]+i, k_off[i+1]+i+1): bg first, then its K classes.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cam_minmax_kernel(const float* __restrict__ cams, float* __restrict__ stats, int P) {
    __shared__ float red[32];
    const int inst = blockIdx.x;
    const float* c = cams + (size_t)inst * P;
    float mn = INFINITY;
    for (int p = threadIdx.x; p < P; p += 256) mn = fminf(mn, c[p]);
    mn = block_min(mn, red);
    float mx = -INFINITY;
    for (int p = threadIdx.x; p < P; p += 256) mx = fmaxf(mx, c[p] - mn);
    mx = block_max(mx, red);
    if (threadIdx.x == 0) { stats[2 * inst] = mn; stats[2 * inst + 1] = 1e-7f + mx; }
}

__global__ void __launch_bounds__(256) cam_upsample_bg_kernel(const float* __restrict__ cams, const float* __restrict__ stats,
                                                              float* __restrict__ out, const __grid_constant__ AffTable tab,
                                                              int h, int w, int Hc, int Wc, float bg_pow) {
    const int X = blockIdx.x * 32 + threadIdx.x, Y = blockIdx.y * 8 + threadIdx.y, img = blockIdx.z;
    if (X >= Wc || Y >= Hc) return;
    const int P = h * w;
    // cv2 resize.cpp: fx = (float)((dx+0.5)*scale - 0.5); sx = floor(fx); fx -= sx; clamp with zero weight
    float fx = (float)(((double)X + 0.5) * ((double)w / (double)Wc) - 0.5);
    float fy = (float)(((double)Y + 0.5) * ((double)h / (double)Hc) - 0.5);
    int sx = (int)floorf(fx), sy = (int)floorf(fy);
    fx -= sx; fy -= sy;
    if (sx < 0) { sx = 0; fx = 0.f; }
    if (sx >= w - 1) { sx = w - 1; fx = 0.f; }
    if (sy < 0) { sy = 0; fy = 0.f; }
    if (sy >= h - 1) { sy = h - 1; fy = 0.f; }
    const int sx1 = min(sx + 1, w - 1), sy1 = min(sy + 1, h - 1);
    const int k0 = tab.k_off[img], K = tab.k_off[img + 1] - k0;
    const size_t hw = (size_t)Hc * Wc;
    float* o = out + (size_t)(k0 + img) * hw + (size_t)Y * Wc + X;
    float mx = -INFINITY;
    for (int k = 0; k < K; ++k) {
        const float* c = cams + (size_t)(k0 + k) * P;
        const float mn = stats[2 * (k0 + k)], den = stats[2 * (k0 + k) + 1];
        const float v00 = (c[sy * w + sx] - mn) / den, v01 = (c[sy * w + sx1] - mn) / den;
        const float v10 = (c[sy1 * w + sx] - mn) / den, v11 = (c[sy1 * w + sx1] - mn) / den;
        const float r0 = v00 * (1.f - fx) + v01 * fx, r1 = v10 * (1.f - fx) + v11 * fx;
        const float val = r0 * (1.f - fy) + r1 * fy;
        o[(size_t)(k + 1) * hw] = val;
        mx = fmaxf(mx, val);
    }
    o[0] = (bg_pow == 1.f) ? (1.f - mx) : powf(1.f - mx, bg_pow);
}

}  // namespace wc

using namespace wc;

extern "C" size_t wc_affinity_workspace_bytes(int N, int h, int w, int total_inst, int nl) {
    const size_t P = (size_t)h * w;
    return sizeof(float) * ((size_t)N * P * P + 3 * (size_t)N * P + 2 * (size_t)total_inst * P) +
           sizeof(double) * (size_t)N * (nl + 1) + 2048;
}

extern "C" int wc_affinity_refine(const float* attn, int nl, const double* attn_sums, const float* attn_pred, int N, int h, int wd, int use_rfm,
                                  const float* cams, const int* k_off_host, double box_thr, float* cams_refined,
                                  float* box_masks, float* rfm_sel, void* workspace, size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(attn && cams && k_off_host && cams_refined && workspace && nl >= 1 && nl <= 16);
    WC_CHECK_ARG(N > 0 && N <= AFF_MAX_IMG && h > 0 && wd > 0 && (!use_rfm || attn_pred));
    cudaStream_t st = as_stream(stream);
    ProfScope prof(PK_AFFINITY, 0.0, st);
    const int P = h * wd, L = P + 1, total = k_off_host[N];
    WC_CHECK_ARG((size_t)P * 21 + 64 <= 200 * 1024);
    if (workspace_bytes < wc_affinity_workspace_bytes(N, h, wd, total, nl))
        return fail(WC_ERR_WORKSPACE, "%s: workspace too small (%ld bytes)", __func__, (long)workspace_bytes);
    AffTable tab; tab.n = N;
    int max_k = 0;
    for (int i = 0; i <= N; ++i) tab.k_off[i] = k_off_host[i];
    for (int i = 0; i < N; ++i) max_k = max(max_k, k_off_host[i + 1] - k_off_host[i]);
    uint8_t* wsb = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    double* sums = reinterpret_cast<double*>(wsb);
    float* Bm = reinterpret_cast<float*>(wsb + (((size_t)N * (nl + 1) * sizeof(double) + 255) & ~(size_t)255));
    float* z = Bm + (size_t)N * P * P;
    float* r = z + (size_t)N * P;
    float* v0 = r + (size_t)N * P;
    float* v1 = v0 + (size_t)total * P;
    if (use_rfm && !attn_sums) {      // no sums from the producers: one extra pass over the nl maps
        WC_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * N * (nl + 1), st));
        aff_layer_sums_kernel<<<dim3(ceil_div(P, 32), nl, N), 256, 0, st>>>(attn, attn_pred, sums, nl, N, L);
        WC_LAUNCH_CHECK();
    }
    aff_build_kernel<<<dim3(ceil_div(P, 16), N), 256, 0, st>>>(attn, attn_pred, attn_sums ? attn_sums : sums, Bm, z, nl, N, L, use_rfm, rfm_sel,
                                                                attn_sums ? 1 : 0);
    WC_LAUNCH_CHECK();
    for (int round = 0; round < 3; ++round) {   // clip_tool.py:67-72
        aff_col_pass_kernel<<<dim3(ceil_div(P, 32), N), 256, 0, st>>>(Bm, round ? r : nullptr, z, P);
        WC_LAUNCH_CHECK();
        aff_row_pass_kernel<<<dim3(ceil_div(P, 8), N), 256, 0, st>>>(Bm, z, r, P);
        WC_LAUNCH_CHECK();
    }
    // box masks -> v0 = mask * cam
    const size_t box_smem = (size_t)P * 21 + 16;
    WC_SET_MAX_SMEM(box_mask_kernel, 200 * 1024);
    box_mask_kernel<<<total, 256, box_smem, st>>>(cams, v0, box_masks, h, wd, box_thr);
    WC_LAUNCH_CHECK();
    // cam' = T_s (T_s v0)
    const float* src = v0;
    float* dsts[2] = {v1, cams_refined};
    for (int pass = 0; pass < 2; ++pass) {
        for (int kb = 0; kb < max_k; kb += AFF_KMAX) {
            tsym_row_kernel<<<dim3(ceil_div(P, 8), N), 256, 0, st>>>(Bm, r, z, src, dsts[pass], tab, P, kb);
            WC_LAUNCH_CHECK();
            tsym_col_kernel<<<dim3(ceil_div(P, 32), N), 256, 0, st>>>(Bm, r, z, src, dsts[pass], tab, P, kb);
            WC_LAUNCH_CHECK();
        }
        src = v1;
    }
    return WC_OK;
}

extern "C" int wc_cam_upsample_bg(const float* cams, const int* k_off_host, int N, int h, int wd, int Hc, int Wc,
                                  float bg_pow, float* out, float* stats_ws, void* stream) {
    WC_CHECK_ARG(cams && k_off_host && out && stats_ws && N > 0 && N <= AFF_MAX_IMG && h > 0 && wd > 0 && Hc > 0 && Wc > 0);
    cudaStream_t st = as_stream(stream);
    ProfScope prof(PK_CAM_UPSAMPLE, 0.0, st);
    AffTable tab; tab.n = N;
    for (int i = 0; i <= N; ++i) tab.k_off[i] = k_off_host[i];
    const int total = k_off_host[N];
    cam_minmax_kernel<<<total, 256, 0, st>>>(cams, stats_ws, h * wd);
    WC_LAUNCH_CHECK();
    cam_upsample_bg_kernel<<<dim3(ceil_div(Wc, 32), ceil_div(Hc, 8), N), dim3(32, 8), 0, st>>>(cams, stats_ws, out, tab, h, wd,
                                                                                              Hc, Wc, bg_pow);
    WC_LAUNCH_CHECK();
    return WC_OK;
}
