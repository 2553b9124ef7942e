/*
 * weclip_b200.h — C ABI of libweclip_sm100.so: the B200 (sm_100a) drop-in for WeCLIP's per-image
 * inference hot path.  The reference (dayae1204/WeCLIP-ViT-CoMer) is pure Python/PyTorch and has no
 * FFI of its own; the "interface each entry point replaces" is therefore the Python function cited
 * beside it (paths relative to the reference root).  INTEGRATION.md shows the ctypes binding a
 * maintainer adds on the reference side.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - plain pointers and sizes only: no torch / C++ types cross this boundary;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises
 *     the device and nothing allocates device memory (callers pass workspaces sized by *_workspace_bytes);
 *   - return value: 0 = OK, negative = error (WC_ERR_*); wc_last_error() returns a thread-local message;
 *   - tensors are dense row-major with the shapes given; fp32 unless stated; "bf16" = __nv_bfloat16 bits.
 */
#ifndef WECLIP_B200_H_
#define WECLIP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WC_OK 0
#define WC_ERR_ARG (-1)
#define WC_ERR_CUDA (-2)
#define WC_ERR_WORKSPACE (-3)
#define WC_ERR_UNSUPPORTED (-4)

#define WC_ABI_VERSION 1
#define WC_PAR_MAX_DIL 8

/* ABI version (bumped whenever a signature changes) and last error text. */
int wc_abi_version(void);
const char* wc_last_error(void);
/* 1 when the current device is compute capability 10.x (the only target); 0 otherwise. */
int wc_device_ok(void);

/* ------------------------------------------------------------------------------------------------
 * PAR — pixel-adaptive refinement.  Replaces WeCLIP_model/PAR.py:64-92 (PAR.forward), incl.
 * get_dilated_neighbors :39-49 and get_pos :51-62.
 *
 * Ragged batch: image b owns mask channels [ch_off_host[b], ch_off_host[b+1]) of the packed
 * [sumKc,H,W] `masks` / `out` tensors (the reference calls PAR once per image with its own K+1).
 * imgs is [B,3,Hi,Wi]; it is bilinearly resized (align_corners=True, PAR.py:67) to HxW when the
 * sizes differ.  w1=0.3, w2=0.01 as PAR.py:36-37.
 * ---------------------------------------------------------------------------------------------- */
size_t wc_par_workspace_bytes(int B, int sum_kc, int H, int W, int n_dil);
int wc_par_fwd(const float* imgs, int Hi, int Wi, const float* masks, float* out,
               const int* ch_off_host, int B, int H, int W,
               const int* dilations_host, int n_dil, int num_iter,
               void* workspace, size_t workspace_bytes, void* stream);
/* The two halves of wc_par_fwd, exposed for the roofline micro-bench (BASELINE.json configs[4]):
 * affinity build (PAR.py:70-86) -> aff [B, 8*n_dil, H, W]; img must already be HxW. */
int wc_par_affinity(const float* img, float* aff, int B, int H, int W,
                    const int* dilations_host, int n_dil, void* stream);
/* num_iter propagation steps (PAR.py:88-90); scratch holds 2*sum_kc*H*W floats. */
int wc_par_propagate(const float* aff, const float* masks, float* out, float* scratch,
                     const int* ch_off_host, int B, int H, int W,
                     const int* dilations_host, int n_dil, int num_iter, void* stream);

/* argmax over each image's channels (first maximum wins, torch semantics) followed by the valid_key
 * LUT.  Replaces _refine_cams' tail, WeCLIP_model/model_attn_aff_voc.py:54-57.
 * lut is int64 [sumKc] (packed like the channels); labels is int64 [B,H,W]. */
int wc_argmax_lut(const float* masks, const int64_t* lut, int64_t* labels,
                  const int* ch_off_host, int B, int H, int W, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Tensor-core GEMM (tcgen05 + TMEM + TMA), bf16 inputs, fp32 accumulation:
 *     out[b] = act(alpha * A[b] @ B[b]^T + bias) (+ resid[b])
 * A [batch,M,K] and B [batch,N,K] are bf16, K contiguous (lda/ldb/a_bs/b_bs in elements, multiples of 8;
 * b_bs = 0 broadcasts B).  out is fp32 or bf16 (out_bf16), row-major [M,N] with row stride ldo, or
 * transposed [N,M] (out_trans: element (m,n) at n*ldo+m).  bias is fp32 [N] or NULL; resid is fp32 with the
 * layout of out (may alias a fp32 out) or NULL.  act: 0 none, 1 ReLU, 2 QuickGELU, 3 sigmoid.
 * This is the building block behind every nn.Linear / F.linear / bmm on the path (clip/model.py:198-213,
 * clip/myAtt.py:53-64,199-201,320-321, segformer_head.py:13-28,69-80, model_attn_aff_voc.py:134-137); it is
 * exported for the numerics tests and the tensor-pipe roofline micro-bench.
 * ---------------------------------------------------------------------------------------------- */
int wc_gemm_bf16(const void* A, const void* B, void* out, const float* bias, const float* resid,
                 int M, int N, int K, int batch, long lda, long ldb, long ldo, long a_bs, long b_bs, long o_bs,
                 int out_bf16, int out_trans, int act, float alpha, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WECLIP_B200_H_ */
