// Fused multi-head attention forward for sm_100a (tcgen05 + TMEM + TMA), head_dim 64 (CLIP blocks) or 32 (decoder blocks).
// Replaces the materialised [N*heads, L, L] score tensor of clip/myAtt.py:53-64.
#pragma once
#include "common.cuh"

namespace wc {

// q, k: [N*heads, L, hd] fp16 (q pre-scaled by 1/sqrt(hd), myAtt.py:54).
// v_or_vt: hd = 64 -> V itself, [N*heads, L, 64] (used as an MN-major tensor-core operand); hd = 32 -> V transposed,
//          [N*heads, 32, Lp] fp16.
// o: [N*L, heads*hd] fp16, head-interleaved (myAtt.py:320).  lse2 (optional): [N*heads, L] fp32, log2-domain
// log-sum-exp of each score row (input of launch_attn_headmean).  hd in {64, 32}.
int launch_attention_fwd(const h16* q, const h16* k, const h16* v_or_vt, h16* o, float* lse2, int n_img, int heads, int L, int Lp,
                         int hd, cudaStream_t st);

// avg[n, i, j] = mean over heads of softmax(q k^T)[n, h, i, j] (myAtt.py:323-326), fp32 [N, L, L], each element written
// once.  Needs lse2 from launch_attention_fwd.
// sums (optional, [n_img] fp64, overwritten): sum of avg[n][1:, 1:] per image, accumulated while the tiles are in registers.
int launch_attn_headmean(const h16* q, const h16* k, const float* lse2, float* avg, int n_img, int heads, int L,
                         cudaStream_t st, double* sums = nullptr);

// Grad-CAM attention backward reduced to column sums (see attention.cu / gradcam.cu headers).
//   dOT: fp16 [n_inst*heads, 80, Lp] = rows 0..63 dO^T of the head, row 64 / 65 = -delta (hi / lo fp16 split), rest zero.
//   colsum: fp32 [n_inst*heads, L] = sum over the patch query rows of dS.  tab.img[inst] = image of each instance.
constexpr int BWD_MAX_INST = 64;
struct BwdInstTable { short img[BWD_MAX_INST]; };
int launch_attn_bwd_colsum(const h16* q, const h16* k, const h16* v, const h16* dOT, const float* lse2, float* colsum,
                           const BwdInstTable& tab, int n_inst, int n_img, int heads, int L, int Lp, cudaStream_t st);

}  // namespace wc
