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
 *   - tensors are dense row-major with the shapes given; fp32 unless stated; "h16" = IEEE fp16 (__half) bits,
 *     the 16-bit tensor-core operand type of this library (fp32 accumulation everywhere).
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

#define WC_ABI_VERSION 13
#define WC_PAR_MAX_DIL 8

/* ABI version (bumped whenever a signature changes) and last error text. */
int wc_abi_version(void);
const char* wc_last_error(void);
/* 1 when the current device is compute capability 10.x (the only target); 0 otherwise. */
int wc_device_ok(void);

/* Opt-in per-kernel-class device timing (no reference counterpart: the reference has no profiling).
 * While enabled, every launch is bracketed by CUDA events on its stream; wc_prof_summary synchronises the device
 * and sums device milliseconds, algorithmic work (flops for gemm_tc, bytes for the memory-bound classes) and
 * launch counts per class into arrays of wc_prof_kinds() entries.  bench.py uses it for the roofline object. */
int wc_prof_kinds(void);
const char* wc_prof_kind_name(int kind);
int wc_prof_enable(int on);
int wc_prof_reset(void);
int wc_prof_summary(double* ms, double* work, long* launches);
/* Raw per-launch records (call after wc_prof_summary): kind, device ms, work, 4 tags (GEMM: M, N, K, batch). */
long wc_prof_records(int* kind, double* ms, double* work, int* tags, long cap);

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
/* aff_f16: storage type of the materialised affinity inside the workspace.  0 = fp32 (exact path: results within 2e-5
 * of PAR.py in fp32); 1 = fp16 (pipeline path: halves the bytes each propagation step streams; pseudo-labels agree
 * with the fp32 path on >= 99.5 % of pixels, masks within 5e-3). */
int wc_par_fwd(const float* imgs, int Hi, int Wi, const float* masks, float* out,
               const int* ch_off_host, int B, int H, int W,
               const int* dilations_host, int n_dil, int num_iter, int aff_f16,
               void* workspace, size_t workspace_bytes, void* stream);
/* The two halves of wc_par_fwd, exposed for the roofline micro-bench (BASELINE.json configs[4]) and the stage tests:
 * affinity build (PAR.py:70-86) -> aff fp32 or fp16 (aff_f16) in the BLOCKED layout [B, H, Wb, 8*n_dil, 64] with
 * Wb = wc_par_aff_pitch(W) / 64 = ceil(W/64): element (b, n, y, x) lives at
 * (((b*H + y)*Wb + x/64)*8*n_dil + n)*64 + x%64 (B*H*wc_par_aff_pitch(W)*8*n_dil elements in all).
 * img must already be HxW. */
int wc_par_aff_pitch(int W);
int wc_par_affinity(const float* img, void* aff, int aff_f16, int B, int H, int W,
                    const int* dilations_host, int n_dil, void* stream);
/* num_iter propagation steps (PAR.py:88-90); scratch holds 2*sum_kc*H*round_up(W,8) floats
 * (2*sum_kc*H*wc_par_aff_pitch(W) is always enough). */
int wc_par_propagate(const void* aff, int aff_f16, const float* masks, float* out, float* scratch,
                     const int* ch_off_host, int B, int H, int W,
                     const int* dilations_host, int n_dil, int num_iter, void* stream);

/* argmax over each image's channels (first maximum wins, torch semantics) followed by the valid_key
 * LUT.  Replaces _refine_cams' tail, WeCLIP_model/model_attn_aff_voc.py:54-57.
 * lut is int64 [sumKc] (packed like the channels); labels is int64 [B,H,W]. */
int wc_argmax_lut(const float* masks, const int64_t* lut, int64_t* labels,
                  const int* ch_off_host, int B, int H, int W, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Tensor-core GEMM (tcgen05 + TMEM + TMA), h16 inputs, fp32 accumulation:
 *     out[b] = act(alpha * A[b] @ B[b]^T + bias) (+ resid[b])
 * A [batch,M,K] and B [batch,N,K] are h16, K contiguous (lda/ldb/a_bs/b_bs in elements, multiples of 8;
 * b_bs = 0 broadcasts B).  out is fp32 or h16 (out_f16), row-major [M,N] with row stride ldo, or
 * transposed [N,M] (out_trans: element (m,n) at n*ldo+m).  bias is fp32 [N] or NULL; resid is fp32 with the
 * layout of out (may alias a fp32 out) or NULL.  act: 0 none, 1 ReLU, 2 QuickGELU, 3 sigmoid.
 * This is the building block behind every nn.Linear / F.linear / bmm on the path (clip/model.py:198-213,
 * clip/myAtt.py:53-64,199-201,320-321, segformer_head.py:13-28,69-80, model_attn_aff_voc.py:134-137); it is
 * exported for the numerics tests and the tensor-pipe roofline micro-bench.
 * ---------------------------------------------------------------------------------------------- */
int wc_gemm_f16(const void* A, const void* B, void* out, const float* bias, const float* resid,
                 int M, int N, int K, int batch, long lda, long ldb, long ldo, long a_bs, long b_bs, long o_bs,
                 int out_f16, int out_trans, int act, float alpha, void* stream);

/* The same product with TWO batch dimensions (b = b0 + nb0*b1; e.g. heads x images of an attention product), each with
 * its own element strides for A, B, out / resid and the bias vector. */
int wc_gemm_f16_ex(const void* A, const void* B, void* out, const float* bias, const float* resid,
                    int M, int N, int K, int nb0, int nb1, long lda, long ldb, long ldo,
                    long a_bs0, long a_bs1, long b_bs0, long b_bs1, long o_bs0, long o_bs1,
                    long bias_bs0, long bias_bs1,
                    int out_f16, int out_trans, int act, float alpha, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Weights.  "h16" pointers are h16 row-major [out_features, in_features] (nn.Linear layout,
 * K contiguous = the GEMM's B operand); biases / LayerNorm parameters are fp32.
 * ---------------------------------------------------------------------------------------------- */
typedef struct wc_block_weights {           /* one ResidualAttentionBlock, clip/model.py:191-214 */
    const void* w_in;   const float* b_in;  /* attn.in_proj_weight [3E,E], in_proj_bias [3E] */
    const void* w_out;  const float* b_out; /* attn.out_proj [E,E] */
    const void* w_fc;   const float* b_fc;  /* mlp.c_fc [4E,E] */
    const void* w_proj; const float* b_proj;/* mlp.c_proj [E,4E] */
    const float *ln1_w, *ln1_b, *ln2_w, *ln2_b;
} wc_block_weights;

typedef struct wc_vit_weights {             /* VisionTransformer, clip/model.py:246-262 */
    const void* conv1_w;                    /* h16 [768, 3*16*16] = conv1.weight flattened */
    const float* class_emb;                 /* [768] */
    const float* pos_emb;                   /* [pos_grid^2+1, 768] */
    int pos_grid;                           /* 14 */
    const float *ln_pre_w, *ln_pre_b;
    wc_block_weights blocks[12];
    const float *ln_post_w, *ln_post_b;
    const float* proj;                      /* fp32 [768,512] */
    /* block 12, transposed copies for the Grad-CAM input-gradient GEMMs (dX = dY @ W): */
    const void* w_proj_t;                   /* h16 [3072,768] = c_proj.weight^T */
    const void* w_fc_t;                     /* h16 [768,3072] = c_fc.weight^T */
    const void* w_out_t;                    /* h16 [768,768]  = out_proj.weight^T */
    const float* w_in_f32;                  /* fp32 [2304,768] = in_proj_weight */
    float logit_scale_exp;                  /* exp(logit_scale), clip/model.py:422 */
} wc_vit_weights;

typedef struct wc_decoder_weights {         /* SegFormerHead (segformer_head.py:49-80) + DecoderTransformer */
    const void* fuse_w1; const float* fuse_b1;   /* h16 [11,256,768], fp32 [11,256]  (linears_modulelist.i.proj) */
    const void* fuse_w2; const float* fuse_b2;   /* h16 [11,256,256], fp32 [11,256]  (.proj_2) */
    const void* fuse_wf; const float* fuse_bf;   /* h16 [256,2816],   fp32 [256]     (linear_fuse, 1x1 conv) */
    wc_block_weights blocks[3];                  /* width 256, 8 heads (TransDecoder.py:104-110) */
    const void* pred_w;  const float* pred_b;    /* h16 [C,256], fp32 [C]            (linear_pred, 1x1 conv) */
    int num_classes;
} wc_decoder_weights;

/* ------------------------------------------------------------------------------------------------
 * Encoder: patch embed, class token, resized positional embedding, ln_pre, 11 of the 12 blocks
 * (clip/model.py:264-287 + :225-243 via clip_tool.py:31-41 / CLIP.encode_image :387-390).
 *   img       fp32 [N,3,H,W]
 *   taps      h16 [11, N*L, 768]   block outputs F_0..F_10 (token-major, image-major rows; L = h*w+1)
 *   x_last    fp32 [N*L, 768]       F_10 in fp32 (input of block 12 / Grad-CAM)
 *   attn_avg  fp32 [n_attn(+slack), attn_imgs, L, L]  head-averaged attention (myAtt.py:323-326) of the LAST n_attn
 *             encoder blocks, oldest first, for the first attn_imgs images of the batch (0 = all N): the CAM path only
 *             runs for the images that have a name (model_attn_aff_voc.py:139); may be NULL when n_attn == 0
 *   attn_sums fp64 [n_attn(+slack), attn_imgs] or NULL: sum of attn_avg[l][n][1:, 1:], accumulated by the kernel that writes
 *             the map (the RFM layer score of clip_tool.py:158-160; saves wc_affinity_refine one pass over every map)
 * ---------------------------------------------------------------------------------------------- */
size_t wc_encoder_workspace_bytes(int N, int H, int W);
int wc_encoder_fwd(const wc_vit_weights* w, const float* img, int N, int H, int W, void* taps, float* x_last,
                   float* attn_avg, double* attn_sums, int n_attn, int attn_imgs, void* workspace, size_t workspace_bytes,
                   void* stream);

/* Decoder: 11 per-layer MLPs + concat + linear_fuse (segformer_head.py:69-80), attn_pred = sigmoid(F^T F)
 * (model_attn_aff_voc.py:134-137), 3 transformer blocks + linear_pred (TransDecoder.py:113-125).
 *   seg fp32 [N,C,h,w]; attn_pred fp32 [N,P,P]; fts (optional, may be NULL) fp32 [N*P,256] token-major */
size_t wc_decoder_workspace_bytes(int N, int h, int w, int num_classes);
int wc_decoder_fwd(const wc_decoder_weights* w, const void* taps, int N, int h, int wd, float* seg, float* attn_pred,
                   float* fts, void* workspace, size_t workspace_bytes, void* stream);

/* Stand-alone pieces exported for the numerics tests (same kernels the two calls above launch). */
int wc_layernorm(const float* x, const float* w, const float* b, void* out_f16, float* out_f32, long rows, int E,
                 void* stream);
int wc_softmax_headmean(const float* S, void* P_f16, float* avg, int n_img, int heads, int L, int Lp, void* stream);
/* Fused attention of the encoder (head_dim 64) and decoder (head_dim 32) blocks: softmax(q k^T) v without materialising
 * the scores (clip/myAtt.py:53-64), output head-interleaved (myAtt.py:320), plus the head-mean probability map
 * (myAtt.py:323-326, head_dim 64 only).
 *   q, k fp16 [N*heads, L, hd] (q pre-scaled by 1/sqrt(hd)); vt: head_dim 64 -> V itself, fp16 [N*heads, L, 64];
 *   head_dim 32 -> V transposed, fp16 [N*heads, 32, Lp] (Lp = round_up(L,8));
 *   o fp16 [N*L, heads*hd]; lse2 fp32 [N*heads, L] or NULL (log2-domain row log-sum-exp);
 *   attn_avg fp32 [N, L, L] or NULL (needs lse2 and head_dim 64). */
int wc_attention_fwd(const void* q, const void* k, const void* vt, void* o, float* lse2, float* attn_avg,
                     int n_img, int heads, int L, int Lp, int head_dim, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Grad-CAM head.  Replaces CLIP.forward_last_layer (clip/model.py:407-429) + GradCAM.__call__
 * (pytorch_grad_cam/base_cam.py:62-108,116-154, grad_cam.py:16-23, activations_and_gradients.py:19-37) as
 * looped per class by perform_single_voc_cam / perform_single_coco_cam (clip/clip_tool.py:139-147, :261-268).
 * Block 12 runs forward ONCE per image; the input-gradient to the ln_1 output is evaluated analytically and
 * batched over all (image, class) instances.
 *   x11        fp32 [N*L,768]    F_10 (x_last of wc_encoder_fwd)
 *   text       fp32 [sumT,512]   per image: its K_i foreground rows (fg_text_features[label_ids]) followed by
 *                                the background rows (clip_tool.py:134-137); rows need not be normalised
 *   text_off_host int [N+1]      row offsets into text;  k_off_host int [N+1]  instance offsets (K_i = diff)
 *   cams       fp32 [sumK,h*w]   raw Grad-CAMs, ReLU'd and min-max normalised twice
 *   attn_last  fp32 [N,L,L]      head-averaged attention of block 12
 *   attn_last_sum fp64 [N] or NULL  sum of attn_last[n][1:, 1:] (see wc_encoder_fwd)
 *   probs      fp32 [sumT] or NULL  softmax(logit_scale * cos) per image
 * ---------------------------------------------------------------------------------------------- */
size_t wc_gradcam_workspace_bytes(int N, int h, int w, int total_inst, int max_k);
int wc_gradcam(const wc_vit_weights* w, const float* x11, int N, int h, int wd, const float* text,
               const int* text_off_host, const int* k_off_host, float* cams, float* attn_last, double* attn_last_sum,
               float* probs, void* workspace, size_t workspace_bytes, void* stream);

/* The reference runs the out-projection of every attention block in fp16 (clip/myAtt.py:321), hence its autograd backward
 * too: with near-uniform class probabilities the gradient there (~1e-7) falls into fp16's subnormal range and the
 * reference keeps 2-3 significant bits of it.  wc_gradcam reproduces those two roundings by default (exactly: the chain is
 * scaled by a power of two); wc_gradcam_set_exact(1) switches them off (exact analytic gradient).  Returns the old mode.
 * Process-wide; set it before launching work. */
int wc_gradcam_set_exact(int exact);

/* Test hook (no reference counterpart): byte offsets, from a 256-byte aligned workspace pointer, of wc_gradcam's
 * intermediate buffers (activations, gradients, column sums) for the last backward chunk; offs has 23 entries in the
 * order documented in csrc/gradcam.cu.  Returns the backward chunk size (instances), negative on error. */
int wc_gradcam_debug_offsets(int N, int h, int w, int total_inst, int max_k, size_t* offs);

/* Attention-affinity refinement.  Replaces clip/clip_tool.py:152-191 (VOC) / :274-314 (COCO): RFM layer
 * selection against attn_pred (use_rfm=1; nl = 6 VOC / 10 COCO) or plain mean of the nl maps (use_rfm=0,
 * nl = 8), compute_trans_mat (:64-80), scoremap2bbox box mask (clip/utils.py:115-142; box_thr 0.4 / 0.7) and
 * cam' = (T * mask) @ cam.
 *   attn fp32 [nl,N,L,L] (oldest first, block 12 last); attn_sums fp64 [nl,N] = sum of attn[l][n][1:,1:] from the producers
 *   (NULL: computed here with one more pass over the maps); attn_pred fp32 [N,P,P]; cams / cams_refined fp32 [sumK,P];
 *   box_masks fp32 [sumK,P] or NULL (exported for the parity tests).  box_thr is a double: the reference evaluates
 *   int(thr * max(uint8)) in float64 (clip/utils.py:118-121); 0.7f * 250 would give 174 instead of 175.
 *   rfm_sel fp32 [N,2,nl] or NULL (test hook): per image the 0/1 layer selection of clip_tool.py:161-165 and the margins
 *   d_l - mean(d) it was decided by.  The layer sums are accumulated in fp64 here; the reference's fp32 torch.sum over
 *   P^2 ~ 1e6 terms resolves ~0.1-0.25 at that magnitude, which is the size of the margins. */
size_t wc_affinity_workspace_bytes(int N, int h, int w, int total_inst, int nl);
int wc_affinity_refine(const float* attn, int nl, const double* attn_sums, const float* attn_pred, int N, int h, int wd, int use_rfm,
                       const float* cams, const int* k_off_host, double box_thr, float* cams_refined,
                       float* box_masks, float* rfm_sel, void* workspace, size_t workspace_bytes, void* stream);

/* generate_cam_label (clip/clip_tool.py:202-216: per-class min-max + cv2.resize INTER_LINEAR to Wc x Hc) and the
 * background channel (model_attn_aff_voc.py:162-163).  out fp32 [sumK+N, Hc, Wc]: image i owns channels
 * [k_off[i]+i, k_off[i+1]+i+1) = (bg, class 1..K_i) — the packed layout wc_par_fwd takes.  stats_ws: 2*sumK floats. */
int wc_cam_upsample_bg(const float* cams, const int* k_off_host, int N, int h, int wd, int Hc, int Wc, float bg_pow,
                       float* out, float* stats_ws, void* stream);

/* ------------------------------------------------------------------------------------------------
 * msc+flip fusion and scoring (the caller that defines the metric).  Replaces the ATen
 * calls of validate(), test_msc_flip_voc.py:60-96 (COCO twin test_msc_flip_coco.py:61-96), and
 * utils/evaluate.py:9-15 (_fast_hist).
 * ---------------------------------------------------------------------------------------------- */
/* F.interpolate(mode='bilinear', align_corners=False) on [planes,Hi,Wi] -> [planes,Ho,Wo]
 * (test_msc_flip_voc.py:63 with size=, :80 with scale_factor=).  scale_factor_{y,x} > 0 reproduces torch's
 * scale_factor= call (source step 1/scale_factor); 0 means the size= call (source step in/out). */
int wc_resize_bilinear(const float* in, float* out, int planes, int Hi, int Wi, int Ho, int Wo,
                       float scale_factor_y, float scale_factor_x, void* stream);
/* torch.cat([x, x.flip(-1)], dim=0) (test_msc_flip_voc.py:68,82): in [planes,H,W] -> out [2,planes,H,W]. */
int wc_flip_concat(const float* in, float* out, int planes, int H, int W, void* stream);
/* One scale of the msc fusion (test_msc_flip_voc.py:74,86-88,90) for B same-size dataset images at once:
 * seg2 [2,B,C,hs,ws] (plain inputs, then their flipped twins) ->
 * acc[B,C,h,w] = (first ? 0 : acc) + weight * (R(seg2[0]) + flip(R(seg2[1]))) / 2, R = bilinear resize to h x w. */
int wc_msc_accumulate(const float* seg2, float* acc, int B, int C, int hs, int ws, int h, int w, float weight,
                      int first, void* stream);
/* argmax_c of the bilinear resize of logits [B,C,h,w] to Ho x Wo (test_msc_flip_voc.py:93-96) -> pred int32 [B,Ho,Wo]
 * (may be NULL) and, when gt (int16 [B,Ho,Wo]) and hist (int64 [C,C]) are given, hist[gt*C+pred] += 1 for 0<=gt<C. */
int wc_upsample_argmax_hist(const float* logits, int B, int C, int h, int w, int Ho, int Wo, int32_t* pred,
                            const int16_t* gt, int64_t* hist, void* stream);
/* _fast_hist over flat label arrays (the CAM pseudo-label histogram, test_msc_flip_voc.py:107):
 * pred int64 [n], gt int16 [n], hist int64 [C,C] accumulated in place. */
int wc_confusion_hist(const int64_t* pred, const int16_t* gt, long n, int C, int64_t* hist, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Training-time label prep and losses on the hot path's outputs: the step right after
 * WeCLIP.forward in scripts/dist_clip_voc.py:246-258.  workspace: wc_labelprep_workspace_bytes(B,H,W) bytes
 * (64 are enough for wc_aff_loss / wc_seg_loss); each call zeroes the part it uses.
 * ---------------------------------------------------------------------------------------------- */
size_t wc_labelprep_workspace_bytes(int B, int H, int W);
/* get_mask_by_radius (scripts/dist_clip_voc.py:116-133): mask uint8 [h*w, h*w], 1 where |dy| <= radius and
 * |dx| <= radius between patches. */
int wc_radius_mask(int h, int w, int radius, uint8_t* mask, void* stream);
/* cams_to_affinity_label (utils/camutils.py:226-247): cam_label int64 [B,H,W] -> nearest down-sample to
 * (H/16, W/16) -> aff_label int64 [B,P,P]: 1 equal labels, 0 different, ignore_index where radius_mask == 0
 * (uint8 [P,P], may be NULL) or either patch label == ignore_index. */
int wc_affinity_label(const int64_t* cam_label, int B, int H, int W, const uint8_t* radius_mask, int ignore_index,
                      int64_t* aff_label, void* workspace, size_t workspace_bytes, void* stream);
/* get_aff_loss (utils/losses.py:11-22): loss[0] = 0.5 sum_{t==1}(1-x)/(n_pos+1) + 0.5 sum_{t==0} x/(n_neg+1),
 * counts = {n_pos+1, n_neg+1}; grad (fp32 [n], may be NULL) = d loss / d inputs.  128-bit accesses when n % 4 == 0
 * and the pointers are 16-byte aligned, scalar otherwise. */
int wc_aff_loss(const float* inputs, const int64_t* targets, long n, float* loss, int64_t* counts, float* grad,
                void* workspace, size_t workspace_bytes, void* stream);
/* The same loss straight from the pseudo-labels: targets are derived on the fly (cams_to_affinity_label with the
 * get_mask_by_radius(radius) mask; radius < 0 = no mask), so the [B,P,P] int64 label tensor is never written or
 * read.  attn_pred fp32 [B,P,P], P = (H/16)*(W/16) (128-bit accesses when P % 4 == 0). */
int wc_aff_loss_fused(const float* attn_pred, const int64_t* cam_label, int B, int H, int W, int radius,
                      int ignore_index, float* loss, int64_t* counts, float* grad, void* workspace,
                      size_t workspace_bytes, void* stream);
/* wc_aff_loss_fused without a workspace argument and without any allocation.  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same inputs, outputs and results as wc_aff_loss_fused; the scratch (accumulator + down-sampled patch
 * labels) is a file-scope static __device__ array of 256 KB + 256 B that every GPU loading the library holds until exit.
 * Limit: B * (H/16) * (W/16) <= wc_aff_loss_static_max_patches() = 65536 (e.g. 64 crops of 512 x 512); larger calls return
 * WC_ERR_WORKSPACE with a message.  The device the stream belongs to must be current.  Calls no cudaMalloc / cudaFree; the first
 * call on a device looks the array's address up (cudaGetSymbolAddress).  All calls share the array: each call clears the
 * accumulator and rewrites the labels it uses, and two calls that overlap on the GPU (different streams) race. */
long wc_aff_loss_static_max_patches(void);
int wc_aff_loss_fused_static(const float* attn_pred, const int64_t* cam_label, int B, int H, int W, int radius,
                             int ignore_index, float* loss, int64_t* counts, float* grad, void* stream);
/* get_seg_loss (utils/losses.py:24-33): 0.5 * (CE over label==0 pixels + CE over foreground pixels), each a mean
 * over its own pixels (nan when it has none, like F.cross_entropy).  pred fp32 [B,C,H,W], label int64 [B,H,W];
 * grad fp32 [B,C,H,W] (may be NULL) = d loss / d pred. */
int wc_seg_loss(const float* pred, const int64_t* label, int B, int C, int H, int W, int ignore_index, float* loss,
                float* grad, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Trainable decoder: the non-GEMM pieces of the backward of decoder_fts_fuse + decoder + attn_pred
 * (scripts/dist_clip_voc.py:246-267 trains them with torch autograd; segformer_head.py:49-80,
 * Decoder/TransDecoder.py:70-125, model_attn_aff_voc.py:134-137).  Every matrix product of the forward, the input
 * gradients and the weight gradients goes through wc_gemm_f16 / wc_gemm_f16_ex; host side: weclip_b200/train.py.
 * ---------------------------------------------------------------------------------------------- */
/* out[b][c][r] = alpha * in[b][r][c] as h16 (in: fp32 when in_f32, else h16) for r < R, c < C; element strides; the
 * pad columns R..ld_out-1 of the output rows are not written (allocate the output zeroed when ld_out > R). */
int wc_transpose_f16(const void* in, int in_f32, void* out, int batch, int R, int C, long ld_in, long ld_out,
                     long in_bs, long out_bs, float alpha, void* stream);
/* out h16 = alpha * in.  ref_scale > 0 additionally reproduces the reference's fp16 rounding of a gradient that is carried
 * here with a power-of-two scale (clip/myAtt.py:321, see wc_gradcam_set_exact): half(alpha*in / ref_scale) * ref_scale. */
int wc_cast_f16(const float* in, void* out, long n, float alpha, float ref_scale, void* stream);
/* LayerNorm backward (eps 1e-5): dx = LN'(dy) [+ add]; dgamma / dbeta [E] are ACCUMULATED (zero them first). E in {256,512,768}. */
int wc_layernorm_bwd(const float* x, const float* gamma, const float* dy, const float* add, float* dx,
                     float* dgamma, float* dbeta, long rows, int E, void* stream);
/* Softmax backward per row: dS = P * (dP - sum_j P_j dP_j); P h16, dP fp32, dS h16, row pitch Lp (pad columns zeroed). */
int wc_softmax_bwd(const void* P, const float* dP, void* dS, long rows, int L, int Lp, void* stream);
/* out[c] = alpha * sum_r x[r][c] (bias gradients); x fp32 or h16 [M, N] with row stride ld. */
int wc_colsum(const void* x, int x_f16, float* out, long M, int N, long ld, float alpha, void* stream);
/* out h16 = alpha * dy * f'(.): kind 1 ReLU (aux = activation output), kind 2 QuickGELU (aux = pre-activation). */
int wc_act_bwd(int kind, const float* dy, const void* aux, void* out, long n, float alpha, void* stream);
/* attn_pred = sigmoid(F F^T): G h16 [N,P,Pp] = alpha * (g + g^T), g = dA * s * (1 - s); then dF = G F. */
int wc_sigmoid_gram_bwd(const float* dA, const float* S, void* G, int N, int P, int Pp, float alpha, void* stream);
/* Dropout2d (segformer_head.py:79): y[n][p][c] = x[n][p][c] * mask[n][c], mask = 0 or 1/(1-p); y fp32 and/or h16. */
int wc_dropout2d(const float* x, const float* mask, float* y, void* y_f16, int N, int P, int C, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Input pipeline: baseline JPEG -> normalised CHW fp32 on the device, label map -> label set.
 * Replaces imageio.imread + normalize_img + HWC->CHW (datasets/voc.py:48-67,253-271, datasets/transforms.py:8-15) and
 * np.unique on the label PNG (clip/clip_tool.py:111-118).  The decoded pixels are bit-identical to PIL / libjpeg's
 * defaults (integer "islow" IDCT, fancy chroma up-sampling, fixed-point YCbCr->RGB): see csrc/jpeg.cu.
 * ---------------------------------------------------------------------------------------------- */
typedef struct wc_jpeg_info {
    int width, height, ncomp;            /* ncomp 1 (grey) or 3 (YCbCr) */
    int hs[3], vs[3], hmax, vmax;        /* sampling factors */
    int mcus_x, mcus_y;
    int blocks_w[3], blocks_h[3];        /* 8x8 blocks per component (padded to whole MCUs) */
    long coef_off[3], coef_count;        /* int16 coefficients: component c starts at coef_off[c]; block-major, natural order */
    int comp_qt[3], restart_interval;
    uint16_t qt[4][64];                  /* quantisation tables, natural order */
} wc_jpeg_info;
/* HOST: parse the markers of a JPEG byte stream (WC_ERR_UNSUPPORTED for progressive / arithmetic / 12-bit / CMYK). */
int wc_jpeg_parse(const uint8_t* data_host, size_t n, wc_jpeg_info* info_host);
/* HOST: sequential Huffman decode into coef_host (info.coef_count int16; pinned memory recommended). Thread-safe. */
int wc_jpeg_entropy_decode(const uint8_t* data_host, size_t n, int16_t* coef_host, long coef_cap);
/* DEVICE: de-quantise + inverse DCT + chroma up-sampling + colour conversion.  coef: DEVICE copy of the coefficients.
 * rgb_hwc (uint8 [H,W,3], optional) = exactly what PIL returns; chw (fp32 [3,H,W], optional) = (rgb - mean) / std per
 * channel evaluated in fp64 and rounded to fp32 like numpy (mean / std: 3 host doubles). */
size_t wc_jpeg_workspace_bytes(const wc_jpeg_info* info_host);
int wc_jpeg_reconstruct(const wc_jpeg_info* info_host, const int16_t* coef, uint8_t* rgb_hwc, float* chw,
                        const double* mean_host, const double* std_host, void* workspace, size_t workspace_bytes, void* stream);
/* uint8 HWC image already on the device -> normalised fp32 CHW (datasets/transforms.py:8-15 + the transpose of voc.py:236). */
int wc_normalize_u8(const uint8_t* hwc, float* chw, int H, int W, const double* mean_host, const double* std_host, void* stream);
/* present256[v] = 1 iff value v occurs in the uint8 label map (np.unique of clip/clip_tool.py:113 as a presence table). */
int wc_label_presence(const uint8_t* label, long n, int* present256, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WECLIP_B200_H_ */
