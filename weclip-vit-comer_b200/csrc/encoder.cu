// Encoder (CLIP ViT-B/16, 11 of 12 blocks) and decoder (SegFormer-style fuse + 3 blocks) orchestration.
// All GEMM-shaped work goes through the tcgen05 kernel (gemm.cu); everything here is launch sequencing.
#include "encoder.cuh"

#include <math.h>

namespace wc {

static GemmOperand op(const h16* p, long rows, long ld, long bs0 = 0, long bs1 = 0) { return GemmOperand{p, rows, ld, bs0, bs1}; }

int run_block(const wc_block_weights& w, const BlockBuffers& b, int n_img, int L, int E, int heads, cudaStream_t st) {
    const int M = n_img * L, hd = E / heads;
    const long Lp = round8(L);
    int rc;
    // ln_1 (clip/model.py:211)
    if ((rc = launch_layernorm(b.x_in, w.ln1_w, w.ln1_b, b.a, b.a_f32, M, E, st, b.tap_in))) return rc;
    // packed in-proj + head split, q pre-scaled by 1/sqrt(hd) (myAtt.py:199-201,257-268,54)
    {
        GemmProblem p; p.M = M; p.N = 3 * E; p.K = E;
        p.A = op(b.a, M, E); p.B = op(reinterpret_cast<const h16*>(w.w_in), 3 * E, E);
        // head_dim 64 on the fused path: V stays head-major like q / k (the attention kernel reads it as an MN-major operand);
        // otherwise V^T for the K-major P V operand
        const bool v_direct = (!b.S && hd == 64);
        if (v_direct && !b.v) return fail(WC_ERR_ARG, "%s: head_dim 64 needs the v buffer", __func__);
        p.epi.mode = GM_QKV; p.epi.bias = w.b_in; p.epi.q = b.q; p.epi.k = b.k; p.epi.v = b.v; p.epi.vt = v_direct ? nullptr : b.vt;
        p.epi.L = L; p.epi.Lp = (int)Lp; p.epi.heads = heads; p.epi.hd = hd; p.epi.qscale = 1.f / sqrtf((float)hd);
        if ((rc = launch_gemm(p, st))) return rc;
    }
    if (!b.S) {
        // fused path: scores and probabilities never leave the SM (attention.cu); softmax(q k^T) v written
        // head-interleaved into [M,E] (myAtt.py:53-64,320), then the head-mean map from the row log-sum-exps (:323-326)
        if (b.attn_avg && (hd != 64 || !b.lse))
            return fail(WC_ERR_UNSUPPORTED, "%s: the head-mean map needs head_dim 64 and an lse buffer", __func__);
        if ((rc = launch_attention_fwd(b.q, b.k, hd == 64 ? b.v : b.vt, b.o, b.lse, n_img, heads, L, (int)Lp, hd, st))) return rc;
        if (b.attn_avg && (rc = launch_attn_headmean(b.q, b.k, b.lse, b.attn_avg, b.attn_imgs > 0 ? b.attn_imgs : n_img, heads, L, st, b.attn_sum))) return rc;
    } else {
        // S = q k^T (myAtt.py:56), fp32 [N*heads, L, Lp]
        {
            GemmProblem p; p.M = L; p.N = L; p.K = hd; p.nb0 = n_img * heads;
            p.A = op(b.q, L, hd, (long)L * hd); p.B = op(b.k, L, hd, (long)L * hd);
            p.epi.out = b.S; p.epi.ldo = Lp; p.epi.out_bs0 = (long)L * Lp;
            if ((rc = launch_gemm(p, st))) return rc;
        }
        // softmax (myAtt.py:59) + mean over heads (myAtt.py:323-326)
        if ((rc = launch_softmax_headmean(b.S, b.P, b.attn_avg, n_img, heads, L, (int)Lp, st))) return rc;
        // O = P V (myAtt.py:63) written head-interleaved into [M,E] (myAtt.py:320)
        {
            GemmProblem p; p.M = L; p.N = hd; p.K = L; p.nb0 = heads; p.nb1 = n_img;
            p.A = op(b.P, L, Lp, (long)L * Lp, (long)heads * L * Lp);
            p.B = op(b.vt, hd, Lp, (long)hd * Lp, (long)heads * hd * Lp);
            p.epi.out = b.o; p.epi.out_f16 = 1; p.epi.ldo = E; p.epi.out_bs0 = hd; p.epi.out_bs1 = (long)L * E;
            if ((rc = launch_gemm(p, st))) return rc;
        }
    }
    // out-proj + residual (myAtt.py:321, model.py:212)
    {
        GemmProblem p; p.M = M; p.N = E; p.K = E;
        p.A = op(b.o, M, E); p.B = op(reinterpret_cast<const h16*>(w.w_out), E, E);
        p.epi.bias = w.b_out; p.epi.out = b.x_mid; p.epi.ldo = E; p.epi.resid = b.x_in; p.epi.ldr = E;
        if ((rc = launch_gemm(p, st))) return rc;
    }
    // ln_2 + c_fc + QuickGELU + c_proj + residual (model.py:198-203,213)
    if ((rc = launch_layernorm(b.x_mid, w.ln2_w, w.ln2_b, b.a, nullptr, M, E, st))) return rc;
    {
        GemmProblem p; p.M = M; p.N = 4 * E; p.K = E;
        p.A = op(b.a, M, E); p.B = op(reinterpret_cast<const h16*>(w.w_fc), 4 * E, E);
        p.epi.bias = w.b_fc; p.epi.act = ACT_QUICKGELU; p.epi.out = b.u; p.epi.out_f16 = 1; p.epi.ldo = 4 * E;
        if (b.u_pre) { p.epi.out2 = b.u_pre; p.epi.ld2 = 4 * E; p.epi.out2_preact = 1; }
        if ((rc = launch_gemm(p, st))) return rc;
    }
    {
        GemmProblem p; p.M = M; p.N = E; p.K = 4 * E;
        p.A = op(b.u, M, 4 * E); p.B = op(reinterpret_cast<const h16*>(w.w_proj), E, 4 * E);
        p.epi.bias = w.b_proj; p.epi.out = b.x_out; p.epi.ldo = E; p.epi.resid = b.x_mid; p.epi.ldr = E;
        if (b.tap) { p.epi.out2 = b.tap; p.epi.ld2 = E; }
        if ((rc = launch_gemm(p, st))) return rc;
    }
    return WC_OK;
}

// Attention scratch shared by every block of one call.
struct AttnScratch { h16 *a, *q, *k, *v, *vt, *P, *o, *u; float *S, *lse; };

// fused: no materialised score / probability tensors, a [N*heads, L] log-sum-exp vector instead
static AttnScratch carve_block(Arena& ar, int n_img, int L, int E, int heads, bool fused) {
    const long M = (long)n_img * L, Lp = round8(L), hd = E / heads;
    AttnScratch s;
    s.a = ar.get<h16>(M * E);
    s.q = ar.get<h16>(M * E);
    s.k = ar.get<h16>(M * E);
    s.v = (fused && hd == 64) ? ar.get<h16>(M * E) : nullptr;
    s.vt = (fused && hd == 64) ? nullptr : ar.get<h16>((long)n_img * heads * hd * Lp);
    s.S = fused ? nullptr : ar.get<float>((long)n_img * heads * L * Lp);
    s.P = fused ? nullptr : ar.get<h16>((long)n_img * heads * L * Lp);
    s.lse = fused ? ar.get<float>((long)n_img * heads * L) : nullptr;
    s.o = ar.get<h16>(M * E);
    s.u = ar.get<h16>(M * 4 * E);
    return s;
}

static size_t encoder_layout(Arena& ar, int N, int H, int W, float** x, float** pos, h16** patches, AttnScratch* s) {
    const int h = H / 16, w = W / 16, P = h * w, L = P + 1, E = 768;
    *x = ar.get<float>((long)N * L * E);
    *pos = ar.get<float>((long)L * E);
    *patches = ar.get<h16>((long)N * P * E);
    *s = carve_block(ar, N, L, E, 12, true);
    return ar.off;
}

static size_t decoder_layout(Arena& ar, int N, int P, int C, h16** z1, h16** cat, float** x, h16** x_bf, AttnScratch* s) {
    const int D = 256;
    *z1 = ar.get<h16>((long)11 * N * P * D);
    *cat = ar.get<h16>((long)N * P * 11 * D);
    *x = ar.get<float>((long)N * P * D);
    *x_bf = ar.get<h16>((long)N * P * D);
    *s = carve_block(ar, N, P, D, 8, true);     // fused attention (head_dim 32); the three maps are unused downstream
    return ar.off;
}

}  // namespace wc

using namespace wc;

extern "C" size_t wc_encoder_workspace_bytes(int N, int H, int W) {
    Arena ar(nullptr, 0);
    float *x, *pos; h16* patches; AttnScratch s;
    return encoder_layout(ar, N, H, W, &x, &pos, &patches, &s) + 512;
}

extern "C" int wc_encoder_fwd(const wc_vit_weights* w, const float* img, int N, int H, int W, void* taps_, float* x_last,
                              float* attn_avg, double* attn_sums, int n_attn, int attn_imgs, void* workspace,
                              size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(w && img && taps_ && x_last && workspace && N > 0 && H >= 16 && W >= 16);
    WC_CHECK_ARG(n_attn >= 0 && n_attn <= 11 && (n_attn == 0 || attn_avg) && attn_imgs >= 0 && attn_imgs <= N);
    if (attn_imgs == 0) attn_imgs = N;
    cudaStream_t st = as_stream(stream);
    const int h = H / 16, wd = W / 16, P = h * wd, L = P + 1, E = 768;
    const long M = (long)N * L;
    Arena ar(workspace, workspace_bytes);
    float *x, *pos; h16* patches; AttnScratch s;
    encoder_layout(ar, N, H, W, &x, &pos, &patches, &s);
    if (!ar.ok) return fail(WC_ERR_WORKSPACE, "%s: workspace too small (%ld bytes)", __func__, (long)workspace_bytes);
    h16* taps = reinterpret_cast<h16*>(taps_);
    int rc;
    // clip/model.py:266 positional embedding resized every forward; :267 conv1 as a GEMM over gathered patches;
    // :270 class token; :271 + pos; :272 ln_pre
    if ((rc = launch_posemb_resize(w->pos_emb, pos, w->pos_grid, h, wd, E, st))) return rc;
    if ((rc = launch_im2col16(img, patches, N, H, W, st))) return rc;
    {
        GemmProblem p; p.M = P; p.N = E; p.K = E; p.nb0 = N;
        p.A = op(patches, P, E, (long)P * E); p.B = op(reinterpret_cast<const h16*>(w->conv1_w), E, E);
        p.epi.out = x + E; p.epi.ldo = E; p.epi.out_bs0 = (long)L * E;
        p.epi.resid = pos + E; p.epi.ldr = E; p.epi.resid_bs0 = 0;
        if ((rc = launch_gemm(p, st))) return rc;
    }
    if ((rc = launch_cls_row(x, w->class_emb, pos, N, L, E, st))) return rc;
    if ((rc = launch_layernorm(x, w->ln_pre_w, w->ln_pre_b, nullptr, x, M, E, st))) return rc;
    for (int i = 0; i < 11; ++i) {
        BlockBuffers b{};
        b.x_in = x; b.x_mid = x; b.x_out = (i == 10) ? x_last : x;
        b.a = s.a; b.q = s.q; b.k = s.k; b.v = s.v; b.vt = s.vt; b.S = s.S; b.P = s.P; b.lse = s.lse; b.o = s.o; b.u = s.u;
        const int slot = i - (11 - n_attn);
        b.attn_avg = (slot >= 0) ? attn_avg + (size_t)slot * attn_imgs * L * L : nullptr;
        b.attn_imgs = attn_imgs;
        b.attn_sum = (slot >= 0 && attn_sums) ? attn_sums + (size_t)slot * attn_imgs : nullptr;
        // tap i = fp16 copy of block i's output: blocks 0..9 add their MLP output into x in place (L2 reduction, no second
        // epilogue output), so the next block's ln_1 - which reads x anyway - writes the copy; block 10 writes its own
        b.tap_in = (i > 0) ? taps + (size_t)(i - 1) * M * E : nullptr;
        b.tap = (i == 10) ? taps + (size_t)i * M * E : nullptr;
        if ((rc = run_block(w->blocks[i], b, N, L, E, 12, st))) return rc;
    }
    return WC_OK;
}

extern "C" size_t wc_decoder_workspace_bytes(int N, int h, int w, int num_classes) {
    Arena ar(nullptr, 0);
    h16 *z1, *cat, *xb; float* x; AttnScratch s;
    return decoder_layout(ar, N, h * w, num_classes, &z1, &cat, &x, &xb, &s) + 512;
}

extern "C" int wc_decoder_fwd(const wc_decoder_weights* w, const void* taps_, int N, int h, int wd, float* seg,
                              float* attn_pred, float* fts, void* workspace, size_t workspace_bytes, void* stream) {
    WC_CHECK_ARG(w && taps_ && seg && attn_pred && workspace && N > 0 && h > 0 && wd > 0 && w->num_classes > 0);
    cudaStream_t st = as_stream(stream);
    const int P = h * wd, L = P + 1, E = 768, D = 256, C = w->num_classes;
    const long M = (long)N * L, MP = (long)N * P;
    Arena ar(workspace, workspace_bytes);
    h16 *z1, *cat, *xb; float* x; AttnScratch s;
    decoder_layout(ar, N, P, C, &z1, &cat, &x, &xb, &s);
    if (!ar.ok) return fail(WC_ERR_WORKSPACE, "%s: workspace too small (%ld bytes)", __func__, (long)workspace_bytes);
    const h16* taps = reinterpret_cast<const h16*>(taps_);
    int rc;
    // segformer_head.py:13-28,69-76: per-layer MLP on the patch tokens (cls row skipped via the +E pointer offset);
    // one grouped launch for all 11 layers x N images
    {
        GemmProblem p; p.M = P; p.N = D; p.K = E; p.nb0 = N; p.nb1 = 11;
        p.A = op(taps + E, P, E, (long)L * E, M * E);
        p.B = op(reinterpret_cast<const h16*>(w->fuse_w1), D, E, 0, (long)D * E);
        p.epi.bias = w->fuse_b1; p.epi.bias_bs1 = D; p.epi.act = ACT_RELU;
        p.epi.out = z1; p.epi.out_f16 = 1; p.epi.ldo = D; p.epi.out_bs0 = (long)P * D; p.epi.out_bs1 = MP * D;
        if ((rc = launch_gemm(p, st))) return rc;
    }
    {   // proj_2, written straight into the channel-concatenated layout (segformer_head.py:77: cat on dim=1)
        GemmProblem p; p.M = (int)MP; p.N = D; p.K = D; p.nb0 = 11;
        p.A = op(z1, MP, D, MP * D);
        p.B = op(reinterpret_cast<const h16*>(w->fuse_w2), D, D, (long)D * D);
        p.epi.bias = w->fuse_b2; p.epi.bias_bs0 = D;
        p.epi.out = cat; p.epi.out_f16 = 1; p.epi.ldo = 11 * D; p.epi.out_bs0 = D;
        if ((rc = launch_gemm(p, st))) return rc;
    }
    {   // linear_fuse 1x1 conv (segformer_head.py:78); Dropout2d is identity in eval
        GemmProblem p; p.M = (int)MP; p.N = D; p.K = 11 * D;
        p.A = op(cat, MP, 11 * D); p.B = op(reinterpret_cast<const h16*>(w->fuse_wf), D, 11 * D);
        p.epi.bias = w->fuse_bf; p.epi.out = x; p.epi.ldo = D; p.epi.out2 = xb; p.epi.ld2 = D;
        if ((rc = launch_gemm(p, st))) return rc;
    }
    if (fts) WC_CUDA(cudaMemcpyAsync(fts, x, sizeof(float) * MP * D, cudaMemcpyDeviceToDevice, st));
    {   // attn_pred = sigmoid(F^T F) (model_attn_aff_voc.py:134-137)
        GemmProblem p; p.M = P; p.N = P; p.K = D; p.nb0 = N;
        p.A = op(xb, P, D, (long)P * D); p.B = op(xb, P, D, (long)P * D);
        p.epi.act = ACT_SIGMOID; p.epi.out = attn_pred; p.epi.ldo = P; p.epi.out_bs0 = (long)P * P;
        if ((rc = launch_gemm(p, st))) return rc;
    }
    for (int i = 0; i < 3; ++i) {   // TransDecoder.py:113-119 (the three attention maps are unused downstream)
        BlockBuffers b{};
        b.x_in = x; b.x_mid = x; b.x_out = x;
        b.a = s.a; b.q = s.q; b.k = s.k; b.v = s.v; b.vt = s.vt; b.S = s.S; b.P = s.P; b.lse = s.lse; b.o = s.o; b.u = s.u;
        b.tap = (i == 2) ? xb : nullptr;
        if ((rc = run_block(w->blocks[i], b, N, P, D, 8, st))) return rc;
    }
    {   // linear_pred 1x1 conv -> [N,C,h,w] (TransDecoder.py:121-122): transposed store
        GemmProblem p; p.M = P; p.N = C; p.K = D; p.nb0 = N;
        p.A = op(xb, P, D, (long)P * D); p.B = op(reinterpret_cast<const h16*>(w->pred_w), C, D);
        p.epi.bias = w->pred_b; p.epi.out = seg; p.epi.out_trans = 1; p.epi.ldo = P; p.epi.out_bs0 = (long)C * P;
        if ((rc = launch_gemm(p, st))) return rc;
    }
    return WC_OK;
}

extern "C" int wc_layernorm(const float* x, const float* w, const float* b, void* out_f16, float* out_f32, long rows,
                            int E, void* stream) {
    return launch_layernorm(x, w, b, reinterpret_cast<h16*>(out_f16), out_f32, rows, E, as_stream(stream));
}

extern "C" int wc_softmax_headmean(const float* S, void* P, float* avg, int n_img, int heads, int L, int Lp, void* stream) {
    return launch_softmax_headmean(S, reinterpret_cast<h16*>(P), avg, n_img, heads, L, Lp, as_stream(stream));
}
