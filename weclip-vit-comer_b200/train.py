"""Trainable decoder behind WeCLIP.forward.

The reference trains ``decoder_fts_fuse`` (SegFormerHead, WeCLIP_model/segformer_head.py:49-80), ``decoder``
(DecoderTransformer, WeCLIP_model/Decoder/TransDecoder.py:70-125) and the ``attn_pred = sigmoid(F^T F)`` branch
(model_attn_aff_voc.py:134-137) with torch autograd (scripts/dist_clip_voc.py:246-267: loss on ``segs`` / ``attn_pred``,
``backward``, ``optimizer.step``).  Here:

* ``SegFormerHead`` / ``DecoderTransformer`` are parameter containers with the reference's module tree, so
  ``state_dict()`` keys, ``get_param_groups()`` and optimisers see live ``nn.Parameter``s;
* ``DecoderFunction`` is ONE ``torch.autograd.Function``: its forward and backward are sequences of launches into
  libweclip_sm100.so — every matrix product (forward, input gradients, weight gradients, the four attention products and
  their backward) on the tcgen05 GEMM (``wc_gemm_f16_ex``, fp16 operands, fp32 accumulation), the rest on the kernels of
  csrc/train.cu.  Gradients are carried with a power-of-two loss scale (fp16 operands) and un-scaled in the weight-gradient
  epilogues; the out-projection's backward reproduces the reference's fp16 rounding (clip/myAtt.py:321).

PyTorch supplies tensors, the autograd tape between this function and the losses, and the optimiser; no torch arithmetic
runs on the decoder path.
"""
from __future__ import annotations

import ctypes
import math
from collections import OrderedDict
from typing import List

import torch
import torch.nn as nn

from . import _lib

D, HEADS, HD, N_TAPS, E_VIT, N_BLOCKS = 256, 8, 32, 11, 768, 3
ACT_NONE, ACT_RELU, ACT_QGELU, ACT_SIGMOID = 0, 1, 2, 3


# ---------------------------------------------------------------------------------------------------------------------
# Parameter containers with the reference's module tree (state_dict keys: SURVEY.md §8b)
# ---------------------------------------------------------------------------------------------------------------------
class _MLP(nn.Module):                      # segformer_head.py:13-28
    def __init__(self, input_dim, embed_dim):
        super().__init__()
        self.proj = nn.Linear(input_dim, embed_dim)
        self.proj_2 = nn.Linear(embed_dim, embed_dim)


class SegFormerHead(nn.Module):             # segformer_head.py:49-80
    def __init__(self, in_channels=E_VIT, embedding_dim=D, index=N_TAPS):
        super().__init__()
        self.linears_modulelist = nn.ModuleList([_MLP(in_channels, embedding_dim) for _ in range(index)])
        self.linear_fuse = nn.Conv2d(embedding_dim * index, embedding_dim, kernel_size=1)
        self.dropout = nn.Dropout2d(0.1)


class _Attention(nn.Module):                # clip/myAtt.py MultiheadAttention parameters
    def __init__(self, width):
        super().__init__()
        self.in_proj_weight = nn.Parameter(torch.empty(3 * width, width))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * width))
        self.out_proj = nn.Linear(width, width)
        nn.init.xavier_uniform_(self.in_proj_weight)


class _Block(nn.Module):                    # TransDecoder.py:70-93
    def __init__(self, width):
        super().__init__()
        self.attn = _Attention(width)
        self.ln_1 = nn.LayerNorm(width)
        self.mlp = nn.Sequential(OrderedDict([("c_fc", nn.Linear(width, width * 4)), ("gelu", nn.Identity()),
                                              ("c_proj", nn.Linear(width * 4, width))]))
        self.ln_2 = nn.LayerNorm(width)


class _Transformer(nn.Module):
    def __init__(self, width, layers):
        super().__init__()
        self.resblocks = nn.Sequential(*[_Block(width) for _ in range(layers)])


class DecoderTransformer(nn.Module):        # TransDecoder.py:104-125
    def __init__(self, width=D, layers=N_BLOCKS, output_dim=21):
        super().__init__()
        self.transformer = _Transformer(width, layers)
        self.linear_pred = nn.Conv2d(width, output_dim, kernel_size=1)


def decoder_parameters(fuse: SegFormerHead, dec: DecoderTransformer) -> List[nn.Parameter]:
    """Flat, fixed order used by DecoderFunction (and only there)."""
    ps = []
    for m in fuse.linears_modulelist:
        ps += [m.proj.weight, m.proj.bias, m.proj_2.weight, m.proj_2.bias]
    ps += [fuse.linear_fuse.weight, fuse.linear_fuse.bias]
    for b in dec.transformer.resblocks:
        ps += [b.attn.in_proj_weight, b.attn.in_proj_bias, b.attn.out_proj.weight, b.attn.out_proj.bias,
               b.ln_1.weight, b.ln_1.bias, b.mlp.c_fc.weight, b.mlp.c_fc.bias, b.mlp.c_proj.weight, b.mlp.c_proj.bias,
               b.ln_2.weight, b.ln_2.bias]
    ps += [dec.linear_pred.weight, dec.linear_pred.bias]
    return ps


# ---------------------------------------------------------------------------------------------------------------------
# Launch helpers (ctypes -> libweclip_sm100.so); tensors are views with the strides the C ABI takes
# ---------------------------------------------------------------------------------------------------------------------
def _r8(v: int) -> int:
    return (v + 7) // 8 * 8


def _c(v):
    return ctypes.c_long(int(v))


def gemm(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, K: int = None, bias=None, resid=None, act=ACT_NONE,
         alpha: float = 1.0, trans: bool = False, bias_bs=(0, 0)):
    """out[b1,b0] = act(alpha * A[b1,b0] @ B[b1,b0]^T + bias) (+ resid).  A [nb1,nb0,M,K], B [nb1|1,nb0|1,N,K] (size-1
    batch dims broadcast), out [nb1,nb0,M,N] ([.., N, M] when trans); all may be strided views with unit inner stride."""
    assert A.dtype == torch.float16 and B.dtype == torch.float16 and A.stride(-1) == 1 and B.stride(-1) == 1
    while A.dim() < 4:
        A = A.unsqueeze(0)
    while B.dim() < 4:
        B = B.unsqueeze(0)
    while out.dim() < 4:
        out = out.unsqueeze(0)
    nb1, nb0, M, Ka = A.shape
    K = Ka if K is None else K
    N = B.shape[2]
    bs = lambda t, d: 0 if t.shape[d] == 1 else t.stride(d)
    assert out.stride(-1) == 1 and out.shape[0] == nb1 and out.shape[1] == nb0
    rc = _lib.lib().wc_gemm_f16_ex(
        _lib.ptr(A), _lib.ptr(B), _lib.ptr(out), _lib.ptr(bias), _lib.ptr(resid), M, N, K, nb0, nb1,
        _c(A.stride(2)), _c(B.stride(2)), _c(out.stride(2)), _c(bs(A, 1)), _c(bs(A, 0)), _c(bs(B, 1)), _c(bs(B, 0)),
        _c(bs(out, 1)), _c(bs(out, 0)), _c(bias_bs[0]), _c(bias_bs[1]), int(out.dtype == torch.float16), int(trans),
        int(act), ctypes.c_float(alpha), _lib.stream_ptr())
    _lib.check(rc, "wc_gemm_f16_ex")
    return out


def transpose16(x: torch.Tensor, R: int, C: int, ld_out: int = None, alpha: float = 1.0) -> torch.Tensor:
    """x [batch, R, C] (any row stride, fp32 or fp16) -> fp16 [batch, C, ld_out] with out[b][c][r] = alpha * x[b][r][c]."""
    if x.dim() == 2:
        x = x.unsqueeze(0)
    batch = x.shape[0]
    ld_out = _r8(R) if ld_out is None else ld_out
    out = (torch.zeros if ld_out != R else torch.empty)(batch, C, ld_out, dtype=torch.float16, device=x.device)
    rc = _lib.lib().wc_transpose_f16(_lib.ptr(x), int(x.dtype == torch.float32), _lib.ptr(out), batch, R, C, _c(x.stride(1)),
                                     _c(ld_out), _c(x.stride(0) if batch > 1 else 0), _c(C * ld_out), ctypes.c_float(alpha),
                                     _lib.stream_ptr())
    _lib.check(rc, "wc_transpose_f16")
    return out


def cast16(x: torch.Tensor, alpha: float = 1.0, ref_scale: float = 0.0) -> torch.Tensor:
    out = torch.empty(x.shape, dtype=torch.float16, device=x.device)
    rc = _lib.lib().wc_cast_f16(_lib.ptr(x.contiguous()), _lib.ptr(out), _c(x.numel()), ctypes.c_float(alpha),
                                ctypes.c_float(ref_scale), _lib.stream_ptr())
    _lib.check(rc, "wc_cast_f16")
    return out


def layernorm(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    out = torch.empty(x.shape, dtype=torch.float16, device=x.device)
    rc = _lib.lib().wc_layernorm(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), None, _c(x.shape[0]), x.shape[1],
                                 _lib.stream_ptr())
    _lib.check(rc, "wc_layernorm")
    return out


def layernorm_bwd(x, gamma, dy, add, dgamma, dbeta) -> torch.Tensor:
    dx = torch.empty_like(x)
    rc = _lib.lib().wc_layernorm_bwd(_lib.ptr(x), _lib.ptr(gamma), _lib.ptr(dy), _lib.ptr(add), _lib.ptr(dx),
                                     _lib.ptr(dgamma), _lib.ptr(dbeta), _c(x.shape[0]), x.shape[1], _lib.stream_ptr())
    _lib.check(rc, "wc_layernorm_bwd")
    return dx


def colsum(x: torch.Tensor, n_cols: int, alpha: float, out: torch.Tensor = None) -> torch.Tensor:
    """x [M, >= n_cols] fp32 / fp16 with row stride x.stride(0) -> out[c] = alpha * sum_r x[r][c]."""
    out = torch.empty(n_cols, dtype=torch.float32, device=x.device) if out is None else out
    rc = _lib.lib().wc_colsum(_lib.ptr(x), int(x.dtype == torch.float16), _lib.ptr(out), _c(x.shape[0]), n_cols,
                              _c(x.stride(0)), ctypes.c_float(alpha), _lib.stream_ptr())
    _lib.check(rc, "wc_colsum")
    return out


def act_bwd(kind: int, dy: torch.Tensor, aux: torch.Tensor) -> torch.Tensor:
    out = torch.empty(dy.shape, dtype=torch.float16, device=dy.device)
    rc = _lib.lib().wc_act_bwd(kind, _lib.ptr(dy), _lib.ptr(aux), _lib.ptr(out), _c(dy.numel()), ctypes.c_float(1.0),
                               _lib.stream_ptr())
    _lib.check(rc, "wc_act_bwd")
    return out


def dropout2d(x: torch.Tensor, mask: torch.Tensor, N: int, P: int, want_f32=True, want_f16=True):
    y = torch.empty_like(x) if want_f32 else None
    y16 = torch.empty(x.shape, dtype=torch.float16, device=x.device) if want_f16 else None
    rc = _lib.lib().wc_dropout2d(_lib.ptr(x), _lib.ptr(mask), _lib.ptr(y), _lib.ptr(y16), N, P, x.shape[-1], _lib.stream_ptr())
    _lib.check(rc, "wc_dropout2d")
    return y, y16


def _heads(t: torch.Tensor, N: int, P: int) -> torch.Tensor:
    """[N*P, 256] (view, any row stride) -> [N, heads, P, 32] view."""
    return t.as_strided((N, HEADS, P, HD), (P * t.stride(0), HD, t.stride(0), 1), t.storage_offset())


def _headsT(t: torch.Tensor, N: int, Pp: int) -> torch.Tensor:
    """[N, 256, Pp] -> [N, heads, 32, Pp] view."""
    return t.view(N, HEADS, HD, Pp)


# ---------------------------------------------------------------------------------------------------------------------
class DecoderFunction(torch.autograd.Function):
    """(taps, dims, dropout mask, *decoder parameters) -> (seg [N,C,h,w], attn_pred [N,P,P]) with a hand-written backward."""

    @staticmethod
    def forward(ctx, taps, dims, mask, *params):
        N, h, w = dims
        P, L, dev = h * w, h * w + 1, taps.device
        M, Pp, Mp = N * P, _r8(h * w), _r8(N * h * w)
        f16 = lambda t: cast16(t.detach().float().contiguous())
        it = iter(params)
        fuse = [(next(it), next(it), next(it), next(it)) for _ in range(N_TAPS)]
        wf, bf = next(it), next(it)
        blocks = [tuple(next(it) for _ in range(12)) for _ in range(N_BLOCKS)]
        wp, bp = next(it), next(it)
        C = wp.shape[0]
        sv = ctx.sv = {}
        # ---- decoder_fts_fuse: 11 x (768 -> 256 -> ReLU -> 256), concat, 1x1 conv (segformer_head.py:69-80)
        W1 = f16(torch.stack([f[0] for f in fuse])); b1 = torch.stack([f[1] for f in fuse]).detach().float().contiguous()
        W2 = f16(torch.stack([f[2] for f in fuse])); b2 = torch.stack([f[3] for f in fuse]).detach().float().contiguous()
        Wf = f16(wf.reshape(D, D * N_TAPS))
        tok = taps.view(N_TAPS, N, L, E_VIT)[:, :, 1:, :]                       # patch tokens [11,N,P,768] (strided view)
        H1 = torch.empty(N_TAPS, N, P, D, dtype=torch.float16, device=dev)
        gemm(tok, W1.view(N_TAPS, 1, D, E_VIT), H1, bias=b1, act=ACT_RELU, bias_bs=(0, D))
        Zcat = torch.empty(M, D * N_TAPS, dtype=torch.float16, device=dev)
        gemm(H1.view(N_TAPS, 1, M, D), W2.view(N_TAPS, 1, D, D), Zcat.view(M, N_TAPS, D).permute(1, 0, 2).unsqueeze(1),
             bias=b2, bias_bs=(0, D))
        F0 = torch.empty(M, D, dtype=torch.float32, device=dev)
        gemm(Zcat, Wf, F0, bias=bf.detach().float().contiguous())
        Fm, F16 = dropout2d(F0, mask, N, P)                                       # Dropout2d(0.1) when training (:79)
        # ---- attn_pred = sigmoid(F^T F) (model_attn_aff_voc.py:134-137)
        attn_pred = torch.empty(N, P, P, dtype=torch.float32, device=dev)
        gemm(F16.view(1, N, P, D), F16.view(1, N, P, D), attn_pred.view(1, N, P, P), act=ACT_SIGMOID)
        # ---- three transformer blocks (TransDecoder.py:86-93 with clip/myAtt.py inside)
        X = Fm
        sv["blocks"] = []
        for (win, bin_, wo, bo, g1, be1, wfc, bfc, wpr, bpr, g2, be2) in blocks:
            s = {}
            s["x_in"] = X
            a1 = layernorm(X, g1.detach().float().contiguous(), be1.detach().float().contiguous())
            qkv = torch.empty(M, 3 * D, dtype=torch.float16, device=dev)
            gemm(a1, f16(win), qkv, bias=bin_.detach().float().contiguous())
            q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
            S = torch.empty(N, HEADS, P, Pp, dtype=torch.float32, device=dev)
            gemm(_heads(q, N, P), _heads(k, N, P), S[..., :P], alpha=1.0 / math.sqrt(HD))
            Pm = torch.zeros(N * HEADS, P, Pp, dtype=torch.float16, device=dev)
            rc = _lib.lib().wc_softmax_headmean(_lib.ptr(S), _lib.ptr(Pm), None, N, HEADS, P, Pp, _lib.stream_ptr())
            _lib.check(rc, "wc_softmax_headmean")
            Vt = transpose16(v.as_strided((N, P, D), (P * 3 * D, 3 * D, 1), v.storage_offset()), P, D, Pp)      # [N,256,Pp]
            O = torch.empty(M, D, dtype=torch.float16, device=dev)
            gemm(Pm.view(N, HEADS, P, Pp), _headsT(Vt, N, Pp), _heads(O, N, P), K=P)
            x_mid = torch.empty(M, D, dtype=torch.float32, device=dev)
            gemm(O, f16(wo), x_mid, bias=bo.detach().float().contiguous(), resid=X)          # fp16 out-proj (myAtt.py:321)
            a2 = layernorm(x_mid, g2.detach().float().contiguous(), be2.detach().float().contiguous())
            Wfc16 = f16(wfc)
            upre = torch.empty(M, 4 * D, dtype=torch.float16, device=dev)
            gemm(a2, Wfc16, upre, bias=bfc.detach().float().contiguous())
            u = torch.empty(M, 4 * D, dtype=torch.float16, device=dev)
            gemm(a2, Wfc16, u, bias=bfc.detach().float().contiguous(), act=ACT_QGELU)
            x_out = torch.empty(M, D, dtype=torch.float32, device=dev)
            gemm(u, f16(wpr), x_out, bias=bpr.detach().float().contiguous(), resid=x_mid)
            s.update(a1=a1, qkv=qkv, Pm=Pm, O=O, x_mid=x_mid, a2=a2, upre=upre, u=u)
            sv["blocks"].append(s)
            X = x_out
        # ---- linear_pred, 1x1 conv (TransDecoder.py:122)
        X16 = cast16(X)
        seg = torch.empty(N, C, h, w, dtype=torch.float32, device=dev)
        gemm(X16.view(1, N, P, D), f16(wp.reshape(C, D)).view(1, 1, C, D), seg.view(1, N, C, P), bias=bp.detach().float().contiguous(),
             trans=True)
        sv.update(dims=(N, h, w, C), tok=tok, H1=H1, Zcat=Zcat, F16=F16, attn_pred=attn_pred, X16=X16, mask=mask)
        ctx.params = params
        return seg, attn_pred

    @staticmethod
    def backward(ctx, dseg, dattn):
        sv, params = ctx.sv, ctx.params
        N, h, w, C = sv["dims"]
        P = h * w
        M, Pp, Mp, Cp = N * P, _r8(P), _r8(N * P), _r8(C)
        dev = sv["F16"].device
        f16 = lambda t: cast16(t.detach().float().contiguous())
        tr_w = lambda wt: transpose16(wt.detach().float().reshape(wt.shape[0], -1), wt.shape[0], wt[0].numel())[0]  # [out,in] -> [in,r8(out)]
        it = iter(params)
        fuse = [(next(it), next(it), next(it), next(it)) for _ in range(N_TAPS)]
        wf, bf = next(it), next(it)
        blocks = [tuple(next(it) for _ in range(12)) for _ in range(N_BLOCKS)]
        wp, bp = next(it), next(it)
        # power-of-two loss scale: the gradient operands of the tensor-core products are fp16.  Eagerly it follows this step's
        # incoming gradients (one host read); a captured step (TrainStep) runs with the scale of the eager step it was captured
        # after — |d loss / d seg| ~ 1 / pixels moves little between batches and max -> 2^8 leaves 2^7 of headroom either way —
        # and leaves the maximum on the device so that the host can look at it now and then.
        am = None
        for t in (dseg, dattn):
            if t is not None:
                a = t.detach().abs().max()
                am = a if am is None else torch.maximum(am, a)
        DecoderFunction.last_amax = am
        if DecoderFunction.scale_override is not None:
            sc = DecoderFunction.scale_override
        else:
            sc = loss_scale_for(float(am) if am is not None else 0.0)
            DecoderFunction.last_scale = sc
        inv = 1.0 / sc
        grads = {}
        zeros = lambda *shape: torch.zeros(*shape, dtype=torch.float32, device=dev)

        def wgrad(dY16, X16, n_out, n_in):
            """dW [n_out, n_in] = dY^T X / sc (contraction over the M tokens on the tensor cores)."""
            dYT = transpose16(dY16, M, n_out, Mp)[0]
            XT = transpose16(X16, M, n_in, Mp)[0]
            out = torch.empty(n_out, n_in, dtype=torch.float32, device=dev)
            return gemm(dYT, XT, out, alpha=inv)

        # ---- linear_pred
        if dseg is None:
            dX = zeros(M, D)
            g_wp, g_bp = zeros(C, D), zeros(C)
        else:
            dY = transpose16(dseg.contiguous().view(N, C, P), C, P, Cp, alpha=sc).view(M, Cp)       # [M, Cp] token-major fp16
            dX = torch.empty(M, D, dtype=torch.float32, device=dev)
            gemm(dY, tr_w(wp), dX, K=Cp)                                                             # dX = dY Wp
            g_wp = wgrad(dY[:, :C], sv["X16"], C, D)
            g_bp = colsum(dY, C, inv)
        # ---- transformer blocks, last to first
        g_blocks = []
        for s, (win, bin_, wo, bo, g1, be1, wfc, bfc, wpr, bpr, g2, be2) in zip(reversed(sv["blocks"]), reversed(blocks)):
            d16 = cast16(dX)
            dU = torch.empty(M, 4 * D, dtype=torch.float32, device=dev)
            gemm(d16, tr_w(wpr), dU)                                                                 # [M,1024] = dX Wproj
            g_wpr, g_bpr = wgrad(d16, s["u"], D, 4 * D), colsum(dX, D, inv)
            dUp = act_bwd(ACT_QGELU, dU, s["upre"])
            dA2 = torch.empty(M, D, dtype=torch.float32, device=dev)
            gemm(dUp, tr_w(wfc), dA2)
            g_wfc, g_bfc = wgrad(dUp, s["a2"], 4 * D, D), colsum(dUp, 4 * D, inv)
            g_g2, g_be2 = zeros(D), zeros(D)
            dXm = layernorm_bwd(s["x_mid"], g2.detach().float().contiguous(), dA2, dX, g_g2, g_be2)
            # out-projection backward in the reference's fp16 (incoming gradient and result rounded to half, myAtt.py:321)
            dm16 = cast16(dXm, ref_scale=sc)
            dO32 = torch.empty(M, D, dtype=torch.float32, device=dev)
            gemm(dm16, tr_w(wo), dO32)
            dO = cast16(dO32, ref_scale=sc)
            g_wo, g_bo = wgrad(dm16, s["O"], D, D), colsum(dXm, D, inv)
            # attention backward (four products per head, batched over heads x images)
            qkv, Pm = s["qkv"], s["Pm"]
            q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
            img = lambda t: t.as_strided((N, P, D), (P * t.stride(0), t.stride(0), 1), t.storage_offset())
            dqkv = torch.empty(M, 3 * D, dtype=torch.float16, device=dev)
            dOT = transpose16(img(dO), P, D, Pp)
            PT = transpose16(Pm, P, P, Pp)
            gemm(PT.view(N, HEADS, P, Pp), _headsT(dOT, N, Pp), _heads(dqkv[:, 2 * D:], N, P), K=P)            # dV = P^T dO
            dP = torch.empty(N * HEADS, P, Pp, dtype=torch.float32, device=dev)
            gemm(_heads(dO, N, P), _heads(v, N, P), dP.view(N, HEADS, P, Pp)[..., :P])
            dS = torch.empty(N * HEADS, P, Pp, dtype=torch.float16, device=dev)
            rc = _lib.lib().wc_softmax_bwd(_lib.ptr(Pm), _lib.ptr(dP), _lib.ptr(dS), _c(N * HEADS * P), P, Pp, _lib.stream_ptr())
            _lib.check(rc, "wc_softmax_bwd")
            KT, QT = transpose16(img(k), P, D, Pp), transpose16(img(q), P, D, Pp)
            rs = 1.0 / math.sqrt(HD)
            gemm(dS.view(N, HEADS, P, Pp), _headsT(KT, N, Pp), _heads(dqkv[:, :D], N, P), K=P, alpha=rs)       # dQ = dS K / sqrt(d)
            dST = transpose16(dS, P, P, Pp)
            gemm(dST.view(N, HEADS, P, Pp), _headsT(QT, N, Pp), _heads(dqkv[:, D:2 * D], N, P), K=P, alpha=rs)  # dK = dS^T Q / sqrt(d)
            dA1 = torch.empty(M, D, dtype=torch.float32, device=dev)
            gemm(dqkv, tr_w(win), dA1)
            g_win, g_bin = wgrad(dqkv, s["a1"], 3 * D, D), colsum(dqkv, 3 * D, inv)
            g_g1, g_be1 = zeros(D), zeros(D)
            dX = layernorm_bwd(s["x_in"], g1.detach().float().contiguous(), dA1, dXm, g_g1, g_be1)
            g_blocks.append([g_win, g_bin, g_wo, g_bo, g_g1.mul_(inv), g_be1.mul_(inv), g_wfc, g_bfc, g_wpr, g_bpr,
                             g_g2.mul_(inv), g_be2.mul_(inv)])
        g_blocks.reverse()
        # ---- attn_pred branch: dF += (g + g^T) F with g = dA * s * (1 - s)
        if dattn is not None:
            G = torch.zeros(N, P, Pp, dtype=torch.float16, device=dev)
            rc = _lib.lib().wc_sigmoid_gram_bwd(_lib.ptr(dattn.contiguous()), _lib.ptr(sv["attn_pred"]), _lib.ptr(G), N, P, Pp,
                                                ctypes.c_float(sc), _lib.stream_ptr())
            _lib.check(rc, "wc_sigmoid_gram_bwd")
            FT = transpose16(sv["F16"].view(N, P, D), P, D, Pp)
            gemm(G.view(1, N, P, Pp), FT.view(1, N, D, Pp), dX.view(1, N, P, D), K=P, resid=dX)
        # ---- Dropout2d backward, linear_fuse, the 11 MLPs
        dF0, dF16 = dropout2d(dX, sv["mask"], N, P)
        g_wf = wgrad(dF16, sv["Zcat"], D, D * N_TAPS).view(D, D * N_TAPS, 1, 1)
        g_bf = colsum(dF0, D, inv)
        dZ = torch.empty(M, D * N_TAPS, dtype=torch.float16, device=dev)
        gemm(dF16, tr_w(wf), dZ)                                                                     # [M,2816]
        g_b2 = colsum(dZ, D * N_TAPS, inv).view(N_TAPS, D)
        dZT = transpose16(dZ, M, D * N_TAPS, Mp)[0].view(N_TAPS, D, Mp)
        H1 = sv["H1"].view(N_TAPS, M, D)
        H1T = transpose16(H1, M, D, Mp)                                                              # [11,256,Mp]
        g_w2 = torch.empty(N_TAPS, D, D, dtype=torch.float32, device=dev)
        gemm(dZT.unsqueeze(1), H1T.unsqueeze(1), g_w2.unsqueeze(1), alpha=inv)
        W2T = transpose16(torch.stack([f[2].detach().float() for f in fuse]), D, D)                  # [11,256(in),256(out)]
        dH1 = torch.empty(N_TAPS, M, D, dtype=torch.float32, device=dev)
        gemm(dZ.view(M, N_TAPS, D).permute(1, 0, 2).unsqueeze(1), W2T.unsqueeze(1), dH1.unsqueeze(1))
        dH16 = act_bwd(ACT_RELU, dH1, H1)
        dH1T = transpose16(dH16, M, D, Mp)                                                           # [11,256,Mp]
        tok = sv["tok"]                                                                              # [11,N,P,768] view
        tokT = torch.zeros(N_TAPS, E_VIT, Mp, dtype=torch.float16, device=dev)
        for i in range(N_TAPS):                                                                      # [P,768] per image -> columns n*P..
            rc = _lib.lib().wc_transpose_f16(_lib.ptr(tok[i]), 0, _lib.ptr(tokT[i]), N, P, E_VIT, _c(tok.stride(2)), _c(Mp),
                                             _c(tok.stride(1)), _c(P), ctypes.c_float(1.0), _lib.stream_ptr())
            _lib.check(rc, "wc_transpose_f16")
        g_w1 = torch.empty(N_TAPS, D, E_VIT, dtype=torch.float32, device=dev)
        gemm(dH1T.unsqueeze(1), tokT.unsqueeze(1), g_w1.unsqueeze(1), alpha=inv)
        g_b1 = torch.empty(N_TAPS, D, dtype=torch.float32, device=dev)
        for i in range(N_TAPS):
            colsum(dH16[i], D, inv, out=g_b1[i])
        out = []
        for i in range(N_TAPS):
            out += [g_w1[i], g_b1[i], g_w2[i], g_b2[i]]
        out += [g_wf, g_bf]
        for g in g_blocks:
            out += g
        out += [g_wp.view(C, D, 1, 1), g_bp]
        out = [g.to(p.dtype) if p.requires_grad else None for g, p in zip(out, params)]
        return (None, None, None, *out)


DecoderFunction.scale_override = None      # float: run backward with this loss scale, no host read (set by TrainStep for a capture)
DecoderFunction.last_scale = 1.0           # the scale the last eager backward chose
DecoderFunction.last_amax = None           # device scalar: max |incoming gradient| of the last backward


def loss_scale_for(amax: float) -> float:
    """Power of two that brings the largest incoming gradient to [2^7, 2^8]."""
    return 2.0 ** (8 - math.ceil(math.log2(amax))) if amax > 0 and math.isfinite(amax) else 1.0


def make_dropout_mask(N: int, p: float, training: bool, device) -> torch.Tensor:
    """Dropout2d(p) channel mask [N, 256] holding 0 or 1/(1-p) (ones when not training)."""
    if not training or p <= 0:
        return torch.ones(N, D, dtype=torch.float32, device=device)
    keep = (torch.rand(N, D, device=device) >= p).float()
    return keep / (1.0 - p)


class TrainStep:
    """One optimisation step of the reference's training loop (scripts/dist_clip_voc.py:238-267) behind one call:

        step = TrainStep(model, optimizer)          # optimizer over model.get_param_groups()[3]
        loss = step(img, img_names)                 # forward mode='train', aff / seg loss, backward, optimizer.step()

    The step is ~300 kernel launches of a few microseconds each; launched from Python it is host-bound (measured 10-16 ms of
    host time for 4.3 ms of device work at batch 4 x 320^2).  With ``use_graphs`` the device half — WeCLIP.run (frozen encoder,
    pseudo-labels, trainable decoder), F.interpolate, the two losses, the whole backward and the optimiser step — is captured
    into a CUDA graph per (batch shape, label-set signature) and replayed; the host half (label PNGs -> class lists,
    WeCLIP.prepare) stays outside and its index vectors are copied into the graph's buffers before each replay.  The first
    batch of a signature runs eagerly (it is a real step) and is captured for the next one.  Graph mode needs an optimiser
    whose state lives on the device (``capturable=True``; a tensor ``lr`` lets a scheduler keep working); otherwise the step
    runs eagerly.  The returned loss is a device scalar (no host synchronisation)."""

    def __init__(self, model, optimizer, radius: int = 8, aff_weight: float = 0.1, ignore_index: int = 255,
                 use_graphs: bool = True, max_graphs: int = 32):
        self.model, self.opt = model, optimizer
        self.radius, self.aff_weight, self.ignore_index = radius, aff_weight, ignore_index
        capturable = all(g.get("capturable", False) for g in optimizer.param_groups)
        self.use_graphs = bool(use_graphs and capturable)
        self.max_graphs = max_graphs
        if self.use_graphs:
            # the parameters' AccumulateGrad nodes were created on the caller's stream, the step runs on its own: torch orders the
            # two with events (captured as graph edges) and would warn about it on every backward
            setter = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
            if setter is not None:
                setter(False)
        self._graphs = {}
        self._stream = None
        self._pool = None
        self.graph_captures = self.graph_replays = 0

    def _device_step(self, img, meta):
        import torch.nn.functional as F
        from . import losses as LS
        segs, cam, attn_pred = self.model.run(img, meta)                                             # dist_clip_voc.py:246
        segs_up = F.interpolate(segs, size=cam.shape[1:], mode="bilinear", align_corners=False)      # :250
        attn_loss, _, _ = LS.aff_loss_from_labels(attn_pred, cam, radius=self.radius, ignore_index=self.ignore_index)
        seg_loss = LS.get_seg_loss(segs_up, cam, ignore_index=self.ignore_index)                     # :256
        loss = seg_loss + self.aff_weight * attn_loss                                                # :258
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        self.opt.step()
        return loss.detach()

    def _device_meta(self, meta_h, dev):
        from .model import ForwardMeta
        m = ForwardMeta()
        m.__dict__.update(meta_h.__dict__)
        m.rows = torch.empty(meta_h.rows_host.shape, dtype=torch.int64, device=dev)
        m.lut = torch.empty(meta_h.lut_host.shape, dtype=torch.int64, device=dev)
        return m

    @staticmethod
    def _refresh(ent, img, meta_h):
        ent["img"].copy_(img, non_blocking=True)
        ent["meta"].rows.copy_(meta_h.rows_host.pin_memory(), non_blocking=True)
        ent["meta"].lut.copy_(meta_h.lut_host.pin_memory(), non_blocking=True)

    def __call__(self, img, img_names):
        _lib.require_device()
        model = self.model
        eng = model.engine
        if eng is None:
            raise _lib.WeclipError("WeCLIP weights are not loaded (encoder + decoder state dict required)")
        dev = eng.device
        img = img.to(dev, torch.float32)
        N, _, H, W = img.shape
        model.iter_num += 1
        if not self.use_graphs:
            with torch.enable_grad():
                return self._device_step(img, model.prepare(img_names, "train", N, H, W))
        meta_h = model.prepare(img_names, "train", N, H, W, device_tensors=False)
        if not meta_h.cam_path:
            raise _lib.WeclipError("the training step needs the pseudo-label path (one name per image)")
        model._text_table(dev)
        key = (N, H, W, meta_h.signature(), id(eng), eng.generation, getattr(model, "_text_key", None), model.training,
               float(model.decoder_fts_fuse.dropout.p))
        if self._stream is None:
            self._stream = torch.cuda.Stream(dev)
            self._pool = torch.cuda.graph_pool_handle()
        caller = torch.cuda.current_stream(dev)
        st = self._stream
        st.wait_stream(caller)
        ent = self._graphs.get(key)
        with torch.cuda.stream(st):
            if ent is None:
                ent = dict(img=torch.empty_like(img), meta=self._device_meta(meta_h, dev))
                self._refresh(ent, img, meta_h)
                with torch.enable_grad():
                    loss = self._device_step(ent["img"], ent["meta"])          # the real step for this batch (and the warm-up)
                    loss = loss.clone()
                    st.synchronize()
                    if len(self._graphs) >= self.max_graphs:
                        self._graphs.pop(next(iter(self._graphs)))
                    g = torch.cuda.CUDAGraph()
                    self.opt.zero_grad(set_to_none=True)      # gradients of the captured step live in the graph's own pool
                    ent["scale"] = DecoderFunction.scale_override = DecoderFunction.last_scale
                    try:
                        with torch.cuda.graph(g, pool=self._pool, stream=st, capture_error_mode="thread_local"):
                            ent["loss"] = self._device_step(ent["img"], ent["meta"])
                    finally:
                        DecoderFunction.scale_override = None
                    ent["amax"] = DecoderFunction.last_amax                    # static tensor of the captured backward
                    ent["amax_h"] = torch.zeros(1).pin_memory()
                    ent["amax_ev"] = None
                ent["graph"] = g
                self._graphs[key] = ent
                self.graph_captures += 1
            else:
                self._refresh(ent, img, meta_h)
                ent["graph"].replay()
                self.graph_replays += 1
                loss = ent["loss"]
                # loss-scale watch: every 32 replays the gradient maximum travels to the host (asynchronously); a graph whose
                # captured scale has drifted more than 2^4 from what that maximum asks for is dropped and captured again
                if ent["amax_ev"] is not None and ent["amax_ev"].query():
                    want = loss_scale_for(float(ent["amax_h"][0]))
                    ent["amax_ev"] = None
                    if not (2.0 ** -4 <= want / ent["scale"] <= 2.0 ** 4):
                        self._graphs.pop(key, None)
                elif ent["amax_ev"] is None and self.graph_replays % 32 == 0 and ent["amax"] is not None:
                    ent["amax_h"].copy_(ent["amax"].reshape(1), non_blocking=True)
                    ent["amax_ev"] = torch.cuda.Event()
                    ent["amax_ev"].record(st)
        img.record_stream(st)
        caller.wait_stream(st)
        return loss
