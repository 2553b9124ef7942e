"""Host-side engine: packs a reference-keyed state dict into the C-ABI weight structs and drives the
coarse C entry points (encoder / decoder / Grad-CAM / affinity refinement / CAM upsample / PAR).

PyTorch is only the device-memory / stream plumbing here; every arithmetic step runs in
libweclip_sm100.so.  There is no fallback path.
"""
from __future__ import annotations

import ctypes
import math
from ctypes import POINTER, Structure, byref, c_double, c_float, c_int, c_size_t, c_void_p
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib

E_VIT, HEADS_VIT, E_DEC, HEADS_DEC, N_TAPS = 768, 12, 256, 8, 11


class BlockW(Structure):
    _fields_ = [(n, c_void_p) for n in ("w_in", "b_in", "w_out", "b_out", "w_fc", "b_fc", "w_proj", "b_proj",
                                        "ln1_w", "ln1_b", "ln2_w", "ln2_b")]


class VitW(Structure):
    _fields_ = [("conv1_w", c_void_p), ("class_emb", c_void_p), ("pos_emb", c_void_p), ("pos_grid", c_int),
                ("ln_pre_w", c_void_p), ("ln_pre_b", c_void_p), ("blocks", BlockW * 12),
                ("ln_post_w", c_void_p), ("ln_post_b", c_void_p), ("proj", c_void_p),
                ("w_proj_t", c_void_p), ("w_fc_t", c_void_p), ("w_out_t", c_void_p), ("w_in_f32", c_void_p),
                ("logit_scale_exp", c_float)]


class DecW(Structure):
    _fields_ = [("fuse_w1", c_void_p), ("fuse_b1", c_void_p), ("fuse_w2", c_void_p), ("fuse_b2", c_void_p),
                ("fuse_wf", c_void_p), ("fuse_bf", c_void_p), ("blocks", BlockW * 3),
                ("pred_w", c_void_p), ("pred_b", c_void_p), ("num_classes", c_int)]


class Engine:
    """Owns device copies of the weights (bf16 GEMM operands, fp32 everything else) and the workspaces."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], num_classes: int, device):
        _lib.require_device()
        self.device = torch.device(device)
        self.num_classes = num_classes
        self._keep: List[torch.Tensor] = []
        self._ws: Dict[tuple, torch.Tensor] = {}
        self.generation = 0          # bumped whenever the device weights are replaced (captured CUDA graphs become stale)
        self.vit = VitW()
        self.dec = DecW()
        self.load_state_dict(state_dict)

    # ------------------------------------------------------------------ weights
    def _bf(self, t):
        t = t.detach().to(self.device, torch.float32).contiguous().to(torch.float16).contiguous()
        self._keep.append(t)
        return t.data_ptr()

    def _f32(self, t):
        t = t.detach().to(self.device, torch.float32).contiguous()
        self._keep.append(t)
        return t.data_ptr()

    def _block(self, bw: BlockW, sd, p):
        bw.w_in, bw.b_in = self._bf(sd[p + ".attn.in_proj_weight"]), self._f32(sd[p + ".attn.in_proj_bias"])
        bw.w_out, bw.b_out = self._bf(sd[p + ".attn.out_proj.weight"]), self._f32(sd[p + ".attn.out_proj.bias"])
        bw.w_fc, bw.b_fc = self._bf(sd[p + ".mlp.c_fc.weight"]), self._f32(sd[p + ".mlp.c_fc.bias"])
        bw.w_proj, bw.b_proj = self._bf(sd[p + ".mlp.c_proj.weight"]), self._f32(sd[p + ".mlp.c_proj.bias"])
        bw.ln1_w, bw.ln1_b = self._f32(sd[p + ".ln_1.weight"]), self._f32(sd[p + ".ln_1.bias"])
        bw.ln2_w, bw.ln2_b = self._f32(sd[p + ".ln_2.weight"]), self._f32(sd[p + ".ln_2.bias"])

    def load_state_dict(self, sd: Dict[str, torch.Tensor]):
        """Keys as in the reference checkpoint: encoder.visual.*, decoder_fts_fuse.*, decoder.*"""
        self._keep.clear()
        self.generation += 1
        v, w = "encoder.visual", self.vit
        w.conv1_w = self._bf(sd[v + ".conv1.weight"].reshape(E_VIT, -1))
        w.class_emb = self._f32(sd[v + ".class_embedding"])
        pos = sd[v + ".positional_embedding"]
        w.pos_emb = self._f32(pos)
        w.pos_grid = int(math.isqrt(pos.shape[0] - 1))
        w.ln_pre_w, w.ln_pre_b = self._f32(sd[v + ".ln_pre.weight"]), self._f32(sd[v + ".ln_pre.bias"])
        for i in range(12):
            self._block(w.blocks[i], sd, f"{v}.transformer.resblocks.{i}")
        w.ln_post_w, w.ln_post_b = self._f32(sd[v + ".ln_post.weight"]), self._f32(sd[v + ".ln_post.bias"])
        w.proj = self._f32(sd[v + ".proj"])
        p11 = f"{v}.transformer.resblocks.11"
        w.w_proj_t = self._bf(sd[p11 + ".mlp.c_proj.weight"].float().t())
        w.w_fc_t = self._bf(sd[p11 + ".mlp.c_fc.weight"].float().t())
        w.w_out_t = self._bf(sd[p11 + ".attn.out_proj.weight"].float().t())
        w.w_in_f32 = self._f32(sd[p11 + ".attn.in_proj_weight"])
        w.logit_scale_exp = float(sd["encoder.logit_scale"].float().exp())
        self.load_decoder_state_dict(sd)

    def refresh_decoder(self, sd):
        """Copy new decoder weights into the EXISTING device buffers (addresses stay fixed)."""
        shadow = Engine.__new__(Engine)
        shadow.device, shadow._keep, shadow.dec, shadow.num_classes = self.device, [], DecW(), self.num_classes
        n0 = len(self._keep) - self._n_dec
        shadow.load_decoder_state_dict(sd)
        old = self._keep[n0:]
        assert len(old) == len(shadow._keep)
        for dst, src in zip(old, shadow._keep):
            dst.copy_(src)

    def load_decoder_state_dict(self, sd):
        n_before = len(self._keep)
        self._load_decoder(sd)
        self._n_dec = len(self._keep) - n_before

    def _load_decoder(self, sd):
        d, f = self.dec, "decoder_fts_fuse.linears_modulelist"
        d.fuse_w1 = self._bf(torch.stack([sd[f"{f}.{i}.proj.weight"] for i in range(N_TAPS)]))
        d.fuse_b1 = self._f32(torch.stack([sd[f"{f}.{i}.proj.bias"] for i in range(N_TAPS)]))
        d.fuse_w2 = self._bf(torch.stack([sd[f"{f}.{i}.proj_2.weight"] for i in range(N_TAPS)]))
        d.fuse_b2 = self._f32(torch.stack([sd[f"{f}.{i}.proj_2.bias"] for i in range(N_TAPS)]))
        d.fuse_wf = self._bf(sd["decoder_fts_fuse.linear_fuse.weight"].flatten(1))
        d.fuse_bf = self._f32(sd["decoder_fts_fuse.linear_fuse.bias"])
        for i in range(3):
            self._block(d.blocks[i], sd, f"decoder.transformer.resblocks.{i}")
        d.pred_w = self._bf(sd["decoder.linear_pred.weight"].flatten(1))
        d.pred_b = self._f32(sd["decoder.linear_pred.bias"])
        d.num_classes = int(sd["decoder.linear_pred.weight"].shape[0])
        self.num_classes = d.num_classes

    # ------------------------------------------------------------------ workspaces
    def _workspace(self, name: str, nbytes: int) -> torch.Tensor:
        # under CUDA-graph capture the scratch must outlive this call at a FIXED address for every replay: it comes from the
        # graph's own memory pool (the shared arena below may be re-allocated when a later, larger call grows it)
        if self.device.type == "cuda" and torch.cuda.is_current_stream_capturing():
            return torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        # one scratch arena per (call kind, CUDA stream): the msc+flip driver runs its scales on separate streams
        key = (name, torch.cuda.current_stream(self.device).cuda_stream if self.device.type == "cuda" else 0)
        ws = self._ws.get(key)
        if ws is None or ws.numel() < nbytes:
            ws = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
            self._ws[key] = ws
        return ws

    # ------------------------------------------------------------------ calls
    def encoder(self, img: torch.Tensor, n_attn: int = 0, attn_slots: Optional[int] = None,
                attn_imgs: Optional[int] = None):
        """clip_tool.py:31-41 generate_clip_fts.  Returns (taps fp16 [11,N*L,768], x_last fp32 [N*L,768],
        attn fp32 [attn_slots or n_attn, attn_imgs or N, L, L] with the last n_attn encoder maps in slots
        0..n_attn-1; attn_imgs = the leading images whose maps the CAM path will read)."""
        L = _lib.lib()
        img = img.contiguous().float()
        N, _, H, W = img.shape
        h, w = H // 16, W // 16
        Lt = h * w + 1
        taps = torch.empty(N_TAPS, N * Lt, E_VIT, dtype=torch.float16, device=self.device)
        x_last = torch.empty(N * Lt, E_VIT, dtype=torch.float32, device=self.device)
        slots = attn_slots if attn_slots is not None else n_attn
        n_ai = N if attn_imgs is None else int(attn_imgs)
        attn = torch.empty(max(slots, 1), n_ai, Lt, Lt, dtype=torch.float32, device=self.device) if slots else None
        # per (map, image) sum of attn[l][n][1:, 1:], accumulated by the kernel that writes the map (RFM layer score)
        self.last_attn_sums = torch.zeros(max(slots, 1), n_ai, dtype=torch.float64, device=self.device) if slots else None
        ws = self._workspace("enc", L.wc_encoder_workspace_bytes(N, H, W))
        rc = L.wc_encoder_fwd(byref(self.vit), _lib.ptr(img), N, H, W, _lib.ptr(taps), _lib.ptr(x_last),
                              _lib.ptr(attn), _lib.ptr(self.last_attn_sums), n_attn, n_ai, _lib.ptr(ws), c_size_t(ws.numel()),
                              _lib.stream_ptr())
        _lib.check(rc, "wc_encoder_fwd")
        return taps, x_last, attn

    def decoder(self, taps: torch.Tensor, N: int, h: int, w: int, want_fts: bool = False):
        """segformer_head + attn_pred + DecoderTransformer.  Returns (seg [N,C,h,w], attn_pred [N,P,P], fts|None)."""
        L = _lib.lib()
        P = h * w
        seg = torch.empty(N, self.num_classes, h, w, dtype=torch.float32, device=self.device)
        attn_pred = torch.empty(N, P, P, dtype=torch.float32, device=self.device)
        fts = torch.empty(N * P, E_DEC, dtype=torch.float32, device=self.device) if want_fts else None
        ws = self._workspace("dec", L.wc_decoder_workspace_bytes(N, h, w, self.num_classes))
        rc = L.wc_decoder_fwd(byref(self.dec), _lib.ptr(taps), N, h, w, _lib.ptr(seg), _lib.ptr(attn_pred),
                              _lib.ptr(fts), _lib.ptr(ws), c_size_t(ws.numel()), _lib.stream_ptr())
        _lib.check(rc, "wc_decoder_fwd")
        return seg, attn_pred, fts

    def gradcam(self, x_last: torch.Tensor, N: int, h: int, w: int, text: torch.Tensor, text_off: Sequence[int],
                k_off: Sequence[int], attn_last_out: Optional[torch.Tensor] = None, want_probs: bool = False,
                attn_last_sum: Optional[torch.Tensor] = None):
        """forward_last_layer + Grad-CAM for every (image, class).  text fp32 [sumT,512] (per image: its K fg rows
        then the bg rows).  Returns (cams fp32 [sumK,h,w], attn_last fp32 [N,L,L], probs|None)."""
        L = _lib.lib()
        P, Lt = h * w, h * w + 1
        total = int(k_off[-1])
        max_k = max(k_off[i + 1] - k_off[i] for i in range(N))
        text = text.to(self.device, torch.float32).contiguous()
        cams = torch.empty(total, h, w, dtype=torch.float32, device=self.device)
        attn_last = attn_last_out if attn_last_out is not None else torch.empty(
            N, Lt, Lt, dtype=torch.float32, device=self.device)
        assert attn_last.is_contiguous() and attn_last.numel() == N * Lt * Lt
        probs = torch.empty(int(text_off[-1]), dtype=torch.float32, device=self.device) if want_probs else None
        ws = self._workspace("gc", L.wc_gradcam_workspace_bytes(N, h, w, total, max_k))
        rc = L.wc_gradcam(byref(self.vit), _lib.ptr(x_last), N, h, w, _lib.ptr(text), _lib.int_array(text_off),
                          _lib.int_array(k_off), _lib.ptr(cams), _lib.ptr(attn_last), _lib.ptr(attn_last_sum), _lib.ptr(probs),
                          _lib.ptr(ws), c_size_t(ws.numel()), _lib.stream_ptr())
        _lib.check(rc, "wc_gradcam")
        return cams, attn_last, probs

    def affinity_refine(self, attn: torch.Tensor, nl: int, attn_pred: torch.Tensor, N: int, h: int, w: int,
                        use_rfm: bool, cams: torch.Tensor, k_off: Sequence[int], box_thr: float, want_masks=False,
                        attn_sums: Optional[torch.Tensor] = None):
        """clip_tool.py:152-191.  attn fp32 [nl,N,L,L]; returns (cams_refined [sumK,h,w], box masks|None); with want_masks the
        RFM layer selection + margins [N,2,nl] are kept in ``self.last_rfm_sel`` (test hook)."""
        L = _lib.lib()
        total = int(k_off[-1])
        out = torch.empty(total, h, w, dtype=torch.float32, device=self.device)
        masks = torch.empty(total, h, w, dtype=torch.float32, device=self.device) if want_masks else None
        sel = torch.zeros(N, 2, nl, dtype=torch.float32, device=self.device) if want_masks else None
        self.last_rfm_sel = sel
        ws = self._workspace("aff", L.wc_affinity_workspace_bytes(N, h, w, total, nl))
        assert attn_sums is None or (attn_sums.dtype == torch.float64 and attn_sums.shape[0] >= nl and attn_sums.shape[1] == N)
        rc = L.wc_affinity_refine(_lib.ptr(attn), nl, _lib.ptr(attn_sums), _lib.ptr(attn_pred), N, h, w, int(use_rfm), _lib.ptr(cams),
                                  _lib.int_array(k_off), c_double(box_thr), _lib.ptr(out), _lib.ptr(masks),
                                  _lib.ptr(sel), _lib.ptr(ws), c_size_t(ws.numel()), _lib.stream_ptr())
        _lib.check(rc, "wc_affinity_refine")
        return out, masks

    def cam_upsample_bg(self, cams: torch.Tensor, k_off: Sequence[int], N: int, h: int, w: int, Hc: int, Wc: int,
                        bg_pow: float = 1.0) -> torch.Tensor:
        """generate_cam_label + bg channel -> packed [sumK+N, Hc, Wc] (PAR's ragged input)."""
        total = int(k_off[-1])
        out = torch.empty(total + N, Hc, Wc, dtype=torch.float32, device=self.device)
        stats = torch.empty(2 * total, dtype=torch.float32, device=self.device)
        rc = _lib.lib().wc_cam_upsample_bg(_lib.ptr(cams), _lib.int_array(k_off), N, h, w, Hc, Wc, c_float(bg_pow),
                                           _lib.ptr(out), _lib.ptr(stats), _lib.stream_ptr())
        _lib.check(rc, "wc_cam_upsample_bg")
        return out
