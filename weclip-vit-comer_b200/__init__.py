"""weclip_b200 — B200-native (sm_100a) implementation of WeCLIP's per-image inference hot path.

Drop-in surface (mirrors the reference's module paths):
  * ``weclip_b200.model.WeCLIP``      <- WeCLIP_model/model_attn_aff_voc.py:60-175 (and the COCO twin)
  * ``weclip_b200.par.PAR``           <- WeCLIP_model/PAR.py:26-92
All arithmetic runs in ``libweclip_sm100.so`` (hand-written CUDA behind a C ABI, see
``include/weclip_b200.h``); there is no CPU or PyTorch fallback on the product path.
"""
__version__ = "0.1.0"
