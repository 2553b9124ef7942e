"""Import the REAL reference (/root/reference, read-only) in the build container.  TEST INFRASTRUCTURE ONLY.

/root/reference does not exist on the GPU box: nothing in ``-m gpu`` tests, ``smoke()`` or ``bench.py``
may import this module.  It exists to (a) pin ``oracle/weclip_oracle.py`` and (b) generate the golden
vectors under ``tests/golden/`` (see ``oracle/make_golden.py``).

Shims: ``sys.modules`` stubs for ftfy / lxml / ttach / mmcv (imported but unused on the hot
path), and a CPU-only no-op for the hard-coded ``.cuda()`` calls (clip_tool.py:32,37,108-109,180,188;
model_attn_aff_voc.py:39,45,89,160-166).  The reference source itself is executed unmodified.
"""
from __future__ import annotations

import os
import sys
import tempfile
import types
from collections import OrderedDict

import numpy as np
import torch

REFERENCE_ROOT = "/root/reference"
_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "WeCLIP_model"))


def _install_shims():
    if "ftfy" not in sys.modules:
        m = types.ModuleType("ftfy"); m.fix_text = lambda s: s; sys.modules["ftfy"] = m
    if "lxml" not in sys.modules:
        m = types.ModuleType("lxml"); e = types.ModuleType("lxml.etree"); m.etree = e
        sys.modules["lxml"] = m; sys.modules["lxml.etree"] = e
    if "ttach" not in sys.modules:
        sys.modules["ttach"] = types.ModuleType("ttach")
    if "mmcv" not in sys.modules:
        m = types.ModuleType("mmcv"); c = types.ModuleType("mmcv.cnn"); c.ConvModule = object; m.cnn = c
        sys.modules["mmcv"] = m; sys.modules["mmcv.cnn"] = c
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
        torch.nn.Module.cuda = lambda self, *a, **k: self
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


def _clip_state_dict(sd):
    """Full CLIP state dict (visual from ``sd``; a tiny random text tower so build_model accepts it)."""
    out = OrderedDict()
    for k, v in sd.items():
        if k.startswith("encoder."):
            out[k[len("encoder."):]] = v.clone()
    g = torch.Generator().manual_seed(1234)
    tw, layers, ctx, vocab = 512, 1, 77, 49408
    out["positional_embedding"] = 0.01 * torch.randn(ctx, tw, generator=g)
    out["text_projection"] = tw ** -0.5 * torch.randn(tw, 512, generator=g)
    out["token_embedding.weight"] = 0.02 * torch.randn(vocab, tw, generator=g)
    out["ln_final.weight"] = torch.ones(tw); out["ln_final.bias"] = torch.zeros(tw)
    for i in range(layers):
        p = f"transformer.resblocks.{i}"
        out[p + ".attn.in_proj_weight"] = tw ** -0.5 * torch.randn(3 * tw, tw, generator=g)
        out[p + ".attn.in_proj_bias"] = torch.zeros(3 * tw)
        out[p + ".attn.out_proj.weight"] = tw ** -0.5 * torch.randn(tw, tw, generator=g)
        out[p + ".attn.out_proj.bias"] = torch.zeros(tw)
        out[p + ".ln_1.weight"] = torch.ones(tw); out[p + ".ln_1.bias"] = torch.zeros(tw)
        out[p + ".mlp.c_fc.weight"] = tw ** -0.5 * torch.randn(4 * tw, tw, generator=g)
        out[p + ".mlp.c_fc.bias"] = torch.zeros(4 * tw)
        out[p + ".mlp.c_proj.weight"] = (4 * tw) ** -0.5 * torch.randn(tw, 4 * tw, generator=g)
        out[p + ".mlp.c_proj.bias"] = torch.zeros(tw)
        out[p + ".ln_2.weight"] = torch.ones(tw); out[p + ".ln_2.bias"] = torch.zeros(tw)
    return out


def build_reference_model(sd, fg_text, bg_text, root, dataset="voc", num_classes=21):
    """Instantiate the reference WeCLIP through its own ctor (clip.load -> build_model), then load our
    decoder weights and substitute the synthetic text features (plain tensor attributes,
    model_attn_aff_voc.py:81-82)."""
    _install_shims()
    import contextlib, io
    if dataset == "voc":
        from WeCLIP_model.model_attn_aff_voc import WeCLIP
    else:
        from WeCLIP_model.model_attn_aff_coco import WeCLIP
    with tempfile.NamedTemporaryFile(suffix=".pt", delete=False) as f:
        ckpt = f.name
    torch.save(_clip_state_dict(sd), ckpt)
    try:
        with contextlib.redirect_stdout(io.StringIO()):   # ctor prints every parameter name
            model = WeCLIP(num_classes=num_classes, clip_model=ckpt, embedding_dim=256,
                           in_channels=[768, 768, 768, 768], dataset_root_path=root, device="cpu")
    finally:
        os.unlink(ckpt)
    own = {k: v for k, v in sd.items() if not k.startswith("encoder.")}
    missing, unexpected = model.load_state_dict(own, strict=False)
    assert not unexpected, unexpected
    assert all(k.startswith("encoder.") or k == "par.kernel" for k in missing), missing
    model.fg_text_features = fg_text.clone()
    model.bg_text_features = bg_text.clone()
    model.eval()
    return model


def reference_modules():
    """Handles to the reference functions every kernel is diffed against (SURVEY.md §8c last row)."""
    _install_shims()
    from WeCLIP_model.PAR import PAR
    from clip.clip_tool import compute_trans_mat, generate_cam_label
    from clip.utils import scoremap2bbox
    from pytorch_grad_cam.utils.image import scale_cam_image
    from utils import evaluate
    return dict(PAR=PAR, compute_trans_mat=compute_trans_mat, scoremap2bbox=scoremap2bbox,
                scale_cam_image=scale_cam_image, generate_cam_label=generate_cam_label, evaluate=evaluate)


def reference_training_fns():
    """The label-prep / loss functions of SURVEY.md §8f-3.  ``get_mask_by_radius`` lives in a script whose top level
    needs omegaconf etc., so only that one function's source is extracted (ast) and executed from where it lies."""
    import ast
    import numpy as np
    _install_shims()
    if "matplotlib" not in sys.modules:          # utils/imutils.py:5 imports pyplot for plotting helpers only
        m = types.ModuleType("matplotlib"); pp = types.ModuleType("matplotlib.pyplot"); m.pyplot = pp
        sys.modules["matplotlib"] = m; sys.modules["matplotlib.pyplot"] = pp
    if "pydensecrf" not in sys.modules:          # utils/dcrf.py (CRF post-processing, off the path)
        m = types.ModuleType("pydensecrf"); d = types.ModuleType("pydensecrf.densecrf")
        u = types.ModuleType("pydensecrf.utils"); u.unary_from_labels = u.unary_from_softmax = None
        m.densecrf = d; m.utils = u
        sys.modules["pydensecrf"] = m; sys.modules["pydensecrf.densecrf"] = d; sys.modules["pydensecrf.utils"] = u
    if "imageio" not in sys.modules:             # utils/camutils.py:6, image dumps only
        sys.modules["imageio"] = types.ModuleType("imageio")
    from utils.camutils import cams_to_affinity_label
    from utils.losses import get_aff_loss, get_seg_loss
    path = os.path.join(REFERENCE_ROOT, "scripts", "dist_clip_voc.py")
    tree = ast.parse(open(path).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "get_mask_by_radius")
    ns = {"np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), path, "exec"), ns)
    return dict(cams_to_affinity_label=cams_to_affinity_label, get_aff_loss=get_aff_loss, get_seg_loss=get_seg_loss,
                get_mask_by_radius=ns["get_mask_by_radius"])


def reference_validate(dataset: str = "voc"):
    """``validate(model, dataset, test_scales)`` of test_msc_flip_voc.py:30-113 (coco twin test_msc_flip_coco.py:31-116),
    executed from where it lies.  The script's top level needs omegaconf / imageio / pydensecrf / datasets.*, so only that
    one function's source is extracted (ast) and executed unmodified in a namespace holding the names it uses
    (torch, F, np, tqdm, the reference's utils.evaluate, and ``args`` with resize_long / work_dir).
    Returns (validate, args): set ``args.resize_long`` / ``args.work_dir`` before calling."""
    import ast
    import torch.nn.functional as F
    _install_shims()
    from utils import evaluate
    path = os.path.join(REFERENCE_ROOT, f"test_msc_flip_{dataset}.py")
    tree = ast.parse(open(path).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "validate")
    args = types.SimpleNamespace(resize_long=512, work_dir=tempfile.mkdtemp(prefix="refval"))
    os.makedirs(os.path.join(args.work_dir, "logit"), exist_ok=True)
    ns = {"np": np, "torch": torch, "F": F, "tqdm": lambda it, **kw: it, "evaluate": evaluate, "args": args, "os": os}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), path, "exec"), ns)
    return ns["validate"], args
