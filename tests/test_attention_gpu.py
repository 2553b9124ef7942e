"""Fused tcgen05 attention (wc_attention_fwd through the C ABI) against a plain PyTorch fp32 reference of the same op
(clip/myAtt.py:53-64 scaled-dot-product attention and :323-326 head-mean map)."""
import pytest
import torch

from weclip_b200 import ops

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["0", "1", "2"], ids=["attn_fwd_kernel", "attn_fwd2_kernel", "attn_fwd3_kernel"], autouse=True)
def head_dim_64_kernel(request, monkeypatch):
    """The head_dim-64 forward kernels (attention.cu, the two-tile TMEM-operand kernel of attention2.cu, WC_ATTN2=1, and the
    two-CTAs-per-SM kernel of attention3.cu, WC_ATTN2=2) run every
    case; the selection is read per launch."""
    monkeypatch.setenv("WC_ATTN2", request.param)


def _ref(q, k, v):
    s = torch.matmul(q.float(), k.float().transpose(-1, -2))
    p = torch.softmax(s, dim=-1)
    o = torch.matmul(p, v.float())                       # [N,heads,L,64]
    N, h, L, d = o.shape
    return o.permute(0, 2, 1, 3).reshape(N, L, h * d), p.mean(1)


@pytest.mark.parametrize("N,heads,L", [(1, 1, 128), (1, 2, 401), (2, 12, 577), (1, 3, 1025), (1, 2, 1601), (2, 1, 50)])
def test_attention_matches_torch(N, heads, L, monkeypatch):
    g = torch.Generator(device="cuda").manual_seed(L + heads)
    q = (torch.randn(N, heads, L, 64, device="cuda", generator=g) * 0.125 * 2.0).half()
    k = (torch.randn(N, heads, L, 64, device="cuda", generator=g) * 1.5).half()
    v = torch.randn(N, heads, L, 64, device="cuda", generator=g).half()
    o, avg = ops.attention_fwd(q, k, v)
    o_ref, avg_ref = _ref(q, k, v)
    # P is rounded to fp16 before the P V product (as the reference's fp16 bmm does): 2^-11 relative on each term
    assert (o.float() - o_ref).abs().max().item() < 4e-3
    assert (avg - avg_ref).abs().max().item() < 2e-5
    assert (avg.sum(-1) - 1).abs().max().item() < 1e-4
    # the head-mean map from the 128 x 128 one-CTA-per-SM kernel (default: 128 x 64 tiles, two CTAs per SM): the same bits
    monkeypatch.setenv("WC_HEADMEAN64", "0")
    _, avg128 = ops.attention_fwd(q, k, v)
    assert torch.equal(avg128, avg)


@pytest.mark.parametrize("N,heads,L", [(1, 1, 64), (2, 8, 400), (1, 8, 576), (1, 3, 1024), (2, 8, 1600), (1, 2, 37)])
def test_attention_head_dim_32_matches_torch(N, heads, L):
    """Decoder blocks (TransDecoder.py:104-125: width 256, 8 heads x 32): same kernel, HD = 32 instantiation."""
    g = torch.Generator(device="cuda").manual_seed(3 * L + heads)
    q = (torch.randn(N, heads, L, 32, device="cuda", generator=g) * 0.35).half()
    k = (torch.randn(N, heads, L, 32, device="cuda", generator=g) * 1.5).half()
    v = torch.randn(N, heads, L, 32, device="cuda", generator=g).half()
    o, _ = ops.attention_fwd(q, k, v, want_avg=False)
    o_ref, _ = _ref(q, k, v)
    assert torch.isfinite(o.float()).all()
    assert (o.float() - o_ref).abs().max().item() < 4e-3


def test_attention_peaked_rows_and_large_scores():
    """Rows dominated by one key (scores ~ +-60): the two-pass log-sum-exp must not overflow or lose the small terms."""
    torch.manual_seed(0)
    N, heads, L = 1, 2, 300
    q = torch.randn(N, heads, L, 64, device="cuda").half()
    k = (torch.randn(N, heads, L, 64, device="cuda") * 1.0).half()
    k[:, :, 7] *= 6
    v = torch.randn(N, heads, L, 64, device="cuda").half()
    o, avg = ops.attention_fwd(q, k, v)
    o_ref, avg_ref = _ref(q, k, v)
    assert torch.isfinite(o.float()).all() and torch.isfinite(avg).all()
    assert (o.float() - o_ref).abs().max().item() < 8e-3
    assert (avg - avg_ref).abs().max().item() < 5e-5


@pytest.mark.parametrize("hd", [64, 32])
def test_attention_growing_maximum_forces_accumulator_rescale(hd):
    """The running maximum jumps by far more than 2^8 in later key tiles (and differently in the two 64-key halves of a
    tile): the lazily rescaled accumulators must still give the exact softmax."""
    torch.manual_seed(1)
    N, heads, L = 1, 2, 700
    q = torch.randn(N, heads, L, hd, device="cuda").half()
    k = (torch.randn(N, heads, L, hd, device="cuda") * 0.5).half()
    for pos, gain in ((130, 3.0), (200, 5.0), (333, 8.0), (480, 12.0), (650, 16.0)):
        k[:, :, pos] *= gain
    v = torch.randn(N, heads, L, hd, device="cuda").half()
    o, avg = ops.attention_fwd(q, k, v, want_avg=(hd == 64))
    o_ref, avg_ref = _ref(q, k, v)
    assert torch.isfinite(o.float()).all()
    assert (o.float() - o_ref).abs().max().item() < 8e-3
    if avg is not None:
        assert (avg - avg_ref).abs().max().item() < 5e-5
