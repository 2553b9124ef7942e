"""tcgen05 GEMM (wc_gemm_f16 through the C ABI) against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

from weclip_b200 import ops

pytestmark = pytest.mark.gpu


def _ref(A, B, bias, resid, act, alpha):
    x = alpha * torch.matmul(A.float(), B.float().transpose(-1, -2))
    if bias is not None:
        x = x + bias
    if act == "relu":
        x = torch.relu(x)
    elif act == "quickgelu":
        x = x * torch.sigmoid(1.702 * x)
    elif act == "sigmoid":
        x = torch.sigmoid(x)
    if resid is not None:
        x = x + resid
    return x


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 128, 256), (256, 384, 768), (401, 768, 768), (2050, 2304, 768),
                                   (130, 21, 256), (100, 64, 32), (77, 200, 401), (1025, 3072, 768), (512, 768, 3072),
                                   # >= 148 tiles of 128 x 256: the CTA-pair kernel (cta_group::2, 256 x 256 tiles), with an odd
                                   # number of m-blocks, a ragged last n-tile and K not a multiple of 64
                                   (9232, 768, 768), (4230, 1100, 328), (16400, 2304, 768),
                                   # A > 48 MB with a small B: n-fastest tile raster
                                   (8200, 520, 3072)])
def test_gemm_shapes(M, N, K):
    torch.manual_seed(M + N + K)
    Kp = (K + 7) // 8 * 8
    A = torch.randn(M, Kp, device="cuda").half()[:, :K]
    B = torch.randn(N, Kp, device="cuda").half()[:, :K]
    bias = torch.randn(N, device="cuda")
    out = ops.gemm_f16(A, B, bias=bias)
    want = _ref(A, B, bias, None, "none", 1.0)
    # bf16 inputs are exact in both; fp32 accumulation order differs: tolerance 1e-3 * sqrt(K)
    assert (out - want).abs().max().item() < 2e-3 * (K ** 0.5)
    rel = (out - want).norm() / want.norm()
    assert rel.item() < 1e-5


@pytest.mark.parametrize("act", ["none", "relu", "quickgelu", "sigmoid"])
@pytest.mark.parametrize("out_f16", [False, True])
def test_gemm_epilogues(act, out_f16):
    torch.manual_seed(3)
    M, N, K = 300, 272, 192
    A = (0.1 * torch.randn(M, K, device="cuda")).half()
    B = torch.randn(N, K, device="cuda").half()
    bias = torch.randn(N, device="cuda")
    resid = torch.randn(M, N, device="cuda")
    out = ops.gemm_f16(A, B, bias=bias, resid=resid, act=act, alpha=0.5,
                        out_dtype=torch.float16 if out_f16 else torch.float32)
    want = _ref(A, B, bias, resid, act, 0.5)
    tol = 3e-2 if out_f16 else 2e-3    # bf16 output rounding: 2^-9 relative on |x| <~ 6
    assert (out.float() - want).abs().max().item() < tol


def test_gemm_batched_broadcast_transposed_inplace_residual():
    torch.manual_seed(4)
    b, M, N, K = 5, 150, 90, 64
    A = torch.randn(b, M, K, device="cuda").half()
    B = torch.randn(b, N, K, device="cuda").half()
    out = ops.gemm_f16(A, B, alpha=0.125)
    assert (out - 0.125 * torch.matmul(A.float(), B.float().transpose(1, 2))).abs().max().item() < 1e-2
    Bs = torch.randn(N, K, device="cuda").half()
    out = ops.gemm_f16(A, Bs, trans=True)
    want = torch.matmul(A.float(), Bs.float().t()).transpose(1, 2)
    assert out.shape == (b, N, M) and (out - want).abs().max().item() < 1e-2
    x = torch.randn(b, M, N, device="cuda")
    want = x + torch.matmul(A.float(), B.float().transpose(1, 2))
    ops.gemm_f16(A, B, resid=x, out=x)     # in-place residual accumulate (clip/model.py:212-213)
    assert (x - want).abs().max().item() < 1e-2
