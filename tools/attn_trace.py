"""Developer trace (GPU box): per-tile clock64 breakdown of attn_fwd_kernel's softmax warps.
    make -C weclip-vit-comer_b200/csrc EXTRA=-DAT_TRACE (after touching attention.cu) && python tools/attn_trace.py
Nothing here is a bench value."""
import sys, os, ctypes
sys.path.insert(0, os.getcwd())
import torch
from weclip_b200 import _lib
L, N, heads = 1601, 16, 12
g = torch.Generator(device="cuda").manual_seed(L)
q = (torch.randn(N, heads, L, 64, device="cuda", generator=g) * 0.25).half()
k = torch.randn(N, heads, L, 64, device="cuda", generator=g).half()
v = torch.randn(N, heads, L, 64, device="cuda", generator=g).half()
o = torch.empty(N, L, heads * 64, dtype=torch.float16, device="cuda")
lse = torch.zeros(N * heads * L + 16384, dtype=torch.float32, device="cuda")
Lp = (L + 7) // 8 * 8
for _ in range(3):
    rc = _lib.lib().wc_attention_fwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(o), _lib.ptr(lse), ctypes.c_void_p(0), N, heads, L, Lp, 64, _lib.stream_ptr())
    assert rc == 0
torch.cuda.synchronize()
raw = lse.view(torch.int32).flatten().cpu()
off = N * heads * L
for name, base in (("warp2", off), ("warp6 (half 1, same scheduler)", off + 4096)):
    t = raw[base: base + 13 * 8].view(13, 8).to(torch.int64) & 0xffffffff
    print(name, "per tile: [ld_wait+probes, max+s_empty, any/rescale, residual barrier waits, turn wait, exp+P+hand-over, fence+arrive] total, gap to next tile start")
    for j in range(13):
        d = [(int(t[j, i + 1]) - int(t[j, i])) & 0xffffffff for i in range(7)]
        tot = (int(t[j, 7]) - int(t[j, 0])) & 0xffffffff
        gap = ((int(t[j + 1, 0]) - int(t[j, 7])) & 0xffffffff) if j < 12 else -1
        print(j, d, tot, gap)

