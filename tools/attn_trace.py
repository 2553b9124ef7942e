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
if os.environ.get("A2"):
    # attention2.cu built with -DA2_TRACE: stamps 0 before s_full wait, 1 after, 2 after sweep 1, 3 after vote/rescale,
    # 4 after sweep 2 (P stored), 5 after hand-over
    for name, base in (("group 0 warp 0", off), ("group 1 warp 4", off + 4096)):
        t = raw[base: base + 13 * 8].view(13, 8).to(torch.int64) & 0xffffffff
        print(name, "per tile: [s_full wait, sweep (exp + P, incl. pv_done wait and a repeat), st wait+arrive] total, start since previous start")
        for j in range(13):
            d = [(int(t[j, b]) - int(t[j, a])) & 0xffffffff for a, b in ((0, 1), (1, 4), (4, 5))]
            tot = (int(t[j, 5]) - int(t[j, 0])) & 0xffffffff
            per = ((int(t[j, 0]) - int(t[j - 1, 0])) & 0xffffffff) if j else -1
            print(j, d, tot, per)
    t = raw[off + 8192: off + 8192 + 13 * 8].view(13, 2, 4).to(torch.int64) & 0xffffffff
    t0 = int(raw[off]) & 0xffffffff
    print("issuer per (tile j, group t): p_full wait start (since group 0's first stamp), wait, issue S + commit, issue P V + commit")
    for j in range(13):
        print(j, [((int(t[j, g, 0]) - t0) & 0xffffffff, (int(t[j, g, 1]) - int(t[j, g, 0])) & 0xffffffff,
                   (int(t[j, g, 3]) - int(t[j, g, 1])) & 0xffffffff, (int(t[j, g, 2]) - int(t[j, g, 3])) & 0xffffffff) for g in range(2)])
    for name, base in (("group 0", off), ("group 1", off + 4096)):
        t = raw[base: base + 13 * 8].view(13, 8).to(torch.int64) & 0xffffffff
        print(name, "absolute [wait start, s_full seen, P handed over]:", [((int(t[j, 0]) - t0) & 0xffffffff, (int(t[j, 1]) - t0) & 0xffffffff, (int(t[j, 5]) - t0) & 0xffffffff) for j in range(13)])
    sys.exit(0)
for name, base in (("warp2", off), ("warp6 (half 1, same scheduler)", off + 4096)):
    t = raw[base: base + 13 * 8].view(13, 8).to(torch.int64) & 0xffffffff
    print(name, "per tile: [ld_wait+probes, max+s_empty, any/rescale, residual barrier waits, turn wait, exp+P+hand-over, fence+arrive] total, gap to next tile start")
    for j in range(13):
        d = [(int(t[j, i + 1]) - int(t[j, i])) & 0xffffffff for i in range(7)]
        tot = (int(t[j, 7]) - int(t[j, 0])) & 0xffffffff
        gap = ((int(t[j + 1, 0]) - int(t[j, 7])) & 0xffffffff) if j < 12 else -1
        print(j, d, tot, gap)

