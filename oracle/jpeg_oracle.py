"""ORACLE — CPU restatement of libjpeg's baseline reconstruction (the arithmetic behind imageio.imread / PIL in
datasets/voc.py:50).  TEST INFRASTRUCTURE ONLY.

libjpeg(-turbo) is a third-party dependency that is not under /root/reference (PIL's bundled copy; the reference pins only
``imageio`` / ``Pillow`` through requirements.txt).  Its published algorithm with the library defaults is restated here in
numpy: jidctint.c (``jpeg_idct_islow``: CONST_BITS 13, PASS1_BITS 2), jdsample.c (``h2v1_fancy_upsample``,
``h2v2_fancy_upsample`` and their edge rules), jdcolor.c (16.16 fixed-point YCbCr->RGB tables).  Parity status: PINNED —
tests/test_io_cpu.py checks this restatement (fed with the coefficients of the library's own host entropy decoder) against
the pixels PIL itself returns, bit for bit, on synthetic JPEGs of every supported sampling mode.
"""
import numpy as np

C = dict(F0_298631336=2446, F0_390180644=3196, F0_541196100=4433, F0_765366865=6270, F0_899976223=7373, F1_175875602=9633,
         F1_501321110=12299, F1_847759065=15137, F1_961570560=16069, F2_053119869=16819, F2_562915447=20995, F3_072711026=25172)


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def _idct8(v, shift):
    """jidctint.c, one 1-D pass over the LAST axis of v (int64 [..., 8])."""
    i = [v[..., k] for k in range(8)]
    z2, z3 = i[2], i[6]
    z1 = (z2 + z3) * C["F0_541196100"]
    tmp2 = z1 - z3 * C["F1_847759065"]
    tmp3 = z1 + z2 * C["F0_765366865"]
    tmp0 = (i[0] + i[4]) << 13
    tmp1 = (i[0] - i[4]) << 13
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = i[7], i[5], i[3], i[1]
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * C["F1_175875602"]
    t0, t1, t2, t3 = t0 * C["F0_298631336"], t1 * C["F2_053119869"], t2 * C["F3_072711026"], t3 * C["F1_501321110"]
    z1, z2, z3, z4 = -z1 * C["F0_899976223"], -z2 * C["F2_562915447"], -z3 * C["F1_961570560"] + z5, -z4 * C["F0_390180644"] + z5
    t0, t1, t2, t3 = t0 + z1 + z3, t1 + z2 + z4, t2 + z2 + z3, t3 + z1 + z4
    out = [tmp10 + t3, tmp11 + t2, tmp12 + t1, tmp13 + t0, tmp13 - t0, tmp12 - t1, tmp11 - t2, tmp10 - t3]
    return np.stack([_descale(o, shift) for o in out], axis=-1)


def idct_plane(coef, qt, bw, bh):
    """coef int16 [bh*bw, 64] natural order, qt [64] -> uint8 plane [bh*8, bw*8]."""
    blk = coef.astype(np.int64).reshape(-1, 8, 8) * qt.astype(np.int64).reshape(1, 8, 8)
    ws = _idct8(blk.transpose(0, 2, 1), 11).transpose(0, 2, 1)          # columns
    px = np.clip(_idct8(ws, 18) + 128, 0, 255).astype(np.uint8)         # rows
    return px.reshape(bh, bw, 8, 8).transpose(0, 2, 1, 3).reshape(bh * 8, bw * 8)


def upsample_h2v1(p):
    """jdsample.c h2v1_fancy_upsample on the real (down-sampled) extent p [h, cw] -> [h, 2 cw]."""
    v = p.astype(np.int64)
    prev = np.concatenate([v[:, :1], v[:, :-1]], 1)
    nxt = np.concatenate([v[:, 1:], v[:, -1:]], 1)
    even, odd = (3 * v + prev + 1) >> 2, (3 * v + nxt + 2) >> 2
    even[:, 0], odd[:, -1] = v[:, 0], v[:, -1]
    out = np.empty((p.shape[0], 2 * p.shape[1]), np.int64)
    out[:, 0::2], out[:, 1::2] = even, odd
    return out


def upsample_h2v2(p):
    """jdsample.c h2v2_fancy_upsample on the real extent p [ch, cw] -> [2 ch, 2 cw]."""
    v = p.astype(np.int64)
    up = np.concatenate([v[:1], v[:-1]], 0)
    dn = np.concatenate([v[1:], v[-1:]], 0)
    rows = np.empty((2 * v.shape[0], v.shape[1]), np.int64)
    rows[0::2], rows[1::2] = 3 * v + up, 3 * v + dn                     # column sums (scaled by 4)
    prev = np.concatenate([rows[:, :1], rows[:, :-1]], 1)
    nxt = np.concatenate([rows[:, 1:], rows[:, -1:]], 1)
    even, odd = (3 * rows + prev + 8) >> 4, (3 * rows + nxt + 7) >> 4
    even[:, 0], odd[:, -1] = (rows[:, 0] * 4 + 8) >> 4, (rows[:, -1] * 4 + 7) >> 4
    out = np.empty((rows.shape[0], 2 * rows.shape[1]), np.int64)
    out[:, 0::2], out[:, 1::2] = even, odd
    return out


def reconstruct(info, coef):
    """info: the fields of wc_jpeg_info as a dict-like / ctypes struct; coef int16 [coef_count] -> uint8 RGB [H,W,3]."""
    H, W, nc = info.height, info.width, info.ncomp
    planes = []
    for c in range(nc):
        bw, bh = info.blocks_w[c], info.blocks_h[c]
        cc = np.asarray(coef[info.coef_off[c]: info.coef_off[c] + bw * bh * 64]).reshape(bw * bh, 64)
        planes.append(idct_plane(cc, np.array(list(info.qt[info.comp_qt[c]])), bw, bh))
    y = planes[0][:H, :W].astype(np.int64)
    if nc == 1:
        return np.repeat(y[..., None], 3, 2).astype(np.uint8)
    hs, vs = info.hs[0], info.vs[0]
    cw, ch = (W + hs - 1) // hs, (H + vs - 1) // vs
    chroma = []
    for p in planes[1:]:
        p = p[:ch, :cw]
        p = p.astype(np.int64) if hs == 1 else upsample_h2v1(p) if vs == 1 else upsample_h2v2(p)
        chroma.append(p[:H, :W] - 128)
    cb, cr = chroma
    r = y + ((91881 * cr + 32768) >> 16)
    g = y + ((-22554 * cb + 32768 - 46802 * cr) >> 16)
    b = y + ((116130 * cb + 32768) >> 16)
    return np.clip(np.stack([r, g, b], -1), 0, 255).astype(np.uint8)
