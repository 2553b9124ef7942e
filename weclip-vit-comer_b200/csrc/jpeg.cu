// Input pipeline: baseline JPEG -> ImageNet-normalised CHW fp32 on the device, and label-map -> label-set.
//
// Replaces, for the evaluation / training loaders, imageio.imread + normalize_img + HWC->CHW
// (datasets/voc.py:48-67,253-271, datasets/transforms.py:8-15) and np.unique on the label PNG (clip/clip_tool.py:111-118).
// The reference decodes through PIL = libjpeg(-turbo) with its defaults (JDCT_ISLOW, fancy up-sampling, the fixed-point
// YCbCr->RGB tables).  libjpeg is not under /root/reference; its published algorithm is restated here so that the decoded
// pixels are BIT-IDENTICAL to PIL's (tests/test_io_gpu.py checks that against PIL on synthetic JPEGs):
//   host   wc_jpeg_parse / wc_jpeg_entropy_decode: markers, quantisation + Huffman tables, the sequential entropy decode
//          (inherently serial bit-stream work: it stays on a host thread; ctypes releases the GIL, loaders run it in parallel)
//   device jpeg_idct_kernel : de-quantise + 8x8 inverse DCT, the "islow" integer algorithm (jidctint.c: CONST_BITS 13,
//                             PASS1_BITS 2, the 12 FIX constants), one thread per block, level shift + clamp to uint8
//          jpeg_color_kernel: chroma up-sampling (h2v2 / h2v1 "fancy" triangle filters with their edge rules, jdsample.c;
//                             1x1 copy), YCbCr->RGB with the 16.16 fixed-point tables of jdcolor.c, then
//                             (u8 - mean) / std evaluated in fp64 and rounded to fp32 exactly as numpy does, CHW planar
// Supported: baseline / extended sequential 8-bit Huffman JPEG (SOF0 / SOF1), 1 or 3 components, luma sampling 1x1, 2x1,
// 2x2 with 1x1 chroma, restart intervals.  Progressive / arithmetic / 12-bit / CMYK streams return WC_ERR_UNSUPPORTED
// (the caller raises; there is no silent fallback).
#include "common.cuh"

#include <stdlib.h>
#include <string.h>

namespace wc {

static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct HuffTable {
    bool present = false;
    uint8_t bits[17];
    uint8_t vals[256];
    // canonical decode tables
    int mincode[17], maxcode[18], valptr[17];
    uint8_t look_nbits[256], look_sym[256];     // 8-bit lookahead
};

static void build_huff(HuffTable& h) {
    int code = 0, k = 0;
    memset(h.look_nbits, 0, sizeof(h.look_nbits));
    for (int l = 1; l <= 16; ++l) {
        h.valptr[l] = k;
        h.mincode[l] = code;
        for (int i = 0; i < h.bits[l]; ++i, ++k, ++code) {
            if (l <= 8) {
                const int first = code << (8 - l), n = 1 << (8 - l);
                for (int j = 0; j < n; ++j) { h.look_nbits[first + j] = (uint8_t)l; h.look_sym[first + j] = h.vals[k]; }
            }
        }
        h.maxcode[l] = h.bits[l] ? code - 1 : -1;
        code <<= 1;
    }
    h.maxcode[17] = 0x7fffffff;
}

struct BitReader {
    const uint8_t* p; const uint8_t* end;
    uint32_t buf = 0; int cnt = 0; bool hit_marker = false;
    void fill() {
        while (cnt <= 24) {
            int b = 0;
            if (!hit_marker && p < end) {
                b = *p;
                if (b == 0xFF) {
                    if (p + 1 < end && p[1] == 0x00) p += 2;            // stuffed zero
                    else { hit_marker = true; b = 0; }                  // a marker: feed zeros, leave p at it
                } else ++p;
            }
            buf |= (uint32_t)b << (24 - cnt);
            cnt += 8;
        }
    }
    inline int peek(int n) { if (cnt < n) fill(); return (int)(buf >> (32 - n)); }
    inline void skip(int n) { buf <<= n; cnt -= n; }
    inline int get(int n) { if (n == 0) return 0; const int v = peek(n); skip(n); return v; }
    void reset() { buf = 0; cnt = 0; hit_marker = false; }
};

static inline int huff_decode(BitReader& br, const HuffTable& h) {
    const int look = br.peek(8);
    int nb = h.look_nbits[look];
    if (nb) { br.skip(nb); return h.look_sym[look]; }
    int code = br.peek(16);
    for (int l = 9; l <= 16; ++l) {
        const int c = code >> (16 - l);
        if (h.maxcode[l] >= 0 && c <= h.maxcode[l] && c >= h.mincode[l]) { br.skip(l); return h.vals[h.valptr[l] + c - h.mincode[l]]; }
    }
    return -1;
}
static inline int extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

struct JpegState {
    wc_jpeg_info info;
    HuffTable dc[4], ac[4];
    int comp_id[3], comp_tq[3], comp_td[3], comp_ta[3];
    const uint8_t* scan = nullptr; const uint8_t* end = nullptr;
};

static int parse(const uint8_t* d, size_t n, JpegState& st) {
    memset(&st.info, 0, sizeof(st.info));
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) return fail(WC_ERR_ARG, "%s: not a JPEG stream", __func__);
    size_t p = 2;
    bool have_sof = false;
    while (p + 4 <= n) {
        if (d[p] != 0xFF) { ++p; continue; }
        const int m = d[p + 1];
        if (m == 0xFF) { ++p; continue; }
        p += 2;
        if (m == 0xD8 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;
        if (m == 0xD9) break;
        if (p + 2 > n) break;
        const size_t len = ((size_t)d[p] << 8) | d[p + 1];
        if (len < 2 || p + len > n) return fail(WC_ERR_ARG, "%s: truncated segment", __func__);
        const uint8_t* s = d + p + 2; const size_t sl = len - 2;
        if (m == 0xDB) {                                   // DQT
            size_t q = 0;
            while (q < sl) {
                const int pq = s[q] >> 4, tq = s[q] & 15; ++q;
                if (tq > 3 || q + (pq ? 128 : 64) > sl) return fail(WC_ERR_ARG, "%s: bad DQT", __func__);
                for (int i = 0; i < 64; ++i) {
                    const int v = pq ? ((s[q] << 8) | s[q + 1]) : s[q];
                    q += pq ? 2 : 1;
                    st.info.qt[tq][kZigzag[i]] = (uint16_t)v;          // stored in natural order
                }
            }
        } else if (m == 0xC0 || m == 0xC1) {               // SOF0 / SOF1
            if (sl < 6 || s[0] != 8) return fail(WC_ERR_UNSUPPORTED, "%s: only 8-bit samples", __func__);
            st.info.height = (s[1] << 8) | s[2]; st.info.width = (s[3] << 8) | s[4];
            st.info.ncomp = s[5];
            if ((st.info.ncomp != 1 && st.info.ncomp != 3) || sl < 6 + 3 * (size_t)st.info.ncomp)
                return fail(WC_ERR_UNSUPPORTED, "%s: %ld components", "", st.info.ncomp);
            for (int c = 0; c < st.info.ncomp; ++c) {
                st.comp_id[c] = s[6 + 3 * c];
                st.info.hs[c] = s[7 + 3 * c] >> 4; st.info.vs[c] = s[7 + 3 * c] & 15;
                st.comp_tq[c] = s[8 + 3 * c];
                if (st.comp_tq[c] > 3) return fail(WC_ERR_ARG, "%s: bad Tq", __func__);
            }
            have_sof = true;
        } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
            return fail(WC_ERR_UNSUPPORTED, "%s: progressive / lossless / arithmetic JPEG (SOF marker 0x%lx)", "", m);
        } else if (m == 0xC4) {                            // DHT
            size_t q = 0;
            while (q + 17 <= sl) {
                const int tc = s[q] >> 4, th = s[q] & 15; ++q;
                if (tc > 1 || th > 3) return fail(WC_ERR_ARG, "%s: bad DHT", __func__);
                HuffTable& h = tc ? st.ac[th] : st.dc[th];
                int cnt = 0;
                h.bits[0] = 0;
                for (int i = 1; i <= 16; ++i) { h.bits[i] = s[q++]; cnt += h.bits[i]; }
                if (cnt > 256 || q + cnt > sl) return fail(WC_ERR_ARG, "%s: bad DHT", __func__);
                memcpy(h.vals, s + q, cnt); q += cnt;
                h.present = true;
                build_huff(h);
            }
        } else if (m == 0xDD) {                            // DRI
            if (sl >= 2) st.info.restart_interval = (s[0] << 8) | s[1];
        } else if (m == 0xDA) {                            // SOS
            if (!have_sof) return fail(WC_ERR_ARG, "%s: SOS before SOF", __func__);
            const int ns = s[0];
            if (ns != st.info.ncomp || sl < 1 + 2 * (size_t)ns + 3) return fail(WC_ERR_UNSUPPORTED, "%s: non-interleaved scan", __func__);
            for (int i = 0; i < ns; ++i) {
                int c = -1;
                for (int j = 0; j < st.info.ncomp; ++j) if (st.comp_id[j] == s[1 + 2 * i]) c = j;
                if (c != i) return fail(WC_ERR_UNSUPPORTED, "%s: scan component order", __func__);
                st.comp_td[c] = s[2 + 2 * i] >> 4; st.comp_ta[c] = s[2 + 2 * i] & 15;
                if (st.comp_td[c] > 3 || st.comp_ta[c] > 3) return fail(WC_ERR_ARG, "%s: bad table selector", __func__);
            }
            st.scan = d + p + len; st.end = d + n;
            break;
        }
        p += len;
    }
    if (!have_sof || !st.scan) return fail(WC_ERR_ARG, "%s: no frame / scan found", __func__);
    wc_jpeg_info& I = st.info;
    if (I.width <= 0 || I.height <= 0) return fail(WC_ERR_ARG, "%s: empty image", __func__);
    I.hmax = 1; I.vmax = 1;
    for (int c = 0; c < I.ncomp; ++c) { I.hmax = I.hs[c] > I.hmax ? I.hs[c] : I.hmax; I.vmax = I.vs[c] > I.vmax ? I.vs[c] : I.vmax; }
    if (I.ncomp == 1) { I.hs[0] = I.vs[0] = I.hmax = I.vmax = 1; }      // a single component is never interleaved
    if (I.ncomp == 3) {
        const bool ok = I.hs[1] == 1 && I.vs[1] == 1 && I.hs[2] == 1 && I.vs[2] == 1 &&
                        ((I.hs[0] == 1 && I.vs[0] == 1) || (I.hs[0] == 2 && I.vs[0] == 1) || (I.hs[0] == 2 && I.vs[0] == 2));
        if (!ok) return fail(WC_ERR_UNSUPPORTED, "%s: chroma sampling other than 4:4:4 / 4:2:2 / 4:2:0", __func__);
    }
    I.mcus_x = (I.width + 8 * I.hmax - 1) / (8 * I.hmax);
    I.mcus_y = (I.height + 8 * I.vmax - 1) / (8 * I.vmax);
    long off = 0;
    for (int c = 0; c < I.ncomp; ++c) {
        I.blocks_w[c] = I.mcus_x * I.hs[c]; I.blocks_h[c] = I.mcus_y * I.vs[c];
        I.coef_off[c] = off;
        off += (long)I.blocks_w[c] * I.blocks_h[c] * 64;
        I.comp_qt[c] = st.comp_tq[c];
    }
    I.coef_count = off;
    return WC_OK;
}

static int entropy_decode(JpegState& st, int16_t* coef) {
    const wc_jpeg_info& I = st.info;
    for (int c = 0; c < I.ncomp; ++c)
        if (!st.dc[st.comp_td[c]].present || !st.ac[st.comp_ta[c]].present) return fail(WC_ERR_ARG, "%s: missing Huffman table", __func__);
    memset(coef, 0, sizeof(int16_t) * I.coef_count);
    BitReader br{st.scan, st.end};
    int pred[3] = {0, 0, 0};
    int restart_left = I.restart_interval;
    for (int my = 0; my < I.mcus_y; ++my)
        for (int mx = 0; mx < I.mcus_x; ++mx) {
            if (I.restart_interval && restart_left == 0) {
                // byte-align, expect RSTn
                const uint8_t* q = br.p;
                while (q + 1 < st.end && !(q[0] == 0xFF && q[1] >= 0xD0 && q[1] <= 0xD7)) ++q;
                if (q + 1 >= st.end) return fail(WC_ERR_ARG, "%s: missing restart marker", __func__);
                br.p = q + 2; br.reset();
                pred[0] = pred[1] = pred[2] = 0;
                restart_left = I.restart_interval;
            }
            for (int c = 0; c < I.ncomp; ++c) {
                const HuffTable& hd = st.dc[st.comp_td[c]];
                const HuffTable& ha = st.ac[st.comp_ta[c]];
                for (int by = 0; by < I.vs[c]; ++by)
                    for (int bx = 0; bx < I.hs[c]; ++bx) {
                        int16_t* blk = coef + I.coef_off[c] + ((long)(my * I.vs[c] + by) * I.blocks_w[c] + (mx * I.hs[c] + bx)) * 64;
                        int s = huff_decode(br, hd);
                        if (s < 0 || s > 11) return fail(WC_ERR_ARG, "%s: corrupt DC code", __func__);
                        int diff = s ? extend(br.get(s), s) : 0;
                        pred[c] += diff;
                        blk[0] = (int16_t)pred[c];
                        for (int k = 1; k < 64;) {
                            const int rs = huff_decode(br, ha);
                            if (rs < 0) return fail(WC_ERR_ARG, "%s: corrupt AC code", __func__);
                            const int r = rs >> 4, sz = rs & 15;
                            if (sz == 0) {
                                if (r == 15) { k += 16; continue; }
                                break;                              // EOB
                            }
                            k += r;
                            if (k > 63) return fail(WC_ERR_ARG, "%s: corrupt AC run", __func__);
                            blk[kZigzag[k]] = (int16_t)extend(br.get(sz), sz);
                            ++k;
                        }
                    }
            }
            if (I.restart_interval) --restart_left;
        }
    return WC_OK;
}

// ---------------------------------------------------------------------------------------------------
// Device side
// ---------------------------------------------------------------------------------------------------
struct JpegDev {
    int width, height, ncomp, hs0, vs0;
    int blocks_w[3], blocks_h[3];
    long coef_off[3];
    long plane_off[3];          // uint8 component planes (pitch = blocks_w * 8)
    uint16_t qt[3][64];
};

#define JFIX_0_298631336 2446
#define JFIX_0_390180644 3196
#define JFIX_0_541196100 4433
#define JFIX_0_765366865 6270
#define JFIX_0_899976223 7373
#define JFIX_1_175875602 9633
#define JFIX_1_501321110 12299
#define JFIX_1_847759065 15137
#define JFIX_1_961570560 16069
#define JFIX_2_053119869 16819
#define JFIX_2_562915447 20995
#define JFIX_3_072711026 25172

__device__ __forceinline__ int jdescale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// jidctint.c (jpeg_idct_islow) on 8 values with stride `st`; SHIFT: CONST_BITS - PASS1_BITS = 11 for the column pass,
// CONST_BITS + PASS1_BITS + 3 = 18 for the row pass.
template <int SHIFT>
__device__ __forceinline__ void idct8(const int* in, int st, int* out, int ost) {
    int z2 = in[2 * st], z3 = in[6 * st];
    int z1 = (z2 + z3) * JFIX_0_541196100;
    int tmp2 = z1 + z3 * (-JFIX_1_847759065);
    int tmp3 = z1 + z2 * JFIX_0_765366865;
    z2 = in[0]; z3 = in[4 * st];
    int tmp0 = (z2 + z3) * (1 << 13), tmp1 = (z2 - z3) * (1 << 13);
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = in[7 * st]; tmp1 = in[5 * st]; tmp2 = in[3 * st]; tmp3 = in[1 * st];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
    int z4 = tmp1 + tmp3;
    const int z5 = (z3 + z4) * JFIX_1_175875602;
    tmp0 *= JFIX_0_298631336; tmp1 *= JFIX_2_053119869; tmp2 *= JFIX_3_072711026; tmp3 *= JFIX_1_501321110;
    z1 *= -JFIX_0_899976223; z2 *= -JFIX_2_562915447; z3 *= -JFIX_1_961570560; z4 *= -JFIX_0_390180644;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    out[0 * ost] = jdescale(tmp10 + tmp3, SHIFT); out[7 * ost] = jdescale(tmp10 - tmp3, SHIFT);
    out[1 * ost] = jdescale(tmp11 + tmp2, SHIFT); out[6 * ost] = jdescale(tmp11 - tmp2, SHIFT);
    out[2 * ost] = jdescale(tmp12 + tmp1, SHIFT); out[5 * ost] = jdescale(tmp12 - tmp1, SHIFT);
    out[3 * ost] = jdescale(tmp13 + tmp0, SHIFT); out[4 * ost] = jdescale(tmp13 - tmp0, SHIFT);
}

__global__ void __launch_bounds__(64) jpeg_idct_kernel(const int16_t* __restrict__ coef, uint8_t* __restrict__ planes,
                                                       const __grid_constant__ JpegDev J, int comp) {
    const long b = (long)blockIdx.x * 64 + threadIdx.x;
    const long nb = (long)J.blocks_w[comp] * J.blocks_h[comp];
    if (b >= nb) return;
    const int16_t* c = coef + J.coef_off[comp] + b * 64;
    int blk[64], ws[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) blk[i] = (int)c[i] * (int)J.qt[comp][i];
#pragma unroll
    for (int col = 0; col < 8; ++col) idct8<11>(blk + col, 8, ws + col, 8);
    const int by = (int)(b / J.blocks_w[comp]), bx = (int)(b % J.blocks_w[comp]);
    const int pitch = J.blocks_w[comp] * 8;
    uint8_t* o = planes + J.plane_off[comp] + (size_t)(by * 8) * pitch + bx * 8;
#pragma unroll
    for (int row = 0; row < 8; ++row) {
        int r[8];
        idct8<18>(ws + row * 8, 1, r, 1);
        uint2 pk = make_uint2(0, 0);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const unsigned v = (unsigned)min(max(r[i] + 128, 0), 255);
            if (i < 4) pk.x |= v << (8 * i); else pk.y |= v << (8 * (i - 4));
        }
        *reinterpret_cast<uint2*>(o + (size_t)row * pitch) = pk;
    }
}

// jdsample.c: value of the up-sampled chroma plane at full-resolution (x, y); cw / chh = the component's real
// (down-sampled) width / height, which is where libjpeg's edge rules apply (not the padded block extent).
__device__ __forceinline__ int chroma_at(const uint8_t* pl, int pitch, int cw, int chh, int x, int y, int hs0, int vs0) {
    if (hs0 == 1) return pl[(size_t)y * pitch + x];                                  // 4:4:4
    const int cx = x >> 1;
    if (vs0 == 1) {                                                                   // h2v1 fancy
        const uint8_t* r = pl + (size_t)y * pitch;
        const int v = r[cx];
        if (cw == 1) return v;
        if (x & 1) return cx == cw - 1 ? v : (3 * v + r[cx + 1] + 2) >> 2;
        return cx == 0 ? v : (3 * v + r[cx - 1] + 1) >> 2;
    }
    // h2v2 fancy: rows weighted 3:1 (nearer : farther, with the first / last real row replicated), then columns 3:1
    const int cy = y >> 1;
    const int cy2 = (y & 1) ? min(cy + 1, chh - 1) : max(cy - 1, 0);
    const uint8_t* r0 = pl + (size_t)cy * pitch;
    const uint8_t* r1 = pl + (size_t)cy2 * pitch;
    const int cs = 3 * r0[cx] + r1[cx];
    if (cw == 1) return (cs * 4 + 8) >> 4;
    if (x & 1) {
        if (cx == cw - 1) return (cs * 4 + 7) >> 4;
        return (cs * 3 + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
    }
    if (cx == 0) return (cs * 4 + 8) >> 4;
    return (cs * 3 + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
}

__global__ void __launch_bounds__(256) jpeg_color_kernel(const uint8_t* __restrict__ planes, uint8_t* __restrict__ rgb_hwc,
                                                         float* __restrict__ chw, const __grid_constant__ JpegDev J,
                                                         double m0, double m1, double m2, double s0, double s1, double s2) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= J.width || y >= J.height) return;
    const int yv = planes[J.plane_off[0] + (size_t)y * (J.blocks_w[0] * 8) + x];
    int r, g, b;
    if (J.ncomp == 1) { r = g = b = yv; }
    else {
        const int cw = (J.width * 1 + J.hs0 - 1) / J.hs0, chh = (J.height + J.vs0 - 1) / J.vs0;
        const int cb = chroma_at(planes + J.plane_off[1], J.blocks_w[1] * 8, cw, chh, x, y, J.hs0, J.vs0) - 128;
        const int cr = chroma_at(planes + J.plane_off[2], J.blocks_w[2] * 8, cw, chh, x, y, J.hs0, J.vs0) - 128;
        // jdcolor.c build_ycc_rgb_table: SCALEBITS 16, ONE_HALF = 1 << 15, FIX(x) = (int)(x * 65536 + 0.5)
        const int cr_r = (91881 * cr + 32768) >> 16, cb_b = (116130 * cb + 32768) >> 16;
        const int cr_g = -46802 * cr, cb_g = -22554 * cb + 32768;
        r = min(max(yv + cr_r, 0), 255);
        g = min(max(yv + ((cb_g + cr_g) >> 16), 0), 255);
        b = min(max(yv + cb_b, 0), 255);
    }
    const size_t px = (size_t)y * J.width + x;
    if (rgb_hwc) { rgb_hwc[3 * px] = (uint8_t)r; rgb_hwc[3 * px + 1] = (uint8_t)g; rgb_hwc[3 * px + 2] = (uint8_t)b; }
    if (chw) {
        const size_t hw = (size_t)J.width * J.height;
        // datasets/transforms.py:8-15: (uint8 - mean) / std in float64, stored as float32
        chw[px] = (float)(((double)r - m0) / s0);
        chw[hw + px] = (float)(((double)g - m1) / s1);
        chw[2 * hw + px] = (float)(((double)b - m2) / s2);
    }
}

// uint8 HWC (already decoded) -> normalised fp32 CHW (datasets/transforms.py:8-15 + np.transpose, voc.py:236)
__global__ void __launch_bounds__(256) normalize_u8_kernel(const uint8_t* __restrict__ hwc, float* __restrict__ chw, long hw,
                                                           double m0, double m1, double m2, double s0, double s1, double s2) {
    const long px = (long)blockIdx.x * 256 + threadIdx.x;
    if (px >= hw) return;
    chw[px] = (float)(((double)hwc[3 * px] - m0) / s0);
    chw[hw + px] = (float)(((double)hwc[3 * px + 1] - m1) / s1);
    chw[2 * hw + px] = (float)(((double)hwc[3 * px + 2] - m2) / s2);
}

// present[v] = 1 if value v occurs in the label map (np.unique of clip_tool.py:113 as a 256-entry presence table)
__global__ void __launch_bounds__(256) label_presence_kernel(const uint8_t* __restrict__ lab, long n, int* __restrict__ present) {
    __shared__ int sp[256];
    sp[threadIdx.x] = 0;
    __syncthreads();
    for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long)gridDim.x * 256) sp[lab[i]] = 1;
    __syncthreads();
    if (sp[threadIdx.x]) present[threadIdx.x] = 1;
}

}  // namespace wc

using namespace wc;

extern "C" int wc_jpeg_parse(const uint8_t* data_host, size_t n, wc_jpeg_info* info_host) {
    WC_CHECK_ARG(data_host && info_host && n > 0);
    JpegState st;
    const int rc = parse(data_host, n, st);
    if (rc) return rc;
    *info_host = st.info;
    return WC_OK;
}

extern "C" int wc_jpeg_entropy_decode(const uint8_t* data_host, size_t n, int16_t* coef_host, long coef_cap) {
    WC_CHECK_ARG(data_host && coef_host && n > 0);
    JpegState st;
    int rc = parse(data_host, n, st);
    if (rc) return rc;
    if (coef_cap < st.info.coef_count) return fail(WC_ERR_WORKSPACE, "%s: coefficient buffer too small (%ld)", __func__, coef_cap);
    return entropy_decode(st, coef_host);
}

extern "C" size_t wc_jpeg_workspace_bytes(const wc_jpeg_info* info_host) {
    if (!info_host) return 0;
    size_t b = 0;
    for (int c = 0; c < info_host->ncomp; ++c) b += (size_t)info_host->blocks_w[c] * info_host->blocks_h[c] * 64;
    return b + 256;
}

extern "C" int wc_jpeg_reconstruct(const wc_jpeg_info* info_host, const int16_t* coef, uint8_t* rgb_hwc, float* chw,
                                   const double* mean_host, const double* std_host, void* workspace, size_t workspace_bytes,
                                   void* stream) {
    WC_CHECK_ARG(info_host && coef && (rgb_hwc || chw) && workspace && (!chw || (mean_host && std_host)));
    if (workspace_bytes < wc_jpeg_workspace_bytes(info_host)) return fail(WC_ERR_WORKSPACE, "%s: workspace too small", __func__);
    const wc_jpeg_info& I = *info_host;
    JpegDev J;
    J.width = I.width; J.height = I.height; J.ncomp = I.ncomp; J.hs0 = I.hs[0]; J.vs0 = I.vs[0];
    long off = 0;
    for (int c = 0; c < 3; ++c) {
        J.blocks_w[c] = c < I.ncomp ? I.blocks_w[c] : 0; J.blocks_h[c] = c < I.ncomp ? I.blocks_h[c] : 0;
        J.coef_off[c] = c < I.ncomp ? I.coef_off[c] : 0;
        J.plane_off[c] = off;
        off += (long)J.blocks_w[c] * J.blocks_h[c] * 64;
        for (int i = 0; i < 64; ++i) J.qt[c][i] = c < I.ncomp ? I.qt[I.comp_qt[c]][i] : 0;
    }
    cudaStream_t st = as_stream(stream);
    uint8_t* planes = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    for (int c = 0; c < I.ncomp; ++c) {
        const long nb = (long)J.blocks_w[c] * J.blocks_h[c];
        jpeg_idct_kernel<<<(unsigned)ceil_div(nb, 64L), 64, 0, st>>>(coef, planes, J, c);
        WC_LAUNCH_CHECK();
    }
    const double one[3] = {1, 1, 1}, zero[3] = {0, 0, 0};
    const double* m = mean_host ? mean_host : zero;
    const double* s = std_host ? std_host : one;
    jpeg_color_kernel<<<dim3(ceil_div(I.width, 32), ceil_div(I.height, 8)), 256, 0, st>>>(planes, rgb_hwc, chw, J, m[0], m[1], m[2],
                                                                                          s[0], s[1], s[2]);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_normalize_u8(const uint8_t* hwc, float* chw, int H, int W, const double* mean_host, const double* std_host,
                               void* stream) {
    WC_CHECK_ARG(hwc && chw && H > 0 && W > 0 && mean_host && std_host);
    const long hw = (long)H * W;
    normalize_u8_kernel<<<(unsigned)ceil_div(hw, 256L), 256, 0, as_stream(stream)>>>(hwc, chw, hw, mean_host[0], mean_host[1], mean_host[2],
                                                                                    std_host[0], std_host[1], std_host[2]);
    WC_LAUNCH_CHECK();
    return WC_OK;
}

extern "C" int wc_label_presence(const uint8_t* label, long n, int* present256, void* stream) {
    WC_CHECK_ARG(label && present256 && n > 0);
    WC_CUDA(cudaMemsetAsync(present256, 0, 256 * sizeof(int), as_stream(stream)));
    const unsigned grid = (unsigned)(ceil_div(n, 256L * 16) < 592 ? ceil_div(n, 256L * 16) : 592);
    label_presence_kernel<<<grid ? grid : 1, 256, 0, as_stream(stream)>>>(label, n, present256);
    WC_LAUNCH_CHECK();
    return WC_OK;
}
