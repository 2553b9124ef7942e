// Internal: transformer block runner shared by the encoder, the decoder and the Grad-CAM block-12 forward.
#pragma once
#include "gemm.cuh"
#include "elementwise.cuh"
#include "attention.cuh"

namespace wc {


inline long round8(long v) { return (v + 7) & ~7L; }

// Bump allocator over a caller-provided workspace (256-byte aligned carves).
struct Arena {
    uint8_t* base; size_t cap; size_t off = 0; bool ok = true;
    Arena(void* p, size_t n) : base(reinterpret_cast<uint8_t*>(p)), cap(n) {
        off = (256 - (reinterpret_cast<uintptr_t>(p) & 255)) & 255;
    }
    template <typename T> T* get(size_t n) {
        size_t bytes = (n * sizeof(T) + 255) & ~(size_t)255;
        if (base == nullptr) { off += bytes; return nullptr; }         // sizing pass
        if (off + bytes > cap) { ok = false; return nullptr; }
        T* r = reinterpret_cast<T*>(base + off);
        off += bytes;
        return r;
    }
};

struct BlockBuffers {
    // residual stream, fp32 [M,E]: x_in -> (attention) x_mid -> (MLP) x_out; pointers may alias
    const float* x_in; float* x_mid; float* x_out;
    h16* a;            // [M,E]   LayerNorm output (GEMM operand)
    float* a_f32;       // optional fp32 copy of the ln_1 output (Grad-CAM hook tensor)
    h16 *q, *k, *vt;   // [N*heads, L, hd], [N*heads, L, hd], [N*heads, hd, Lp]
    h16* v;            // optional [N*heads, L, hd]
    float* S;           // [N*heads, L, Lp]   materialised scores (unfused path only)
    h16* P;            // [N*heads, L, Lp]   materialised probabilities (unfused path only; Grad-CAM's backward reads them)
    int attn_imgs = 0;  // images (a prefix of the batch) whose head-mean map is wanted; 0 = all
    float* lse;         // [N*heads, L]       fused path (S == nullptr, head_dim 64): log2-domain row log-sum-exp
    h16* o;            // [M,E]
    h16* u;            // [M,4E]  QuickGELU(c_fc)
    h16* u_pre;        // optional [M,4E] pre-activation
    float* attn_avg;    // optional [N,L,L]
    double* attn_sum = nullptr;   // optional [attn_imgs]: sum of attn_avg[n][1:,1:] (fused path only)
    h16* tap;          // optional h16 copy of x_out
    h16* tap_in;       // optional h16 copy of x_in, written by ln_1 (the previous block's tap when its c_proj adds in place)
};

// clip/model.py:210-214 with clip/myAtt.py:199-326 inside.
int run_block(const wc_block_weights& w, const BlockBuffers& b, int n_img, int L, int E, int heads, cudaStream_t st);

}  // namespace wc
