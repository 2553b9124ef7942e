// Internal launchers of the memory-bound helper kernels (elementwise.cu).
#pragma once
#include "common.cuh"

namespace wc {
int launch_layernorm(const float* x, const float* w, const float* b, h16* out_h, float* out_f, long rows,
                     int E, cudaStream_t st, h16* raw_h = nullptr);   // raw_h: optional fp16 copy of x itself
int launch_softmax_headmean(const float* S, h16* P, float* avg, int n_img, int heads, int L, int Lp,
                            cudaStream_t st);
int launch_im2col16(const float* img, h16* out, int N, int H, int W, cudaStream_t st);
int launch_posemb_resize(const float* pos, float* out, int gs, int h, int w, int E, cudaStream_t st);
int launch_cls_row(float* x, const float* cls, const float* pos, int N, int L, int E, cudaStream_t st);
int launch_cast_bf16(const float* x, h16* y, size_t n, cudaStream_t st);
}  // namespace wc
