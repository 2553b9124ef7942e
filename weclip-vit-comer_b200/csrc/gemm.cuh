// tcgen05 / TMEM / TMA GEMM for sm_100a:  D[b][m][n] = epilogue( sum_k A[b][m][k] * B[b][n][k] ).
// A and B are h16, K-major (row-major [rows, K]); accumulation is fp32 in tensor memory.
// Internal C++ interface; the C ABI wrappers live in gemm.cu / encoder.cu.
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace wc {

enum GemmAct : int { ACT_NONE = 0, ACT_RELU = 1, ACT_QUICKGELU = 2, ACT_SIGMOID = 3 };
enum GemmMode : int {
    GM_STORE = 0,        // out = act(alpha*acc + bias) [+ resid]; optional h16 copy out2; optional transposed store
    GM_QKV = 1,          // in-proj scatter: q (scaled), k, v, vT head-major layouts
    GM_MUL_QGELU_GRAD = 2,   // out = acc * d/du quickgelu(u) with u = aux (backward of the MLP activation)
};

// One operand: a rank-4 tensor (k, row, b0, b1); strides in ELEMENTS (h16).  b = blockIdx.z -> (b % nb0, b / nb0).
struct GemmOperand {
    const h16* ptr;
    long rows;          // extent of the row dimension per batch entry (TMA zero-fills beyond it)
    long ld;            // row stride
    long bs0, bs1;      // batch strides (0 = broadcast)
};

struct GemmEpilogue {
    int mode = GM_STORE;
    int act = ACT_NONE;
    float alpha = 1.f;
    const float* bias = nullptr; long bias_bs0 = 0, bias_bs1 = 0;
    void* out = nullptr; int out_f16 = 0; int out_trans = 0; long ldo = 0, out_bs0 = 0, out_bs1 = 0;
    const float* resid = nullptr; long ldr = 0, resid_bs0 = 0, resid_bs1 = 0;
    h16* out2 = nullptr; long ld2 = 0, out2_bs0 = 0, out2_bs1 = 0; int out2_preact = 0;   // h16 copy (before act+resid if preact)
    const h16* aux = nullptr; long ldaux = 0; long aux_bs = 0; const int* aux_map = nullptr;
    // GM_QKV
    h16 *q = nullptr, *k = nullptr, *v = nullptr, *vt = nullptr;
    int L = 0, Lp = 0, heads = 0, hd = 0; float qscale = 1.f;
};

struct GemmProblem {
    int M, N, K;        // per batch entry
    int nb0 = 1, nb1 = 1;
    GemmOperand A, B;
    GemmEpilogue epi;
};

// Rank-4 K-major tensor map (k, row, b0, b1) over `op` with a (64 x box_rows) box and 128-byte swizzle (cached per
// thread).  use_b0 / use_b1 report whether the batch coordinates are live (0: pass coordinate 0).
int make_tma_map(CUtensorMap* out, const GemmOperand& op, int K, int nb0, int nb1, int box_rows, int* use_b0, int* use_b1);

// Enqueue on `st`.  Returns WC_OK or a negative error (message in wc_last_error()).
int launch_gemm(const GemmProblem& p, cudaStream_t st);

}  // namespace wc
