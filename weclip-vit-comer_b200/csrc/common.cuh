// Shared helpers for libweclip_sm100.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>

#include "../../include/weclip_b200.h"

namespace wc {

// 16-bit GEMM operand type.  fp16, not bf16: the reference's own GPU weights are fp16 (clip/model.py:461-476), so real
// checkpoints load bit-exactly, and the 11-bit significand keeps activations 4x closer to the fp32 oracle at the same
// tensor-core rate.  Accumulation is always fp32 (TMEM); gradients are power-of-two scaled (gradcam.cu).
using h16 = __half;
using h16x2 = __half2;

extern thread_local char g_err[512];

inline int fail(int code, const char* fmt, const char* a = "", long b = 0, long c = 0) {
    snprintf(g_err, sizeof(g_err), fmt, a, b, c);
    return code;
}

#define WC_CHECK_ARG(cond)                                                                     \
    do {                                                                                       \
        if (!(cond)) return ::wc::fail(WC_ERR_ARG, "%s: bad argument: " #cond, __func__);      \
    } while (0)

#define WC_CUDA(expr)                                                                          \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess)                                                                \
            return ::wc::fail(WC_ERR_CUDA, "%s: CUDA error %ld", cudaGetErrorString(e__), (long)e__); \
    } while (0)

#define WC_LAUNCH_CHECK() WC_CUDA(cudaGetLastError())

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device (per-context) attribute: set it once per call site AND
// device, so that one process driving several GPUs gets it on each of them (bit d of the mask = device d done).
#define WC_SET_MAX_SMEM(func, bytes)                                                                       \
    do {                                                                                                   \
        static std::atomic<unsigned long long> done__{0ull};                                               \
        int dev__ = 0;                                                                                     \
        WC_CUDA(cudaGetDevice(&dev__));                                                                    \
        const unsigned long long bit__ = 1ull << (dev__ & 63);                                             \
        if (!(done__.load(std::memory_order_acquire) & bit__)) {                                           \
            WC_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            done__.fetch_or(bit__, std::memory_order_release);                                             \
        }                                                                                                  \
    } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

template <typename T>
__host__ __device__ constexpr T ceil_div(T a, T b) { return (a + b - 1) / b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide reductions for blockDim.x <= 1024 (scratch: 32 floats of shared memory).
__device__ __forceinline__ float block_sum(float v, float* scratch) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? scratch[lane] : 0.f;
    return warp_sum(r);
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? scratch[lane] : -INFINITY;
    return warp_max(r);
}
__device__ __forceinline__ float block_min(float v, float* scratch) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_min(v);
    __syncthreads();
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? scratch[lane] : INFINITY;
    return warp_min(r);
}

// ---------------------------------------------------------------------------------------------------
// Opt-in per-kernel-class device timing (wc_prof_*): CUDA events recorded on the launching stream around each
// launch, summed per class after a synchronise.  Off by default; bench.py turns it on for a separate pass so the
// headline timed region is never perturbed.
// ---------------------------------------------------------------------------------------------------
enum ProfKind : int {
    PK_GEMM = 0, PK_LAYERNORM, PK_SOFTMAX, PK_PAR_AFF, PK_PAR_ITER, PK_AFFINITY, PK_GRADCAM_MISC, PK_ENC_MISC,
    PK_CAM_UPSAMPLE, PK_EVAL, PK_ATTENTION, PK_COUNT
};
extern bool g_prof_on;
void prof_begin(int kind, double work, cudaStream_t st);
void prof_tag(int a, int b, int c, int d);   // annotate the open record (GEMM: M, N, K, batch)
void prof_end(cudaStream_t st);
struct ProfScope {
    cudaStream_t st; bool on;
    ProfScope(int kind, double work, cudaStream_t s) : st(s), on(g_prof_on) { if (on) prof_begin(kind, work, st); }
    ~ProfScope() { if (on) prof_end(st); }
};

#define WC_PROF_LAUNCH(kind, work, st, ...)            \
    do {                                               \
        ::wc::ProfScope ps__(kind, work, st);          \
        __VA_ARGS__;                                   \
    } while (0)

}  // namespace wc
