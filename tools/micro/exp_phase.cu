// Developer micro-benchmark (GPU box): the exp + pack + store phase of the attention softmax warps in isolation.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/micro/exp_phase.cu -o tools/micro/exp_phase && tools/micro/exp_phase
// Per thread and round: 64 scores in registers -> p = exp2(s * log2e - m) -> row sum -> fp16 pairs -> 8 x st.shared.v4 (swizzled).
// MODE 0: scalar FFMA / FADD (what attention.cu issues), MODE 1: fma.rn.f32x2 / add.f32x2 for the scale-subtract and the sums.
// Prints clk per 64-score round per warp for 1 / 2 / 4 warps per SM sub-partition (MUFU floor: 64 x 8 = 512 clk per warp-round,
// shared by the warps of a sub-partition).  Nothing here is a bench value.
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int MODE>
__global__ void k(const float* __restrict__ in, float* out, int rounds, long long* cyc) {
    extern __shared__ __align__(16) uint8_t sm[];
    float s[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) s[i] = in[(threadIdx.x * 64 + i) & 4095];
    const int row = threadIdx.x & 127;
    uint8_t* blk = sm + (threadIdx.x >> 7) * 16384 + row * 128;
    float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f, m = 3.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int r = 0; r < rounds; ++r) {
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            float p0, p1;
            if (MODE == 0) {
                p0 = ex2(fmaf(s[2 * i], 1.4426950408889634f, -m));
                p1 = ex2(fmaf(s[2 * i + 1], 1.4426950408889634f, -m));
                if (i & 1) { l2 += p0; l3 += p1; } else { l0 += p0; l1 += p1; }
            } else {
                float x0 = s[2 * i], x1 = s[2 * i + 1];
                const float c = 1.4426950408889634f, nm = -m;
                asm("{\n\t.reg .b64 a, b, d;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %2};\n\tmov.b64 d, {%3, %3};\n\t"
                    "fma.rn.f32x2 a, a, b, d;\n\tmov.b64 {%0, %1}, a;\n\t}" : "+f"(x0), "+f"(x1) : "f"(c), "f"(nm));
                p0 = ex2(x0); p1 = ex2(x1);
                float& la = (i & 1) ? l2 : l0;
                float& lb = (i & 1) ? l3 : l1;
                asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tadd.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
                    : "+f"(la), "+f"(lb) : "f"(p0), "f"(p1));
            }
            const __half2 h = __floats2half2_rn(p0, p1);
            pk[i] = *reinterpret_cast<const uint32_t*>(&h);
        }
#pragma unroll
        for (int t = 0; t < 8; ++t)
            *reinterpret_cast<uint4*>(blk + ((t ^ (row & 7)) << 4)) = make_uint4(pk[4 * t], pk[4 * t + 1], pk[4 * t + 2], pk[4 * t + 3]);
        m += 1e-7f * l0;          // keeps the rounds dependent on each other without changing the values much
#pragma unroll
        for (int i = 0; i < 64; i += 16) s[i] += 1e-7f;
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = l0 + l1 + l2 + l3 + m;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(int threads) {
    float *in, *out; long long* cyc;
    cudaMalloc(&in, 4096 * 4); cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    cudaMemset(in, 0, 4096 * 4);
    const int rounds = 2000;
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    k<MODE><<<148, threads, 65536>>>(in, out, 10, cyc);
    k<MODE><<<148, threads, 65536>>>(in, out, rounds, cyc);
    long long h;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return; }
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("mode %d, %d warps per sub-partition: %7.1f clk per 64-score round of a warp (%.1f exp / clk / SM)\n", MODE, threads / 128,
           (double)h / rounds, 64.0 * threads * rounds / (double)h);
    cudaFree(in); cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int th : {128, 256, 512}) { run<0>(th); run<1>(th); }
    return 0;
}
