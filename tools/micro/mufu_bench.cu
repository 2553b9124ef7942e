// Micro-benchmark (GPU box): ex2.approx / rcp / tanh / FFMA throughput per SM per clock, and a polynomial exp2 on the
// FMA pipe.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/mufu_bench tools/micro/mufu_bench.cu && /tmp/mufu_bench
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void k(float* out, int iters, long long* cyc) {
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = -0.001f * (threadIdx.x + i + 1);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 1) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 2) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 3) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(x[i]));
            if (OP == 4) {   // packed half2 exp2
                unsigned u = __float_as_uint(x[i]);
                asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u));
                x[i] = __uint_as_float(u);
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int threads) {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 4096;
    k<OP><<<148, threads>>>(out, iters, cyc);
    k<OP><<<148, threads>>>(out, iters, cyc);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = 0; for (int i = 0; i < 148; ++i) c += h[i]; c /= 148;
    printf("%-16s threads/SM %4d: %.2f lane-ops/clk/SM\n", name, threads, (double)iters * 8 * threads / c);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int t : {128, 256, 512, 1024}) {
        run<0>("ex2.approx.f32", t); run<1>("rcp.approx.f32", t); run<2>("tanh.approx.f32", t);
        run<4>("ex2.f16x2(ops)", t); run<3>("fma.f32", t);
    }
    return 0;
}
