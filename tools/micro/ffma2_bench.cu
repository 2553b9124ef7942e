// FFMA vs FFMA2 (fma.rn.f32x2) issue rate on sm_100a, and LDS.128 wavefront rate: per-SM throughput in thread-ops per clock.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench ffma2_bench.cu && ./ffma2_bench
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, long long* clk, int iters) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    const float w0 = 1.0001f, w1 = 0.9999f, m0 = 0.5f, m1 = 0.25f;
    __shared__ float4 sm[1024];
    if (MODE == 2) { for (int i = threadIdx.x; i < 1024; i += 256) sm[i] = make_float4(i, 1, 2, 3); __syncthreads(); }
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], w0, m0);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 16; i += 2)
                asm volatile("{\n\t.reg .b64 ra, rw, rm;\n\tmov.b64 ra, {%0, %1};\n\tmov.b64 rw, {%2, %3};\n\tmov.b64 rm, {%4, %5};\n\t"
                             "fma.rn.f32x2 ra, ra, rw, rm;\n\tmov.b64 {%0, %1}, ra;\n\t}"
                             : "+f"(a[i]), "+f"(a[i + 1]) : "f"(w0), "f"(w1), "f"(m0), "f"(m1));
        } else {
#pragma unroll
            for (int i = 0; i < 16; i += 4) {
                const float4 v = sm[(threadIdx.x + 16 * i + it) & 1023];
                a[i] += v.x; a[i + 1] += v.y; a[i + 2] += v.z; a[i + 3] += v.w;
            }
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[blockIdx.x * 256 + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

int main() {
    float* out; long long* clk;
    cudaMalloc(&out, 148 * 4 * 256 * 4); cudaMalloc(&clk, 148 * 4 * 8);
    const int iters = 20000;
    for (int ctas = 1; ctas <= 4; ctas *= 2) {
        for (int mode = 0; mode < 3; ++mode) {
            if (mode == 0) k<0><<<148 * ctas, 256>>>(out, clk, iters);
            if (mode == 1) k<1><<<148 * ctas, 256>>>(out, clk, iters);
            if (mode == 2) k<2><<<148 * ctas, 256>>>(out, clk, iters);
            cudaDeviceSynchronize();
            long long h[4 * 148];
            cudaMemcpy(h, clk, sizeof(long long) * 148 * ctas, cudaMemcpyDeviceToHost);
            double avg = 0; for (int i = 0; i < 148 * ctas; ++i) avg += h[i]; avg /= 148 * ctas;
            // per SM: ctas * 256 threads; ops per thread per iteration
            const double ops = mode == 0 ? 16 : mode == 1 ? 16 : 4;    // FMAs (mode 0, 1) or LDS.128 (mode 2) per thread-iteration
            printf("%s  %d CTA/SM (%d warps): %.1f clk/iter -> %.1f %s per clk per SM\n",
                   mode == 0 ? "FFMA   " : mode == 1 ? "FFMA2  " : "LDS.128", ctas, ctas * 8, avg / iters,
                   ops * ctas * 256 / (avg / iters), mode == 2 ? "LDS.128 thread-ops (x16 B)" : "fp32 FMAs");
        }
    }
    return 0;
}
