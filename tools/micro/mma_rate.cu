// Developer micro-benchmark (GPU box): issue rate of tcgen05.mma.cta_group::1.kind::f16 at the attention kernel's shapes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I weclip-vit-comer_b200/csrc \
//        tools/micro/mma_rate.cu -o tools/micro/mma_rate -lcuda && tools/micro/mma_rate
// One CTA per SM, one thread issues `rounds` batches back to back (operands are zeros: only the time matters), then commits and
// waits.  Modes: 0 = SS M128 N128 K16 x4 (scores), 1 = TS M128 N64 K16 x8, B MN-major (P V), 2 = SS M128 N64 K16 x8 B MN-major,
// 3 = TS M128 N64 with B K-major, 4 = mode 0 + mode 1 interleaved (one attention tile), 5 = SS M128 N256 K16 x4 (GEMM tile).
// Nothing here is a bench value.
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include "common.cuh"
#include "tc_ptx.cuh"

using namespace wc;

__device__ __forceinline__ void umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

// warps 4.. (when launched with more than 128 threads) stream tcgen05.ld over the S columns (and, ldst = 2, tcgen05.st into the P
// columns) until the issuer is done: what the softmax warps do to the tensor memory while the products run
// data = 1: operands are pseudo-random fp16 in (-1, 1) instead of zeros (real switching activity in the tensor pipe; the first product of
// every batch overwrites its accumulator as in the attention kernel, so the values stay finite)
__global__ void __launch_bounds__(384, 1) mma_rate_kernel(int mode, int rounds, int ldst, int commits, int data, long long* clk) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar, bar2[2];
    __shared__ uint32_t slot;
    __shared__ volatile int done_flag;
    if (threadIdx.x == 0) done_flag = 0;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) {
        uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        // two halves with exponent field <= 14 (|x| < 1), random sign and mantissa
        const uint32_t v = (h & 0x83ff83ffu) | ((((h >> 10) & 7u) + 7u) << 10) | ((((h >> 26) & 7u) + 7u) << 26);
        reinterpret_cast<uint32_t*>(smem)[i] = data ? v : 0u;
    }
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&bar2[0], 1); mbar_init(&bar2[1], 1); mbar_fence_init(); }
    if (threadIdx.x < 32) tmem_alloc(&slot, 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = slot;
    if (data && threadIdx.x < 128) {      // the TMEM A operand (P, columns 128..191) gets the same kind of values
        uint32_t r[32];
        for (int c = 0; c < 2; ++c) {
            for (int e = 0; e < 32; ++e) r[e] = reinterpret_cast<uint32_t*>(smem)[(threadIdx.x * 64 + c * 32 + e) & 8191];
            tmem_st32(tb + ((uint32_t)((threadIdx.x >> 5) * 32) << 16) + 128 + c * 32, r);
        }
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x == 0) {
        const uint64_t da = make_smem_desc(smem_u32(smem)), db = make_smem_desc(smem_u32(smem + 16384)),
                       dv = make_smem_desc(smem_u32(smem + 32768)), dw = make_smem_desc(smem_u32(smem + 49152));
        constexpr uint32_t id_s = make_idesc(128, 128), id_o = make_idesc(128, 64) | (1u << 16), id_ok = make_idesc(128, 64),
                           id_g = make_idesc(128, 256);
        const long long t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
            if (mode == 0 || mode == 4) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tb, da + 2 * k, db + 2 * k, id_s, k != 0);
                if (commits) umma_commit(&bar2[0]);          // (nobody waits: the phase just flips)
            }
            if (mode == 1 || mode == 4) {
#pragma unroll
                for (int k = 0; k < 8; ++k) umma_ts(tb + 256, tb + 128 + 8 * k, dv + 128 * k, id_o, k != 0);
                if (commits) umma_commit(&bar2[1]);
            }
            if (mode == 2) {
#pragma unroll
                for (int k = 0; k < 8; ++k) umma_f16(tb + 256, da + 2 * (k & 3), dv + 128 * k, id_o, k != 0);
            }
            if (mode == 3) {
#pragma unroll
                for (int k = 0; k < 8; ++k) umma_ts(tb + 256, tb + 128 + 8 * k, dv + 2 * (k & 3), id_ok, k != 0);
            }
            if (mode == 5) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tb, da + 2 * k, dv + 2 * k, id_g, k != 0);
                if (commits) umma_commit(&bar2[0]);
            }
            (void)dw;
        }
        const long long t1 = clock64();
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        const long long t2 = clock64();
        if (blockIdx.x == 0) { clk[0] = t1 - t0; clk[1] = t2 - t0; }
        done_flag = 1;
    } else if (threadIdx.x >= 128 && ldst) {
        const int warp = threadIdx.x >> 5;
        const uint32_t lanes = (uint32_t)((warp & 3) * 32) << 16;
        uint32_t r[32];
        long long n = 0;
        const long long t0 = clock64();
        while (!done_flag) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                tmem_ld32_issue(tb + lanes + c * 32, r);
                tmem_ld_wait();
                if (ldst == 2 && (c & 1)) tmem_st32(tb + lanes + 384 + (c >> 1) * 32, r);
            }
            if (ldst == 2) tmem_st_wait();
            ++n;
        }
        const long long t1 = clock64();
        if (blockIdx.x == 0 && threadIdx.x == 128) { clk[2] = n * 4 * 4096; clk[3] = t1 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

int main() {
    long long* clk;
    cudaMalloc(&clk, 32);
    cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    const char* names[] = {"SS 128x128x16 x4 (scores)", "TS 128x64x16 x8, B MN-major (P V)", "SS 128x64x16 x8, B MN-major",
                           "TS 128x64x16 x8, B K-major", "scores + P V (one attention tile)", "SS 128x256x16 x4 (GEMM k-block)"};
    for (int cfg = 0; cfg < 9; ++cfg) {
        const int data = cfg >= 7, c = data ? (cfg == 7 ? 1 : 6) : cfg;       // cfg 7 / 8 = cfg 1 / 6 with random operands
        const int grid = c == 0 ? 1 : 148, ldst = (c < 2 || c == 5) ? 0 : (c == 4 || c == 6 ? 2 : 1),
                  threads = (c < 2 || c == 5) ? 128 : (c == 2 ? 256 : 384), commits = c >= 5;
        printf("-- grid %d, %d warps streaming tcgen05.ld%s beside the issuer\n", grid, (threads - 128) / 32, ldst == 2 ? " + st" : "");
        if (commits) printf("   (tcgen05.commit after every batch of 4 / 8)\n");
        if (data) printf("   (pseudo-random fp16 operands instead of zeros)\n");
        for (int mode = 0; mode < 6; ++mode) {
            const int rounds = 2000;
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            mma_rate_kernel<<<grid, threads, 100 * 1024>>>(mode, 10, ldst, commits, data, clk);
            cudaEventRecord(e0);
            mma_rate_kernel<<<grid, threads, 100 * 1024>>>(mode, rounds, ldst, commits, data, clk);
            cudaEventRecord(e1);
            cudaError_t err = cudaDeviceSynchronize();
            if (err != cudaSuccess) { printf("mode %d: %s\n", mode, cudaGetErrorString(err)); return 1; }
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            long long h[4]; cudaMemcpy(h, clk, 32, cudaMemcpyDeviceToHost);
            printf("grid %3d  %-36s: issue %7.1f clk / batch, complete %7.1f clk / batch, %.3f ms (%.2f GHz)\n", grid, names[mode],
                   (double)h[0] / rounds, (double)h[1] / rounds, ms, h[1] / (ms * 1e6));
            if (ldst) printf("            one streaming warp: %.1f B/clk of tcgen05.ld\n", (double)h[2] / (double)h[3]);
        }
    }
    return 0;
}
