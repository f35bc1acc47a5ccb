// Micro-benchmark: MUFU throughput of ex2.approx for f32, f16x2 and bf16x2 operands on one SM sub-partition.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

template <int MODE>
__global__ void k(uint32_t* out, long long* cycles, int iters) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = 0x3c003c00u ^ (threadIdx.x * 2654435761u >> 20) ^ i;   // small finite values
    float f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = -0.001f * (threadIdx.x + i);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
            if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(a[i]));
            if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(a[i]));
        }
    }
    long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= a[i] ^ __float_as_uint(f[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
    const int iters = 2000;
    for (int warps = 4; warps <= 16; warps *= 2) {
        for (int mode = 0; mode < 3; ++mode) {
            if (mode == 0) k<0><<<1, warps * 32>>>(out, cyc, iters);
            if (mode == 1) k<1><<<1, warps * 32>>>(out, cyc, iters);
            if (mode == 2) k<2><<<1, warps * 32>>>(out, cyc, iters);
            cudaDeviceSynchronize();
            long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            const double ops = (double)iters * 8 * warps;       // warp-instructions on one SM
            printf("warps %2d mode %s: %.2f cycles per warp-instruction per SM (%.1f results/clk/SM)\n", warps,
                   mode == 0 ? "f32   " : mode == 1 ? "f16x2 " : "bf16x2", c / ops, (mode ? 64.0 : 32.0) * ops / c);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
