// Does the MUFU take half-precision exp2 any faster than fp32?  ex2.approx.f16x2 compiles to two MUFU.EX2.F16 and a
// PRMT (cuobjdump), so the packed form saves nothing by itself; this measures cycles per MUFU warp instruction for
// f32, f16 (scalar and packed) and bf16x2, one and two warps per SM sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ex2_half ex2_half.cu && ./ex2_half
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

// MODE 0: 64 x ex2.approx.ftz.f32 per iteration; 1: 64 x ex2.approx.f16x2 (= 128 MUFU.EX2.F16); 2: 64 x ex2.approx.ftz.bf16x2;
// 3: 64 x ex2.approx.f16 (scalar)
template <int MODE>
__global__ void __launch_bounds__(256, 1) k(uint32_t* out, long long* cycles, int iters) {
    uint32_t r[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) r[i] = MODE == 0 ? __float_as_uint(-0.01f * ((threadIdx.x + i) & 63)) : 0xb800b400u + ((threadIdx.x + i) & 63) * 0x00010001u;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            if (MODE == 0) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(__uint_as_float(r[i]))); r[i] = __float_as_uint(y) | 0x80000000u; }
            else if (MODE == 1) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(r[i])); r[i] = y | 0x80008000u; }
            else if (MODE == 2) { uint32_t y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(r[i])); r[i] = y | 0x80008000u; }
            else { unsigned short y, x = (unsigned short)r[i]; asm volatile("ex2.approx.f16 %0, %1;" : "=h"(y) : "h"(x)); r[i] = y | 0x8000u; }
        }
    }
    const long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 64; ++i) acc ^= r[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int mufu_per_iter, uint32_t* out, long long* cyc) {
    const int iters = 4000;
    for (int warps = 4; warps <= 8; warps += 4) {
        k<MODE><<<1, warps * 32>>>(out, cyc, iters);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("%-28s %d warp(s) per sub-partition: %7.1f cycles per iteration, %5.2f cycles per MUFU warp instruction per sub-partition\n", name,
               warps / 4, (double)c / iters, (double)c / iters / mufu_per_iter / (warps / 4));
    }
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
    run<0>("ex2.approx.ftz.f32", 64, out, cyc);
    run<1>("ex2.approx.f16x2", 128, out, cyc);
    run<2>("ex2.approx.ftz.bf16x2", 128, out, cyc);
    run<3>("ex2.approx.f16", 64, out, cyc);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
