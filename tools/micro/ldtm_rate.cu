// Micro-benchmark: TMEM -> register bandwidth of tcgen05.ld (32x32b shape) on one SM: how long a warpgroup needs to
// pull a 128 x 128 fp32 score tile out of tensor memory, as a function of the repeat count per instruction and of
// how many warps share an SM sub-partition's TMEM port.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ldtm_rate ldtm_rate.cu && ./ldtm_rate
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

#define LD_REGS32(r, o) "=r"(r[o+0]), "=r"(r[o+1]), "=r"(r[o+2]), "=r"(r[o+3]), "=r"(r[o+4]), "=r"(r[o+5]), "=r"(r[o+6]), "=r"(r[o+7]), \
    "=r"(r[o+8]), "=r"(r[o+9]), "=r"(r[o+10]), "=r"(r[o+11]), "=r"(r[o+12]), "=r"(r[o+13]), "=r"(r[o+14]), "=r"(r[o+15]), \
    "=r"(r[o+16]), "=r"(r[o+17]), "=r"(r[o+18]), "=r"(r[o+19]), "=r"(r[o+20]), "=r"(r[o+21]), "=r"(r[o+22]), "=r"(r[o+23]), \
    "=r"(r[o+24]), "=r"(r[o+25]), "=r"(r[o+26]), "=r"(r[o+27]), "=r"(r[o+28]), "=r"(r[o+29]), "=r"(r[o+30]), "=r"(r[o+31])

__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : LD_REGS32(r, 0) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void ld64(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : LD_REGS32(r, 0), LD_REGS32(r, 32) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// MODE 0: 4 x (x32) then wait; 1: 2 x (x64) then wait; 2: 8 x (x16) then wait; 3: (x32, wait) x 4
template <int MODE>
__global__ void k(uint32_t* out, long long* cycles, int iters) {
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_ptr)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = tmem_ptr + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;
    uint32_t r[128];
    uint32_t acc = 0;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) { ld32(base, r); ld32(base + 32, r + 32); ld32(base + 64, r + 64); ld32(base + 96, r + 96); wait_ld(); }
        if (MODE == 1) { ld64(base, r); ld64(base + 64, r + 64); wait_ld(); }
        if (MODE == 2) {
#pragma unroll
            for (int c = 0; c < 8; ++c) ld16(base + 16 * c, r + 16 * c);
            wait_ld();
        }
        if (MODE == 3) {
#pragma unroll
            for (int c = 0; c < 4; ++c) { ld32(base + 32 * c, r + 32 * c); wait_ld(); }
        }
#pragma unroll
        for (int i = 0; i < 128; i += 16) acc ^= r[i];
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_ptr) : "memory");
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
    const int iters = 1000;
    const int warp_counts[] = {1, 4, 8, 16};
    for (int warps : warp_counts) {
        for (int mode = 0; mode < 4; ++mode) {
            if (mode == 0) k<0><<<1, warps * 32>>>(out, cyc, iters);
            if (mode == 1) k<1><<<1, warps * 32>>>(out, cyc, iters);
            if (mode == 2) k<2><<<1, warps * 32>>>(out, cyc, iters);
            if (mode == 3) k<3><<<1, warps * 32>>>(out, cyc, iters);
            cudaDeviceSynchronize();
            long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            const double bytes = (double)iters * warps * 32 * 128 * 4;
            printf("warps %2d mode %d: %.0f cycles per 128-column row load per warp, %.1f B/clk/SM\n", warps, mode,
                   (double)c / iters, bytes / c);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
