// Micro-benchmark of the softmax exp2 loop of the prefill kernel: one thread = one row of 128 scores in registers,
// per pair of scores: scale (FFMA2), two MUFU.EX2, row sum (FADD2), conversion (F2FP).  Question: how many cycles
// per pair does ONE warp per SM sub-partition reach (the kernel measures 24 against 16 of MUFU time), how many with
// TWO warps per sub-partition, and does the distance between a pair's MUFUs and their consumers matter?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o exp_loop exp_loop.cu && ./exp_loop
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

__device__ __forceinline__ uint64_t pack_f2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack_f2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) { uint64_t r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t cvt2(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }

// MODE 0: MUFU only (128 independent ex2 per iteration).  MODE 1: the kernel's loop, consumers DELAY pairs behind.
// MODE 2: blocks of 32 pairs: all MUFUs of a block, then all its consumers.
template <int MODE, int DELAY>
__global__ void __launch_bounds__(256, 1) k(uint32_t* out, long long* cycles, int iters) {
    float s[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) s[i] = -0.01f * ((threadIdx.x + i) & 63);
    uint64_t l0 = 0, l1 = 0;
    uint32_t acc = 0;
    const uint64_t sc = pack_f2(1.0009f, 1.0009f);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const float shift = -1e-3f * (float)(acc & 1);       // loop-carried: nothing hoists out of the loop
        const uint64_t sh = pack_f2(shift, shift);
        uint32_t pk[64];
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 128; ++i) s[i] = ex2(s[i] + shift) - 1.0f;
#pragma unroll
            for (int i = 0; i < 64; ++i) pk[i] = __float_as_uint(s[2 * i]) ^ __float_as_uint(s[2 * i + 1]);
        } else if (MODE == 1) {
            uint64_t done[DELAY > 0 ? DELAY : 1];
#pragma unroll
            for (int i = 0; i < 64 + DELAY; ++i) {
                if (i < 64) {
                    float x0, x1;
                    unpack_f2(ffma2(pack_f2(s[2 * i], s[2 * i + 1]), sc, sh), x0, x1);
                    const uint64_t e = pack_f2(ex2(x0), ex2(x1));
                    if (DELAY == 0) {
                        float p0, p1; unpack_f2(e, p0, p1);
                        if (i & 1) l1 = fadd2(l1, e); else l0 = fadd2(l0, e);
                        pk[i] = cvt2(p0, p1);
                    } else {
                        if (i >= DELAY) {
                            const uint64_t d = done[i % DELAY];
                            float p0, p1; unpack_f2(d, p0, p1);
                            if ((i - DELAY) & 1) l1 = fadd2(l1, d); else l0 = fadd2(l0, d);
                            pk[i - DELAY] = cvt2(p0, p1);
                        }
                        done[i % DELAY] = e;
                    }
                } else if (DELAY > 0) {
                    const uint64_t d = done[i % DELAY];
                    float p0, p1; unpack_f2(d, p0, p1);
                    if ((i - DELAY) & 1) l1 = fadd2(l1, d); else l0 = fadd2(l0, d);
                    pk[i - DELAY] = cvt2(p0, p1);
                }
            }
        } else {
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                uint64_t e[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float x0, x1;
                    unpack_f2(ffma2(pack_f2(s[64 * b + 2 * i], s[64 * b + 2 * i + 1]), sc, sh), x0, x1);
                    e[i] = pack_f2(ex2(x0), ex2(x1));
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float p0, p1; unpack_f2(e[i], p0, p1);
                    if (i & 1) l1 = fadd2(l1, e[i]); else l0 = fadd2(l0, e[i]);
                    pk[32 * b + i] = cvt2(p0, p1);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 64; ++i) acc ^= pk[i];
    }
    const long long t1 = clock64();
    float a, b; unpack_f2(fadd2(l0, l1), a, b);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc ^ __float_as_uint(a + b);
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE, int DELAY>
void run(const char* name, uint32_t* out, long long* cyc) {
    const int iters = 2000;
    for (int warps = 4; warps <= 8; warps += 4) {
        k<MODE, DELAY><<<1, warps * 32>>>(out, cyc, iters);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("%-34s %d warp(s) per sub-partition: %6.1f cycles per 128-score row tile = %5.2f cycles per pair per warp\n", name,
               warps / 4, (double)c / iters, (double)c / iters / 64);
    }
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
    run<0, 0>("MUFU only", out, cyc);
    run<1, 0>("kernel loop, consumers delay 0", out, cyc);
    run<1, 1>("kernel loop, consumers delay 1", out, cyc);
    run<1, 2>("kernel loop, consumers delay 2", out, cyc);
    run<1, 4>("kernel loop, consumers delay 4", out, cyc);
    run<1, 8>("kernel loop, consumers delay 8", out, cyc);
    run<2, 0>("32 MUFU pairs, then 32 consumers", out, cyc);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
