// Micro-benchmark: how much of the softmax's exp2 work can move from the MUFU to the FMA pipe, and in what instruction
// arrangement.  One thread = one row of 128 scores; per pair of scores the kernel's exp2 loop does scale (FFMA2), two
// exponentials, row sum (FADD2), conversion (F2FP).  Variants: MUFU for every pair; a degree-3 polynomial on the FMA
// pipe for POLY of every 8 pairs, either as one block per pair (ARR 0) or with the polynomial's steps spread between
// the MUFU pairs of its group (ARR 1), packed (f32x2) or scalar arithmetic.  1 and 2 warps per SM sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o exp_poly exp_poly.cu && ./exp_poly
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

#define V asm volatile
__device__ __forceinline__ uint64_t pack_f2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack_f2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t r; V("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) { uint64_t r; V("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float ex2(float x) { float y; V("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t cvt2(float lo, float hi) { uint32_t r; V("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ float fmax_v(float a, float b) { float r; V("max.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float fadd_v(float a, float b) { float r; V("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float ffma_v(float a, float b, float c) { float r; V("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ float scale_exp(float r, float t) {   // r * 2^n with n in the low bits of t: one integer multiply-add
    int o; V("mad.lo.s32 %0, %1, 8388608, %2;" : "=r"(o) : "r"(__float_as_int(t)), "r"(__float_as_int(r))); return __int_as_float(o);
}
constexpr float kMagic = 12582912.f, kC3 = 0.055171459913253784f, kC2 = 0.2426108568906784f, kC1 = 0.6932609677314758f, kC0 = 0.9999281167984009f;

// one polynomial pair, packed arithmetic, as resumable steps
struct PolyP {
    uint64_t x, t, f, r;
    __device__ __forceinline__ void step(int s) {
        if (s == 0) { float a, b; unpack_f2(x, a, b); x = pack_f2(fmax_v(a, -125.f), fmax_v(b, -125.f)); }
        if (s == 1) t = fadd2(x, pack_f2(kMagic, kMagic));
        if (s == 2) f = fadd2(t, pack_f2(-kMagic, -kMagic));
        if (s == 3) f = fadd2(x, f ^ 0x8000000080000000ull);
        if (s == 4) r = ffma2(pack_f2(kC3, kC3), f, pack_f2(kC2, kC2));
        if (s == 5) r = ffma2(r, f, pack_f2(kC1, kC1));
        if (s == 6) r = ffma2(r, f, pack_f2(kC0, kC0));
        if (s == 7) { float t0, t1, r0, r1; unpack_f2(t, t0, t1); unpack_f2(r, r0, r1); r = pack_f2(scale_exp(r0, t0), scale_exp(r1, t1)); }
    }
};
// the same with scalar arithmetic (two elements side by side)
struct PolyS {
    float x0, x1, t0, t1, f0, f1, r0, r1;
    __device__ __forceinline__ void step(int s) {
        if (s == 0) { x0 = fmax_v(x0, -125.f); x1 = fmax_v(x1, -125.f); }
        if (s == 1) { t0 = fadd_v(x0, kMagic); t1 = fadd_v(x1, kMagic); }
        if (s == 2) { f0 = fadd_v(t0, -kMagic); f1 = fadd_v(t1, -kMagic); }
        if (s == 3) { f0 = fadd_v(x0, -f0); f1 = fadd_v(x1, -f1); }
        if (s == 4) { r0 = ffma_v(kC3, f0, kC2); r1 = ffma_v(kC3, f1, kC2); }
        if (s == 5) { r0 = ffma_v(r0, f0, kC1); r1 = ffma_v(r1, f1, kC1); }
        if (s == 6) { r0 = ffma_v(r0, f0, kC0); r1 = ffma_v(r1, f1, kC0); }
        if (s == 7) { r0 = scale_exp(r0, t0); r1 = scale_exp(r1, t1); }
    }
};

// POLY of every 8 pairs by polynomial; ARR 0: whole polynomial in place; ARR 1: steps spread between the MUFU pairs; PACKED
template <int POLY, int ARR, bool PACKED>
__global__ void __launch_bounds__(256, 1) k(uint32_t* out, long long* cycles, int iters) {
    float s[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) s[i] = -0.013f * (float)((threadIdx.x * 7 + i * 13) & 255) - 1e-4f * i;
    uint64_t l0 = 0, l1 = 0;
    uint32_t acc = 0;
    const uint64_t sc = pack_f2(1.0009f, 1.0009f);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const float shift = -1e-3f * (float)(acc & 1);       // loop-carried: nothing hoists out of the loop
        const uint64_t sh = pack_f2(shift, shift);
        uint32_t pk[64];
        auto consume = [&](int i, uint64_t e) {
            float p0, p1; unpack_f2(e, p0, p1);
            if (i & 1) l1 = fadd2(l1, e); else l0 = fadd2(l0, e);
            pk[i] = cvt2(p0, p1);
        };
#pragma unroll
        for (int g = 0; g < 8; ++g) {            // groups of 8 pairs: the first 8 - POLY by MUFU, the last POLY by polynomial
            uint64_t xs[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) xs[i] = ffma2(pack_f2(s[16 * g + 2 * i], s[16 * g + 2 * i + 1]), sc, sh);
            PolyP pp[POLY > 0 ? POLY : 1];
            PolyS ps[POLY > 0 ? POLY : 1];
#pragma unroll
            for (int j = 0; j < POLY; ++j) { pp[j].x = xs[8 - POLY + j]; unpack_f2(xs[8 - POLY + j], ps[j].x0, ps[j].x1); }
            constexpr int kMufu = 8 - POLY;
            constexpr int kSteps = 8 * POLY;                  // polynomial steps of the group (round robin over its pairs)
            int done_steps = 0;
#pragma unroll
            for (int i = 0; i < kMufu; ++i) {
                float x0, x1; unpack_f2(xs[i], x0, x1);
                const uint64_t e = pack_f2(ex2(x0), ex2(x1));
                if (ARR == 1) {                               // this MUFU pair's share of the polynomial steps
                    const int upto = kSteps * (i + 1) / kMufu;
#pragma unroll
                    for (; done_steps < upto; ++done_steps) {
                        if (PACKED) pp[done_steps % POLY].step(done_steps / POLY); else ps[done_steps % POLY].step(done_steps / POLY);
                    }
                }
                consume(8 * g + i, e);
            }
            if (ARR == 0) {
#pragma unroll
                for (int j = 0; j < POLY; ++j)
#pragma unroll
                    for (int st = 0; st < 8; ++st) { if (PACKED) pp[j].step(st); else ps[j].step(st); }
            }
#pragma unroll
            for (int j = 0; j < POLY; ++j) consume(8 * g + kMufu + j, PACKED ? pp[j].r : pack_f2(ps[j].r0, ps[j].r1));
        }
#pragma unroll
        for (int i = 0; i < 64; ++i) acc ^= pk[i];
    }
    const long long t1 = clock64();
    float a, b; unpack_f2(fadd2(l0, l1), a, b);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc ^ __float_as_uint(a + b);
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int POLY, int ARR, bool PACKED>
void run(uint32_t* out, long long* cyc) {
    const int iters = 2000;
    for (int warps = 4; warps <= 8; warps += 4) {
        k<POLY, ARR, PACKED><<<1, warps * 32>>>(out, cyc, iters);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("poly %d/8 %-10s %-6s %d warp(s)/sub-partition: %6.1f cycles per row tile per warp = %5.2f per pair; per sub-partition %5.2f cycles per pair\n",
               POLY, ARR ? "spread" : "in place", PACKED ? "packed" : "scalar", warps / 4, (double)c / iters, (double)c / iters / 64,
               (double)c / iters / 64 / (warps / 4));
    }
}

int main() {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
    run<0, 0, true>(out, cyc);
    run<1, 0, true>(out, cyc); run<1, 1, true>(out, cyc); run<1, 0, false>(out, cyc); run<1, 1, false>(out, cyc);
    run<2, 0, true>(out, cyc); run<2, 1, true>(out, cyc); run<2, 0, false>(out, cyc); run<2, 1, false>(out, cyc);
    run<3, 0, true>(out, cyc); run<3, 1, true>(out, cyc); run<3, 0, false>(out, cyc); run<3, 1, false>(out, cyc);
    run<4, 1, true>(out, cyc); run<4, 1, false>(out, cyc);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
