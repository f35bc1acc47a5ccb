// Merge of split-KV partials by one thread block (the reference's flash_combine_kernel,
// src/flash_attn.cu:877-935, in fp32 and with the split axis spread over the block so that a row with a few
// hundred splits is not a serial chain of loads).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace sfa {

constexpr int kMergeGroups = 8;   // blockDim.x = D * kMergeGroups; thread = (group g, head_dim element d)

template <int D>
struct MergeScratch {
    float red[32];                       // cross-warp reductions
    float acc[kMergeGroups][D];
};

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// o_part (rows, S, D) un-normalised partial outputs, ml_part (rows, S) = (m in log2 units, l).  On return every
// thread holds M and L of the row; threads of group 0 hold acc = sum_s 2^(m_s - M) o_s[d].  w_sm: S floats.
template <int D>
__device__ __forceinline__ void merge_splits(const float* __restrict__ o_part, const float2* __restrict__ ml_part,
                                             int64_t row0, int S, float* w_sm, MergeScratch<D>& sc, float& acc,
                                             float& L, float& M) {
    constexpr int T = D * kMergeGroups, W = T / 32;
    constexpr int G = kMergeGroups;
    constexpr int kPre = 8;               // partial rows per thread whose loads go out before the weights are known
    const int tid = threadIdx.x, g = tid / D, d = tid - g * D, warp = tid >> 5, lane = tid & 31;
    // One global round trip instead of three: every (m, l) is read once (S <= T: at most one per thread, more only in
    // the fallback loop), and the first kPre * G rows of o_part are requested before M and the weights exist.
    const float* op = o_part + row0 * D + d;
    const float2 ml0 = tid < S ? ml_part[row0 + tid] : make_float2(-INFINITY, 0.f);
    float v[kPre];
#pragma unroll
    for (int i = 0; i < kPre; ++i) v[i] = (g + i * G < S) ? op[(int64_t)(g + i * G) * D] : 0.f;
    float m = ml0.x;
    for (int s = tid + T; s < S; s += T) m = fmaxf(m, ml_part[row0 + s].x);
    m = warp_max(m);
    if (lane == 0) sc.red[warp] = m;
    __syncthreads();
    m = lane < W ? sc.red[lane] : -INFINITY;
    M = warp_max(m);
    __syncthreads();
    float l = 0.f;
    if (tid < S) {
        const float w = (ml0.x == -INFINITY) ? 0.f : exp2f(ml0.x - M);
        w_sm[tid] = w;
        l = w * ml0.y;
    }
    for (int s = tid + T; s < S; s += T) {
        const float2 ml = ml_part[row0 + s];
        const float w = (ml.x == -INFINITY) ? 0.f : exp2f(ml.x - M);
        w_sm[s] = w;
        l += w * ml.y;
    }
    l = warp_sum(l);
    if (lane == 0) sc.red[warp] = l;
    __syncthreads();                      // also publishes w_sm
    l = lane < W ? sc.red[lane] : 0.f;
    L = warp_sum(l);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int i = 0; i < kPre; i += 2) {
        if (g + i * G < S) a0 += w_sm[g + i * G] * v[i];
        if (g + (i + 1) * G < S) a1 += w_sm[g + (i + 1) * G] * v[i + 1];
    }
    // the rest, four independent loads in flight per thread
    int s = g + kPre * G;
    for (; s + 3 * G < S; s += 4 * G) {
        const float v0 = op[(int64_t)s * D], v1 = op[(int64_t)(s + G) * D];
        const float v2 = op[(int64_t)(s + 2 * G) * D], v3 = op[(int64_t)(s + 3 * G) * D];
        a0 += w_sm[s] * v0; a1 += w_sm[s + G] * v1; a2 += w_sm[s + 2 * G] * v2; a3 += w_sm[s + 3 * G] * v3;
    }
    for (; s < S; s += G) a0 += w_sm[s] * op[(int64_t)s * D];
    a0 += a2; a1 += a3;
    sc.acc[g][d] = a0 + a1;
    __syncthreads();
    acc = 0.f;
    if (g == 0) {
#pragma unroll
        for (int i = 0; i < kMergeGroups; ++i) acc += sc.acc[i][d];
    }
}

}  // namespace sfa
