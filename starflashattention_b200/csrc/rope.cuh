// Interleaved-pair rotary embedding shared by the stand-alone RoPE/append kernel (rope_cache.cu) and the fused
// decoder step (decode_sm100.cu).  Reference: rotary_embedding_transform / apply_rotary_embedding,
// src/flash_attn.cu:167-257: (x, y) -> (c x - s y, c y + s x) with angle pos / 10000^(2i / rot_dim).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace sfa {

template <typename T> struct Vec2;
template <> struct Vec2<__half> {
    using type = __half2;
    static __device__ __forceinline__ float2 to_f(type v) { return __half22float2(v); }
    static __device__ __forceinline__ type from_f(float a, float b) { return __floats2half2_rn(a, b); }
    static __device__ __forceinline__ float one_to_f(__half v) { return __half2float(v); }
    static __device__ __forceinline__ __half one_from_f(float v) { return __float2half_rn(v); }
};
template <> struct Vec2<__nv_bfloat16> {
    using type = __nv_bfloat162;
    static __device__ __forceinline__ float2 to_f(type v) { return __bfloat1622float2(v); }
    static __device__ __forceinline__ type from_f(float a, float b) { return __floats2bfloat162_rn(a, b); }
    static __device__ __forceinline__ float one_to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
    static __device__ __forceinline__ __nv_bfloat16 one_from_f(float v) { return __float2bfloat16_rn(v); }
};

// angle_i(pos) = pos * fl32(10000^(-2i/rot_dim)); the inverse frequency is rounded once from
// double so it is the correctly rounded fp32 value, the product is an fp32 multiply as in the
// Python example (examples/python/testFlashDecoder.py:10,20-25).
__device__ __forceinline__ float rope_angle(int pos, int pair, int rot_dim) {
    const float inv_freq = (float)(1.0 / pow(10000.0, (double)(2 * pair) / (double)rot_dim));
    return (float)pos * inv_freq;
}

// fp32 evaluation of (cos, sin) when no table is given.  Deliberately not inlined: the double-precision pow and
// the slow path of sincosf are several hundred instructions, and the fused decode kernel would otherwise carry
// dozens of copies of them in front of its streaming loop.
static __device__ __noinline__ float2 rope_cos_sin_eval(int pos, int pair, int rot_dim) {
    float s, c;
    sincosf(rope_angle(pos, pair, rot_dim), &s, &c);
    return make_float2(c, s);
}

// cos/sin of pair `pair` at position `pos`: from the (max_len, rot_dim/2) tables when given, else in fp32.
template <typename T>
__device__ __forceinline__ void rope_cos_sin(const void* cos_table, const void* sin_table, int pos, int pair,
                                             int rot_dim, float& c, float& s) {
    if (cos_table) {
        const int64_t ti = (int64_t)pos * (rot_dim / 2) + pair;
        c = Vec2<T>::one_to_f(reinterpret_cast<const T*>(cos_table)[ti]);
        s = Vec2<T>::one_to_f(reinterpret_cast<const T*>(sin_table)[ti]);
    } else {
        const float2 cs = rope_cos_sin_eval(pos, pair, rot_dim);
        c = cs.x; s = cs.y;
    }
}

template <typename T>
__device__ __forceinline__ uint32_t rope_rotate_pair(uint32_t xy, float c, float s) {
    using V = Vec2<T>;
    using T2 = typename V::type;
    const float2 v = V::to_f(*reinterpret_cast<T2*>(&xy));
    T2 r = V::from_f(v.x * c - v.y * s, v.y * c + v.x * s);
    return *reinterpret_cast<uint32_t*>(&r);
}

// Rotates the four pairs held in eight consecutive elements starting at element index `elem0` (a multiple of 8)
// of a head.
template <typename T>
__device__ __forceinline__ uint4 rope_rotate8(uint4 w, int elem0, int pos, int rot_dim, const void* cos_table,
                                              const void* sin_table) {
    using V = Vec2<T>;
    using T2 = typename V::type;
    if (cos_table && elem0 + 8 <= rot_dim && rot_dim % 8 == 0) {
        // the four pairs' cos and sin are 8 contiguous, 8-byte aligned bytes each
        const int64_t ti = (int64_t)pos * (rot_dim / 2) + elem0 / 2;
        uint2 cw = *reinterpret_cast<const uint2*>(reinterpret_cast<const T*>(cos_table) + ti);
        uint2 sw = *reinterpret_cast<const uint2*>(reinterpret_cast<const T*>(sin_table) + ti);
        const float2 c01 = V::to_f(*reinterpret_cast<T2*>(&cw.x)), c23 = V::to_f(*reinterpret_cast<T2*>(&cw.y));
        const float2 s01 = V::to_f(*reinterpret_cast<T2*>(&sw.x)), s23 = V::to_f(*reinterpret_cast<T2*>(&sw.y));
        return make_uint4(rope_rotate_pair<T>(w.x, c01.x, s01.x), rope_rotate_pair<T>(w.y, c01.y, s01.y),
                          rope_rotate_pair<T>(w.z, c23.x, s23.x), rope_rotate_pair<T>(w.w, c23.y, s23.y));
    }
    uint32_t x0 = w.x, x1 = w.y, x2 = w.z, x3 = w.w;
    const int pair0 = elem0 / 2;
    float c, s;
    if (2 * (pair0 + 0) < rot_dim) { rope_cos_sin<T>(cos_table, sin_table, pos, pair0 + 0, rot_dim, c, s); x0 = rope_rotate_pair<T>(x0, c, s); }
    if (2 * (pair0 + 1) < rot_dim) { rope_cos_sin<T>(cos_table, sin_table, pos, pair0 + 1, rot_dim, c, s); x1 = rope_rotate_pair<T>(x1, c, s); }
    if (2 * (pair0 + 2) < rot_dim) { rope_cos_sin<T>(cos_table, sin_table, pos, pair0 + 2, rot_dim, c, s); x2 = rope_rotate_pair<T>(x2, c, s); }
    if (2 * (pair0 + 3) < rot_dim) { rope_cos_sin<T>(cos_table, sin_table, pos, pair0 + 3, rot_dim, c, s); x3 = rope_rotate_pair<T>(x3, c, s); }
    return make_uint4(x0, x1, x2, x3);
}

}  // namespace sfa
