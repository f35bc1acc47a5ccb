// RoPE + KV-cache append prologue of the decoder step, the rotary table builder and the fill utility.
//
// Reference counterparts (src/flash_attn.cu): rotary_embedding_transform :167-188 and
// apply_rotary_embedding :191-257 (interleaved pairs (x,y) -> (c x - s y, c y + s x), position =
// seq_len[b]); the cache append :696-701; rotary_table_kernel / compute_rotary_table<half> :512-538;
// init_half_array_kernel / init_half_array :493-510.  Cache layout is the documented
// (batch, num_layer, memory_max_len, heads, head_dim) of src/params.h:22-25.
#include "internal.h"
#include "rope.cuh"

namespace sfa {
namespace {

struct RopeParams {
    const void* qkv;
    void* q_rot;
    void* k_cache;
    void* v_cache;
    const int32_t* seq_len;
    const void* cos_table;
    const void* sin_table;
    int64_t qkv_stride;
    int heads, head_dim, rotary_dim, max_len, num_layer, idx_layer;
};

// grid (heads, batch), block head_dim/2: thread i owns the element pair (2i, 2i+1) of q, k and v.
template <typename T>
__global__ void rope_append_kernel(const RopeParams p) {
    using V = Vec2<T>;
    using T2 = typename V::type;
    const int h = blockIdx.x, b = blockIdx.y, i = threadIdx.x;
    const int D = p.head_dim;
    const int pos = p.seq_len[b];
    const T* tok = reinterpret_cast<const T*>(p.qkv) + (int64_t)b * p.qkv_stride;
    const int64_t hd = (int64_t)h * D + 2 * i;
    const int64_t plane = (int64_t)p.heads * D;
    float2 q = V::to_f(*reinterpret_cast<const T2*>(tok + hd));
    float2 k = V::to_f(*reinterpret_cast<const T2*>(tok + plane + hd));
    const T2 v = *reinterpret_cast<const T2*>(tok + 2 * plane + hd);
    if (2 * i < p.rotary_dim) {
        float c, s;
        rope_cos_sin<T>(p.cos_table, p.sin_table, min(max(pos, 0), p.max_len - 1), i, p.rotary_dim, c, s);   // tables hold max_len rows
        q = make_float2(q.x * c - q.y * s, q.y * c + q.x * s);
        k = make_float2(k.x * c - k.y * s, k.y * c + k.x * s);
    }
    T* qo = reinterpret_cast<T*>(p.q_rot) + ((int64_t)b * p.heads + h) * D + 2 * i;
    *reinterpret_cast<T2*>(qo) = V::from_f(q.x, q.y);
    if (pos >= 0 && pos < p.max_len) {
        const int64_t off = ((((int64_t)b * p.num_layer + p.idx_layer) * p.max_len + pos) * p.heads + h) * D + 2 * i;
        *reinterpret_cast<T2*>(reinterpret_cast<T*>(p.k_cache) + off) = V::from_f(k.x, k.y);
        *reinterpret_cast<T2*>(reinterpret_cast<T*>(p.v_cache) + off) = v;
    }
}

template <typename T>
__global__ void rotary_table_kernel(T* cos_t, T* sin_t, int rot_dim) {
    const int pos = blockIdx.x;
    for (int i = threadIdx.x; i < rot_dim / 2; i += blockDim.x) {
        float s, c;
        sincosf(rope_angle(pos, i, rot_dim), &s, &c);
        const int64_t ti = (int64_t)pos * (rot_dim / 2) + i;
        cos_t[ti] = Vec2<T>::one_from_f(c);
        sin_t[ti] = Vec2<T>::one_from_f(s);
    }
}

template <typename T>
__global__ void fill_kernel(T* arr, T value, int64_t n) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) arr[i] = value;
}

}  // namespace

int launch_rope_append(const RopeAppendLaunch& a, cudaStream_t stream) {
    if (a.head_dim % 2 || a.rotary_dim % 2 || a.rotary_dim > a.head_dim || a.rotary_dim < 0)
        return fail(SFA_ERR_INVALID, "rotary_embedding_dim must be even and <= head_dim");
    if ((a.cos_table == nullptr) != (a.sin_table == nullptr))
        return fail(SFA_ERR_INVALID, "rotary cos and sin tables must be given together");
    if (a.batch == 0) return SFA_OK;
    RopeParams p;
    p.qkv = a.qkv; p.q_rot = a.q_rot; p.k_cache = a.k_cache; p.v_cache = a.v_cache; p.seq_len = a.seq_len;
    p.cos_table = a.cos_table; p.sin_table = a.sin_table; p.qkv_stride = a.qkv_stride;
    p.heads = a.heads; p.head_dim = a.head_dim; p.rotary_dim = a.rotary_dim; p.max_len = a.max_len;
    p.num_layer = a.num_layer; p.idx_layer = a.idx_layer;
    dim3 grid(a.heads, a.batch);
    if (a.dtype == SFA_FP16) rope_append_kernel<__half><<<grid, a.head_dim / 2, 0, stream>>>(p);
    else rope_append_kernel<__nv_bfloat16><<<grid, a.head_dim / 2, 0, stream>>>(p);
    SFA_CUDA_CHECK(cudaGetLastError());
    count_launch();
    return SFA_OK;
}

int launch_rotary_table(void* cos_t, void* sin_t, int max_seq_len, int rot_dim, int dtype, cudaStream_t stream) {
    if (!cos_t || !sin_t || max_seq_len < 0 || rot_dim <= 0 || rot_dim % 2)
        return fail(SFA_ERR_INVALID, "rotary table: null pointer or bad shape");
    if (max_seq_len == 0) return SFA_OK;
    const int threads = rot_dim / 2 < 128 ? ((rot_dim / 2 + 31) / 32) * 32 : 128;
    if (dtype == SFA_FP16)
        rotary_table_kernel<__half><<<max_seq_len, threads, 0, stream>>>((__half*)cos_t, (__half*)sin_t, rot_dim);
    else
        rotary_table_kernel<__nv_bfloat16><<<max_seq_len, threads, 0, stream>>>((__nv_bfloat16*)cos_t, (__nv_bfloat16*)sin_t, rot_dim);
    SFA_CUDA_CHECK(cudaGetLastError());
    count_launch();
    return SFA_OK;
}

int launch_fill(void* arr, float value, int64_t n, int dtype, cudaStream_t stream) {
    if (n < 0 || (!arr && n > 0)) return fail(SFA_ERR_INVALID, "fill: null pointer or negative size");
    if (n == 0) return SFA_OK;
    const DeviceInfo* dev = current_device();
    if (!dev) return SFA_ERR_NO_DEVICE;
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)dev->sm_count * 16;
    if (blocks > cap) blocks = cap;
    if (dtype == SFA_FP16) fill_kernel<__half><<<(int)blocks, 256, 0, stream>>>((__half*)arr, __float2half_rn(value), n);
    else fill_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, stream>>>((__nv_bfloat16*)arr, __float2bfloat16_rn(value), n);
    SFA_CUDA_CHECK(cudaGetLastError());
    count_launch();
    return SFA_OK;
}

}  // namespace sfa
