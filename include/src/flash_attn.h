// C++ face of libStarFlashAttention.so — drop-in for the reference's src/flash_attn.h:1-11.
// The three host functions keep the reference's names, signatures and C++ mangling
// (_Z17run_flash_decoderI6__halfEvR19Flash_decoder_inputR20Flash_decoder_paramsP11CUstream_st,
//  _Z20compute_rotary_tableI6__halfEvPT_S2_ii, _Z15init_half_arrayP6__halfS_iii).
// The reference also declared its two __global__ kernels here; they are an implementation detail
// of the library now (and this header is included by host-only translation units).
#pragma once
#include <src/params.h>
#include <cuda_bf16.h>

// Rotate + append + attend for one new token per sequence; asynchronous on `stream`.
// T in {half, __nv_bfloat16}.  Errors are reported on stderr (the reference's behaviour,
// src/flash_attn.cu:979-991); use the C ABI (include/star_flash_attn.h) for status codes.
template <typename T>
void run_flash_decoder(Flash_decoder_input &input, Flash_decoder_params &params, cudaStream_t stream);

// cos/sin tables of shape (max_seq_len, rot_embed_dim / 2).
template <typename T>
void compute_rotary_table(T *rotary_cos_table, T *rotary_sin_table, int max_seq_len, int rot_embed_dim);

// Test utility: array[i] = value for i < n.  numBlocks/blockSize are accepted for source
// compatibility; the launch shape is chosen by the library.
void init_half_array(half *array, half value, int n, int numBlocks, int blockSize);

// ---- additions: prefill forward and GQA / bf16 decode (no reference counterpart) ----
struct Flash_fwd_params
{
    const void *q = nullptr, *k = nullptr, *v = nullptr;  // (B,S,H,D), strides in elements
    void *o = nullptr;
    float *lse = nullptr;                                  // (B,Hq,Sq) fp32 or null
    int batch = 0, seqlen_q = 0, seqlen_k = 0, num_heads = 0, num_heads_k = 0, head_dim = 0;
    int64_t q_batch_stride = 0, q_row_stride = 0, q_head_stride = 0;
    int64_t k_batch_stride = 0, k_row_stride = 0, k_head_stride = 0;
    int64_t v_batch_stride = 0, v_row_stride = 0, v_head_stride = 0;
    int64_t o_batch_stride = 0, o_row_stride = 0, o_head_stride = 0;
    float softmax_scale = 0.f;                             // <= 0: 1/sqrt(head_dim)
    bool is_causal = false;
};
template <typename T>
void run_flash_fwd(Flash_fwd_params &params, cudaStream_t stream);

struct Flash_decode_params
{
    const void *q = nullptr, *k_cache = nullptr, *v_cache = nullptr;
    const int *seqlens = nullptr;                          // null => seqlen_k for all
    void *o = nullptr;
    float *lse = nullptr;
    int batch = 0, seqlen_k = 0, num_heads = 0, num_heads_k = 0, head_dim = 0;
    int64_t q_batch_stride = 0, q_head_stride = 0;
    int64_t k_batch_stride = 0, k_row_stride = 0, k_head_stride = 0;
    int64_t v_batch_stride = 0, v_row_stride = 0, v_head_stride = 0;
    int64_t o_batch_stride = 0, o_head_stride = 0;
    float softmax_scale = 0.f;
    int num_splits = 0;
};
template <typename T>
void run_flash_decode(Flash_decode_params &params, cudaStream_t stream);
