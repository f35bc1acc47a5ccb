// Parameter structs of the decoder step — the drop-in contract of the reference's src/params.h:10-68.
// Field names, types and order of the reference are kept so that callers which fill the structs
// member by member (examples/cpp/testFlashDecoder.cc:15-33, src/flash_api.cpp:18-32) compile
// unchanged; new members are appended at the end and default to "reference behaviour".
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime_api.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <src/traits.h>

struct Flash_decoder_input
{
    // New token, (batch_size, 3, num_heads, head_dim); row stride `stride` elements.
    void *__restrict__ qkv = nullptr;
    // Accepted and ignored (as the reference does).
    void *__restrict__ q_bias = nullptr;
    void *__restrict__ k_bias = nullptr;
    void *__restrict__ v_bias = nullptr;
    // Output, (batch_size, num_heads, head_dim).
    void *__restrict__ o = nullptr;
    // int32 (batch_size): tokens already in the cache == position of the new token.
    void *__restrict__ seq_len = nullptr;
    // KV caches, (batch_size, num_layer, memory_max_len, num_heads, head_dim).
    void *__restrict__ k_cache_table = nullptr;
    void *__restrict__ v_cache_table = nullptr;
    // cos / sin tables, (memory_max_len, rotary_embedding_dim / 2) PACKED (row stride rotary_embedding_dim / 2, as
    // this library's compute_rotary_table<T> writes them; the reference kernel strides its rows by
    // rotary_embedding_dim - a table built for it is not interchangeable); both null => angles are evaluated in fp32
    // inside the kernel.
    void *__restrict__ rotary_cos_table = nullptr;
    void *__restrict__ rotary_sin_table = nullptr;

    int batch_size = 0;
    int memory_max_len = 0;
    int num_heads = 0;
    int head_dim = 0;
    float head_dim_inv = 0;          // softmax scale, 1/sqrt(head_dim)
    int rotary_embedding_dim = 0;
    int max_input_length = 0;
    int stride = 0;                  // 3 * num_heads * head_dim
    int num_layer = 0;
    int idx_layer = 0;

    // ---- additions (not in the reference) ----
    // Optional fp32 log-sum-exp output, (batch_size, num_heads).
    float *lse = nullptr;
};

struct Flash_decoder_params
{
    int kBlockN;     // kept for source compatibility; tile shapes are chosen by the launcher
    int num_splits;  // <= 0: launcher's choice
    int kNThreads;   // kept for source compatibility; ignored
};

struct Flash_decoder_buffers
{
    // (batch_size, num_heads, n_split, head_dim) fp32 un-normalised partial outputs
    void *o_split = nullptr;
    // (batch_size, num_heads, n_split) fp32 partial denominators
    void *ell = nullptr;
    // (batch_size, num_heads, n_split) fp32 running maxima (log2 domain)
    void *m_formula = nullptr;
};
