// TEST INFRASTRUCTURE (oracle/): runs the REFERENCE's own CUDA decode kernel - compiled unchanged from
// /root/reference/src/flash_attn.cu into oracle/_ref/libStarFlashAttention_ref.so by oracle/Makefile - on the five safe
// all-ones fixtures of its example (/root/reference/examples/cpp/testFlashDecoder.cc:138-150 without the (4096, 4096)
// case that overruns its cache), with rotary tables allocated at the size the kernel indexes
// (M x D halves), and prints what it computes and how long a call takes: whole call (its launcher
// mallocs, copies and synchronises per call, src/flash_attn.cu:937-1018) and the kernels alone (CUDA events around
// a second call measure the same thing: the launcher synchronises inside).  The kernel is unfinished (SURVEY.md
// appendix A), so its output is reported, not used as an oracle.  The product never links or loads this.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <src/flash_attn.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s\n", cudaGetErrorString(e_), #x); return 1; } } while (0)

static int run_case(int batch, int heads, int dim, int max_len, int seq_len, int splits, int layers, int layer) {
    Flash_decoder_input in;
    Flash_decoder_params prm;
    prm.num_splits = splits; prm.kNThreads = 32; prm.kBlockN = 128;
    in.batch_size = batch; in.num_heads = heads; in.head_dim = dim; in.head_dim_inv = 1.0f / std::sqrt((float)dim);
    in.memory_max_len = max_len; in.max_input_length = max_len; in.rotary_embedding_dim = dim;
    in.stride = 3 * heads * dim; in.idx_layer = layer; in.num_layer = layers;
    const size_t tok = (size_t)batch * 3 * heads * dim, out_n = (size_t)batch * heads * dim;
    const size_t cache = (size_t)batch * layers * max_len * heads * dim;
    CK(cudaMalloc((void**)&in.qkv, tok * 2)); CK(cudaMalloc((void**)&in.o, out_n * 2));
    CK(cudaMalloc((void**)&in.k_cache_table, cache * 2)); CK(cudaMalloc((void**)&in.v_cache_table, cache * 2));
    CK(cudaMalloc((void**)&in.seq_len, batch * sizeof(int)));
    CK(cudaMalloc((void**)&in.rotary_cos_table, (size_t)max_len * dim * 2));   // full size: the kernel indexes pos * D + t
    CK(cudaMalloc((void**)&in.rotary_sin_table, (size_t)max_len * dim * 2));
    CK(cudaMemset(in.rotary_cos_table, 0, (size_t)max_len * dim * 2)); CK(cudaMemset(in.rotary_sin_table, 0, (size_t)max_len * dim * 2));
    compute_rotary_table<half>((half*)in.rotary_cos_table, (half*)in.rotary_sin_table, max_len, dim);
    std::vector<int> pos(batch, seq_len);
    CK(cudaMemcpy(in.seq_len, pos.data(), batch * sizeof(int), cudaMemcpyHostToDevice));
    const half one = __float2half(1.0f);
    init_half_array((half*)in.qkv, one, (int)tok, (int)((tok + 255) / 256), 256);
    init_half_array((half*)in.k_cache_table, one, (int)cache, (int)((cache + 255) / 256), 256);
    init_half_array((half*)in.v_cache_table, one, (int)cache, (int)((cache + 255) / 256), 256);
    CK(cudaMemset(in.o, 0, out_n * 2));
    CK(cudaDeviceSynchronize());
    cudaStream_t stream;
    CK(cudaStreamCreate(&stream));
    for (int i = 0; i < 5; ++i) run_flash_decoder<half>(in, prm, stream);
    CK(cudaDeviceSynchronize());
    const int iters = 20;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto t0 = std::chrono::steady_clock::now();
    CK(cudaEventRecord(e0, stream));
    for (int i = 0; i < iters; ++i) run_flash_decoder<half>(in, prm, stream);
    CK(cudaEventRecord(e1, stream));
    CK(cudaStreamSynchronize(stream));
    auto t1 = std::chrono::steady_clock::now();
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double wall_us = std::chrono::duration<double, std::micro>(t1 - t0).count() / iters;
    std::vector<half> o(out_n);
    CK(cudaMemcpy(o.data(), in.o, out_n * 2, cudaMemcpyDeviceToHost));
    double worst = 0.0; int nan = 0;
    for (size_t i = 0; i < out_n; ++i) {
        const float v = __half2float(o[i]);
        if (v != v) ++nan; else worst = std::fmax(worst, std::fabs(v - 1.0f));
    }
    printf("[ref kernel] max_len %5d seq_len %5d : whole call %.1f us (host clock), %.1f us (events); o[0..3] = %.4f %.4f %.4f %.4f; "
           "max |o - 1| = %.3e, NaNs %d of %zu\n", max_len, seq_len, wall_us, 1e3 * ms / iters, __half2float(o[0]), __half2float(o[1]),
           __half2float(o[2]), __half2float(o[3]), worst, nan, out_n);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("[ref kernel] sticky error: %s\n", cudaGetErrorString(e));
    return 0;
}

int main() {
    const int max_len[5] = {512, 1024, 2048, 8192, 8192};
    const int seq_len[5] = {511, 1023, 2047, 6143, 8191};
    for (int i = 0; i < 5; ++i)
        if (run_case(2, 32, 128, max_len[i], seq_len[i], 4, 4, 0)) return 1;
    return 0;
}
