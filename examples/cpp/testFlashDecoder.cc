// C++ caller of the drop-in decoder step, equivalent to the reference's examples/cpp/testFlashDecoder.cc
// (same six shapes, B=2 H=32 D=128, 4 layers, all-ones fixture, 100 warm-up + 2 NVTX-ranged launches), but
// it checks the answer instead of printing it: with V == 1 every output element must be 1.
// The reference's fourth shape (max_seq_len 4096, seq_len 4096) overruns its cache; here the
// position is clamped to the last cache row.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <nvtx3/nvToolsExt.h>
#include <src/flash_attn.h>
#include <src/utils.h>

#include <cmath>
#include <cstdio>
#include <vector>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } \
    } while (0)

static int run_case(int batch, int heads, int dim, int max_len, int seq_len, int splits, int layers, int layer) {
    Flash_decoder_input in;
    Flash_decoder_params prm;
    prm.kBlockN = 128; prm.num_splits = splits; prm.kNThreads = 32;
    in.batch_size = batch; in.num_heads = heads; in.head_dim = dim;
    in.head_dim_inv = 1.0f / std::sqrt((float)dim);
    in.memory_max_len = max_len; in.max_input_length = max_len; in.rotary_embedding_dim = dim;
    in.stride = 3 * heads * dim; in.num_layer = layers; in.idx_layer = layer;

    const size_t tok = (size_t)batch * 3 * heads * dim, out_n = (size_t)batch * heads * dim;
    const size_t cache = (size_t)batch * layers * max_len * heads * dim;
    CK(cudaMalloc((void **)&in.qkv, tok * sizeof(half)));
    CK(cudaMalloc((void **)&in.o, out_n * sizeof(half)));
    CK(cudaMalloc((void **)&in.k_cache_table, cache * sizeof(half)));
    CK(cudaMalloc((void **)&in.v_cache_table, cache * sizeof(half)));
    CK(cudaMalloc((void **)&in.seq_len, batch * sizeof(int)));
    CK(cudaMalloc((void **)&in.rotary_cos_table, (size_t)max_len * dim / 2 * sizeof(half)));
    CK(cudaMalloc((void **)&in.rotary_sin_table, (size_t)max_len * dim / 2 * sizeof(half)));
    compute_rotary_table<half>((half*)in.rotary_cos_table, (half*)in.rotary_sin_table, max_len, dim);
    std::vector<int> pos(batch, seq_len < max_len ? seq_len : max_len - 1);
    CK(cudaMemcpy(in.seq_len, pos.data(), batch * sizeof(int), cudaMemcpyHostToDevice));
    const half one = __float2half(1.0f);
    init_half_array((half*)in.qkv, one, (int)tok, 0, 0);
    init_half_array((half*)in.k_cache_table, one, (int)cache, 0, 0);
    init_half_array((half*)in.v_cache_table, one, (int)cache, 0, 0);

    cudaStream_t stream;
    CK(cudaStreamCreate(&stream));
    CK(cudaDeviceSynchronize());
    for (int i = 0; i < 100; ++i) run_flash_decoder<half>(in, prm, stream);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    nvtxRangePushA("run_flash_decoder Range");
    CK(cudaEventRecord(e0, stream));
    for (int i = 0; i < 2; ++i) run_flash_decoder<half>(in, prm, stream);
    CK(cudaEventRecord(e1, stream));
    nvtxRangePop();
    CK(cudaStreamSynchronize(stream));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));

    std::vector<half> o(out_n);
    CK(cudaMemcpy(o.data(), in.o, out_n * sizeof(half), cudaMemcpyDeviceToHost));
    float worst = 0.f;
    for (size_t i = 0; i < out_n; ++i) worst = std::fmax(worst, std::fabs(__half2float(o[i]) - 1.0f));
    printf("[case] max_len %5d seq_len %5d : %.1f us per step, max |o - 1| = %.2e  %s\n", max_len, pos[0], 500.f * ms,
           worst, worst <= 1e-3f ? "ok" : "MISMATCH");
    cudaFree(in.qkv); cudaFree(in.o); cudaFree(in.k_cache_table); cudaFree(in.v_cache_table); cudaFree(in.seq_len);
    cudaFree(in.rotary_cos_table); cudaFree(in.rotary_sin_table);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(stream);
    return worst <= 1e-3f ? 0 : 2;
}

int main() {
    int dev_count = 0;
    if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0) {
        fprintf(stderr, "no CUDA device: this library has no CPU fallback\n");
        return 3;
    }
    const int max_len[6] = {512, 1024, 2048, 4096, 8192, 8192};
    const int seq_len[6] = {511, 1023, 2047, 4096, 6143, 8191};
    int rc = 0;
    for (int i = 0; i < 6; ++i) rc |= run_case(2, 32, 128, max_len[i], seq_len[i], 4, 4, 0);
    return rc;
}
