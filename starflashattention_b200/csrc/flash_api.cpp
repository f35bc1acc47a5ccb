// PyTorch extension module `star_flash_attn` — the reference's Python entry point (reference
// src/flash_api.cpp:42-80: mha_fwd_cuda with the same 16 keyword arguments, `o` written in place and
// returned, biases accepted and ignored, work queued on the current CUDA stream), bound to the C ABI of
// libStarFlashAttention.so.  Host-only translation unit: g++ compiles it, no CUDA code lives here.
//
// Differences from the reference, all additive:
//   * arguments are validated (the reference has no checks) and errors raise instead of printing;
//   * bf16 is accepted next to fp16;
//   * the split count is chosen by the library's launcher (the reference hard-codes 4, :37);
//   * `fwd` and `decode` expose the prefill forward and the GQA split-KV decode, which the reference lacks.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <cmath>
#include <tuple>

#include <star_flash_attn.h>

namespace {

int dtype_code(const at::Tensor& t) {
    if (t.scalar_type() == at::kHalf) return SFA_FP16;
    if (t.scalar_type() == at::kBFloat16) return SFA_BF16;
    TORCH_CHECK(false, "star_flash_attn: only float16 and bfloat16 tensors are supported, got ", t.scalar_type());
}

void check_rc(int rc) { TORCH_CHECK(rc == SFA_OK, "libStarFlashAttention error ", rc, ": ", sfa_last_error()); }

void* current_stream() { return at::cuda::getCurrentCUDAStream().stream(); }

}  // namespace

at::Tensor mha_fwd_cuda(at::Tensor& qkv, at::Tensor& q_bias, at::Tensor& k_bias, at::Tensor& v_bias,
                        at::Tensor& k_cache_table, at::Tensor& v_cache_table, at::Tensor& seq_len, at::Tensor& o,
                        int batch_size, int memory_max_len, int num_heads, int head_dim, int rotary_embedding_dim,
                        int max_input_length, int num_layer, int idx_layer) {
    (void)q_bias; (void)k_bias; (void)v_bias; (void)max_input_length;   // accepted and ignored, as in the reference
    TORCH_CHECK(qkv.is_cuda() && k_cache_table.is_cuda() && v_cache_table.is_cuda() && seq_len.is_cuda() && o.is_cuda(),
                "star_flash_attn: all tensors must be CUDA tensors (there is no CPU fallback)");
    TORCH_CHECK(seq_len.scalar_type() == at::kInt && seq_len.is_contiguous() && seq_len.numel() == batch_size,
                "seq_len must be a contiguous int32 tensor of shape (batch_size,)");
    TORCH_CHECK(qkv.is_contiguous() && qkv.numel() == (int64_t)batch_size * 3 * num_heads * head_dim,
                "qkv must be a contiguous (batch_size, 3, num_heads, head_dim) tensor");
    const int64_t cache_elems = (int64_t)batch_size * num_layer * memory_max_len * num_heads * head_dim;
    TORCH_CHECK(k_cache_table.is_contiguous() && v_cache_table.is_contiguous() &&
                    k_cache_table.numel() == cache_elems && v_cache_table.numel() == cache_elems,
                "caches must be contiguous (batch_size, num_layer, memory_max_len, num_heads, head_dim) tensors");
    TORCH_CHECK(o.is_contiguous() && o.numel() == (int64_t)batch_size * num_heads * head_dim,
                "o must be a contiguous (batch_size, num_heads, head_dim) tensor");
    TORCH_CHECK(qkv.scalar_type() == k_cache_table.scalar_type() && qkv.scalar_type() == v_cache_table.scalar_type() &&
                    qkv.scalar_type() == o.scalar_type(), "qkv, caches and o must share one dtype");
    c10::cuda::CUDAGuard guard(qkv.device());
    sfa_decoder_step_args a{};
    a.qkv = qkv.data_ptr();
    a.o = o.data_ptr();
    a.seq_len = seq_len.data_ptr<int>();
    a.k_cache_table = k_cache_table.data_ptr();
    a.v_cache_table = v_cache_table.data_ptr();
    a.batch_size = batch_size;
    a.memory_max_len = memory_max_len;
    a.num_heads = num_heads;
    a.head_dim = head_dim;
    a.head_dim_inv = 1.0f / std::sqrt((float)head_dim);
    a.rotary_embedding_dim = rotary_embedding_dim;
    a.qkv_stride = 3 * num_heads * head_dim;
    a.num_layer = num_layer;
    a.idx_layer = idx_layer;
    a.num_splits = 0;
    a.dtype = dtype_code(qkv);
    check_rc(sfa_decoder_step(&a, current_stream()));
    return o;
}

std::tuple<at::Tensor, at::Tensor> fwd(const at::Tensor& q, const at::Tensor& k, const at::Tensor& v, bool causal,
                                       double softmax_scale) {
    TORCH_CHECK(q.is_cuda() && k.is_cuda() && v.is_cuda(), "star_flash_attn.fwd: CUDA tensors required");
    TORCH_CHECK(q.dim() == 4 && k.dim() == 4 && v.dim() == 4, "q, k, v must be (batch, seqlen, heads, head_dim)");
    TORCH_CHECK(q.stride(3) == 1 && k.stride(3) == 1 && v.stride(3) == 1, "the head_dim axis must be contiguous");
    TORCH_CHECK(k.sizes() == v.sizes() && k.size(0) == q.size(0) && k.size(3) == q.size(3), "k/v shape mismatch");
    TORCH_CHECK(q.scalar_type() == k.scalar_type() && q.scalar_type() == v.scalar_type(), "q, k, v must share one dtype");
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor out = at::empty(q.sizes(), q.options());
    at::Tensor lse = at::empty({q.size(0), q.size(2), q.size(1)}, q.options().dtype(at::kFloat));
    sfa_fwd_args a{};
    a.q = q.data_ptr(); a.k = k.data_ptr(); a.v = v.data_ptr(); a.o = out.data_ptr();
    a.lse = lse.data_ptr<float>();
    a.batch = (int)q.size(0); a.seqlen_q = (int)q.size(1); a.seqlen_k = (int)k.size(1);
    a.heads_q = (int)q.size(2); a.heads_kv = (int)k.size(2); a.head_dim = (int)q.size(3);
    a.q_batch_stride = q.stride(0); a.q_row_stride = q.stride(1); a.q_head_stride = q.stride(2);
    a.k_batch_stride = k.stride(0); a.k_row_stride = k.stride(1); a.k_head_stride = k.stride(2);
    a.v_batch_stride = v.stride(0); a.v_row_stride = v.stride(1); a.v_head_stride = v.stride(2);
    a.o_batch_stride = out.stride(0); a.o_row_stride = out.stride(1); a.o_head_stride = out.stride(2);
    a.softmax_scale = (float)softmax_scale;
    a.causal = causal ? 1 : 0;
    a.dtype = dtype_code(q);
    check_rc(sfa_fwd(&a, current_stream()));
    return {out, lse};
}

std::tuple<at::Tensor, at::Tensor> decode(const at::Tensor& q, const at::Tensor& k_cache, const at::Tensor& v_cache,
                                          const c10::optional<at::Tensor>& seqlens, double softmax_scale,
                                          int num_splits) {
    TORCH_CHECK(q.is_cuda() && k_cache.is_cuda() && v_cache.is_cuda(), "star_flash_attn.decode: CUDA tensors required");
    TORCH_CHECK(q.dim() == 3 && k_cache.dim() == 4 && k_cache.sizes() == v_cache.sizes(),
                "q must be (batch, heads, head_dim); caches (batch, max_len, heads_kv, head_dim)");
    TORCH_CHECK(q.stride(2) == 1 && k_cache.stride(3) == 1 && v_cache.stride(3) == 1, "the head_dim axis must be contiguous");
    TORCH_CHECK(q.scalar_type() == k_cache.scalar_type() && q.scalar_type() == v_cache.scalar_type(), "dtype mismatch");
    if (seqlens.has_value())
        TORCH_CHECK(seqlens->is_cuda() && seqlens->scalar_type() == at::kInt && seqlens->is_contiguous() &&
                        seqlens->numel() == q.size(0), "seqlens must be a contiguous int32 CUDA tensor of shape (batch,)");
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor out = at::empty(q.sizes(), q.options());
    at::Tensor lse = at::empty({q.size(0), q.size(1)}, q.options().dtype(at::kFloat));
    sfa_decode_args a{};
    a.q = q.data_ptr(); a.k_cache = k_cache.data_ptr(); a.v_cache = v_cache.data_ptr();
    a.seqlens = seqlens.has_value() ? seqlens->data_ptr<int>() : nullptr;
    a.o = out.data_ptr(); a.lse = lse.data_ptr<float>();
    a.batch = (int)q.size(0); a.seqlen_k = (int)k_cache.size(1);
    a.heads_q = (int)q.size(1); a.heads_kv = (int)k_cache.size(2); a.head_dim = (int)q.size(2);
    a.q_batch_stride = q.stride(0); a.q_head_stride = q.stride(1);
    a.k_batch_stride = k_cache.stride(0); a.k_row_stride = k_cache.stride(1); a.k_head_stride = k_cache.stride(2);
    a.v_batch_stride = v_cache.stride(0); a.v_row_stride = v_cache.stride(1); a.v_head_stride = v_cache.stride(2);
    a.o_batch_stride = out.stride(0); a.o_head_stride = out.stride(1);
    a.softmax_scale = (float)softmax_scale;
    a.num_splits = num_splits;
    a.dtype = dtype_code(q);
    check_rc(sfa_decode(&a, current_stream()));
    return {out, lse};
}

// Decode over a paged cache: pools (pages, page_size, heads_kv, head_dim), block_table int32 (batch, blocks).
std::tuple<at::Tensor, at::Tensor> decode_paged(const at::Tensor& q, const at::Tensor& k_pool, const at::Tensor& v_pool,
                                                const at::Tensor& block_table, const at::Tensor& seqlens,
                                                double softmax_scale, int num_splits) {
    TORCH_CHECK(q.is_cuda() && k_pool.is_cuda() && v_pool.is_cuda() && block_table.is_cuda() && seqlens.is_cuda(),
                "star_flash_attn.decode_paged: CUDA tensors required");
    TORCH_CHECK(q.dim() == 3 && k_pool.dim() == 4 && k_pool.sizes() == v_pool.sizes() && k_pool.size(3) == q.size(2),
                "q must be (batch, heads, head_dim); pools (pages, page_size, heads_kv, head_dim)");
    TORCH_CHECK(q.stride(2) == 1 && k_pool.stride(3) == 1 && v_pool.stride(3) == 1, "the head_dim axis must be contiguous");
    TORCH_CHECK(q.scalar_type() == k_pool.scalar_type() && q.scalar_type() == v_pool.scalar_type(), "dtype mismatch");
    TORCH_CHECK(block_table.scalar_type() == at::kInt && block_table.dim() == 2 && block_table.size(0) == q.size(0) &&
                    block_table.stride(1) == 1, "block_table must be int32 (batch, blocks) with contiguous rows");
    TORCH_CHECK(seqlens.scalar_type() == at::kInt && seqlens.is_contiguous() && seqlens.numel() == q.size(0),
                "seqlens must be a contiguous int32 CUDA tensor of shape (batch,)");
    c10::cuda::CUDAGuard guard(q.device());
    at::Tensor out = at::empty(q.sizes(), q.options());
    at::Tensor lse = at::empty({q.size(0), q.size(1)}, q.options().dtype(at::kFloat));
    sfa_decode_paged_args p{};
    sfa_decode_args& a = p.base;
    a.q = q.data_ptr(); a.k_cache = k_pool.data_ptr(); a.v_cache = v_pool.data_ptr();
    a.seqlens = seqlens.data_ptr<int>();
    a.o = out.data_ptr(); a.lse = lse.data_ptr<float>();
    a.batch = (int)q.size(0); a.seqlen_k = (int)(block_table.size(1) * k_pool.size(1));
    a.heads_q = (int)q.size(1); a.heads_kv = (int)k_pool.size(2); a.head_dim = (int)q.size(2);
    a.q_batch_stride = q.stride(0); a.q_head_stride = q.stride(1);
    a.k_batch_stride = k_pool.stride(0); a.k_row_stride = k_pool.stride(1); a.k_head_stride = k_pool.stride(2);
    a.v_batch_stride = v_pool.stride(0); a.v_row_stride = v_pool.stride(1); a.v_head_stride = v_pool.stride(2);
    a.o_batch_stride = out.stride(0); a.o_head_stride = out.stride(1);
    a.softmax_scale = (float)softmax_scale;
    a.num_splits = num_splits;
    a.dtype = dtype_code(q);
    p.block_table = block_table.data_ptr<int>(); p.block_table_stride = block_table.stride(0);
    p.page_size = (int)k_pool.size(1); p.num_pages = (int)k_pool.size(0);
    check_rc(sfa_decode_paged(&p, current_stream()));
    return {out, lse};
}

PYBIND11_MODULE(star_flash_attn, m) {
    m.doc() = "B200-native FlashAttention forward and split-KV decode behind StarFlashAttention's Python entry point";
    m.def("mha_fwd_cuda", &mha_fwd_cuda, "One decoder step: RoPE + KV-cache append + split-KV attention (in-place on o)",
          py::arg("qkv"), py::arg("q_bias"), py::arg("k_bias"), py::arg("v_bias"), py::arg("k_cache_table"),
          py::arg("v_cache_table"), py::arg("seq_len"), py::arg("o"), py::arg("batch_size"), py::arg("memory_max_len"),
          py::arg("num_heads"), py::arg("head_dim"), py::arg("rotary_embedding_dim"), py::arg("max_input_length"),
          py::arg("num_layer"), py::arg("idx_layer"));
    m.def("fwd", &fwd, "Prefill attention forward: returns (out, lse)", py::arg("q"), py::arg("k"), py::arg("v"),
          py::arg("causal") = false, py::arg("softmax_scale") = 0.0);
    m.def("decode", &decode, "Split-KV decode attention: returns (out, lse)", py::arg("q"), py::arg("k_cache"),
          py::arg("v_cache"), py::arg("seqlens") = c10::nullopt, py::arg("softmax_scale") = 0.0, py::arg("num_splits") = 0);
    m.def("decode_paged", &decode_paged, "Split-KV decode attention over a paged cache: returns (out, lse)", py::arg("q"),
          py::arg("k_pool"), py::arg("v_pool"), py::arg("block_table"), py::arg("seqlens"), py::arg("softmax_scale") = 0.0,
          py::arg("num_splits") = 0);
    m.def("launch_count", []() { return (uint64_t)sfa_launch_count(); });
}
