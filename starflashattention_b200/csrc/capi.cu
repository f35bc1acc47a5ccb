// Entry points of libStarFlashAttention.so: the C ABI of include/star_flash_attn.h and the
// C++-mangled drop-ins of the reference's src/flash_attn.h (run_flash_decoder<half>,
// compute_rotary_table<half>, init_half_array) plus the typed additions.
#include "internal.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <src/flash_attn.h>

namespace sfa {
namespace {

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int check_dtype(int dtype) {
    if (dtype != SFA_FP16 && dtype != SFA_BF16) return fail(SFA_ERR_INVALID, "dtype must be SFA_FP16 or SFA_BF16");
    return SFA_OK;
}

int resolve_splits(int requested, int batch, int heads_kv, int max_len, int sm_count) {
    if (requested > 0) return requested > 256 ? 256 : requested;
    return choose_decode_splits(batch, heads_kv, max_len, sm_count);
}

struct FusedStep {   // new-token inputs of the fused decoder step (see DecodeLaunch)
    const void* qkv; const void* cos_table; const void* sin_table;
    int64_t qkv_stride; int rotary_dim;
};

struct PagedKV {
    const int32_t* block_table; int64_t bt_stride; int page_size;
};

int decode_impl(const sfa_decode_args& a, int seqlen_add, cudaStream_t stream, const FusedStep* fused = nullptr,
                const PagedKV* paged = nullptr) {
    int rc = check_dtype(a.dtype);
    if (rc) return rc;
    if ((!a.q && !fused) || !a.k_cache || !a.v_cache || !a.o) return fail(SFA_ERR_INVALID, "decode: null tensor pointer");
    if (a.batch < 0 || a.seqlen_k < 0 || a.heads_q <= 0 || a.heads_kv <= 0)
        return fail(SFA_ERR_INVALID, "decode: negative or zero dimension");
    if (a.heads_q % a.heads_kv) return fail(SFA_ERR_INVALID, "decode: heads_q must be a multiple of heads_kv");
    const int64_t strides[] = {a.q_batch_stride, a.q_head_stride, a.k_batch_stride, a.k_row_stride, a.k_head_stride,
                               a.v_batch_stride, a.v_row_stride, a.v_head_stride};
    for (int64_t s : strides)
        if (s % 8 != 0 || s < 0) return fail(SFA_ERR_INVALID, "decode: q/k/v strides must be non-negative multiples of 8 elements");
    const uintptr_t ptrs[] = {(uintptr_t)a.q, (uintptr_t)a.k_cache, (uintptr_t)a.v_cache};
    for (uintptr_t x : ptrs)
        if (x % 16) return fail(SFA_ERR_INVALID, "decode: q/k/v base pointers must be 16-byte aligned");
    if (a.batch == 0) return SFA_OK;
    const DeviceInfo* dev = current_device();
    if (!dev) return SFA_ERR_NO_DEVICE;

    DecodeLaunch L;
    L.q = a.q; L.k = a.k_cache; L.v = a.v_cache; L.seqlens = a.seqlens; L.o = a.o; L.lse = a.lse;
    L.batch = a.batch; L.seqlen_k = a.seqlen_k; L.heads_q = a.heads_q; L.heads_kv = a.heads_kv; L.head_dim = a.head_dim;
    L.q_bs = a.q_batch_stride; L.q_hs = a.q_head_stride;
    L.k_bs = a.k_batch_stride; L.k_rs = a.k_row_stride; L.k_hs = a.k_head_stride;
    L.v_bs = a.v_batch_stride; L.v_rs = a.v_row_stride; L.v_hs = a.v_head_stride;
    L.o_bs = a.o_batch_stride; L.o_hs = a.o_head_stride;
    L.scale = a.softmax_scale > 0.f ? a.softmax_scale : 1.f / sqrtf((float)a.head_dim);
    L.seqlen_add = seqlen_add;
    L.num_splits = resolve_splits(a.num_splits, a.batch, a.heads_kv * decode_head_chunks(a.heads_q, a.heads_kv), a.seqlen_k + seqlen_add, dev->sm_count);
    L.dtype = a.dtype;
    L.o_part = nullptr; L.ml_part = nullptr;
    if (fused) {
        L.fused_qkv = fused->qkv; L.cos_table = fused->cos_table; L.sin_table = fused->sin_table;
        L.qkv_stride = fused->qkv_stride; L.rotary_dim = fused->rotary_dim;
    }
    if (paged) { L.block_table = paged->block_table; L.bt_stride = paged->bt_stride; L.page_size = paged->page_size; }
    if (L.num_splits > 1) {
        const size_t need = decode_workspace_bytes(a.batch, a.heads_q, a.head_dim, L.num_splits);
        uint8_t* ws = reinterpret_cast<uint8_t*>(a.workspace);
        if (ws) {
            if (a.workspace_bytes < need) return fail(SFA_ERR_INVALID, "decode: workspace too small, need " + std::to_string(need) + " bytes");
            if ((uintptr_t)ws % 16) return fail(SFA_ERR_INVALID, "decode: workspace must be 16-byte aligned");
        } else {
            ws = reinterpret_cast<uint8_t*>(get_workspace(need, stream));
            if (!ws) return SFA_ERR_CUDA;
        }
        const size_t rows = (size_t)a.batch * a.heads_q * L.num_splits;
        L.o_part = reinterpret_cast<float*>(ws);
        L.ml_part = reinterpret_cast<float2*>(ws + align_up(rows * a.head_dim * sizeof(float), 16));
    }
    return launch_decode(L, stream);
}

int decoder_step_impl(const sfa_decoder_step_args& a, cudaStream_t stream) {
    int rc = check_dtype(a.dtype);
    if (rc) return rc;
    if (!a.qkv || !a.o || !a.seq_len || !a.k_cache_table || !a.v_cache_table)
        return fail(SFA_ERR_INVALID, "decoder step: null tensor pointer");
    if (a.batch_size < 0 || a.memory_max_len <= 0 || a.num_heads <= 0 || a.num_layer <= 0 || a.idx_layer < 0 ||
        a.idx_layer >= a.num_layer)
        return fail(SFA_ERR_INVALID, "decoder step: bad dimension or layer index");
    if (a.head_dim != 64 && a.head_dim != 128) return fail(SFA_ERR_UNSUPPORTED, "decoder step: head_dim must be 64 or 128");
    if (a.qkv_stride < 3 * a.num_heads * a.head_dim || a.qkv_stride % 8)
        return fail(SFA_ERR_INVALID, "decoder step: qkv stride must be >= 3*num_heads*head_dim and a multiple of 8");
    if (a.batch_size == 0) return SFA_OK;
    const DeviceInfo* dev = current_device();
    if (!dev) return SFA_ERR_NO_DEVICE;

    const int B = a.batch_size, H = a.num_heads, D = a.head_dim;
    if (a.rotary_embedding_dim % 2 || a.rotary_embedding_dim > D || a.rotary_embedding_dim < 0)
        return fail(SFA_ERR_INVALID, "rotary_embedding_dim must be even and <= head_dim");
    if ((a.rotary_cos_table == nullptr) != (a.rotary_sin_table == nullptr))
        return fail(SFA_ERR_INVALID, "rotary cos and sin tables must be given together");
    // One kernel (RoPE + append inside the split-KV kernel) unless SFA_DECODER_FUSED=0 asks for the three-launch
    // form (stand-alone RoPE/append kernel first), kept for A/B checks.
    const bool fused = decode_tuning().fused_step != 0;
    const int splits = resolve_splits(a.num_splits, B, H, a.memory_max_len, dev->sm_count);
    const size_t q_bytes = fused ? 0 : align_up((size_t)B * H * D * 2, 256);
    const size_t part_bytes = decode_workspace_bytes(B, H, D, splits);
    uint8_t* ws = nullptr;
    if (q_bytes + part_bytes) {
        if (a.workspace) {
            if (a.workspace_bytes < q_bytes + part_bytes)
                return fail(SFA_ERR_INVALID, "decoder step: workspace too small, need " + std::to_string(q_bytes + part_bytes) + " bytes");
            if ((uintptr_t)a.workspace % 256) return fail(SFA_ERR_INVALID, "decoder step: workspace must be 256-byte aligned");
            ws = reinterpret_cast<uint8_t*>(a.workspace);
        } else {
            ws = reinterpret_cast<uint8_t*>(get_workspace(q_bytes + part_bytes, stream));
            if (!ws) return SFA_ERR_CUDA;
        }
    }

    if (!fused) {
        RopeAppendLaunch R;
        R.qkv = a.qkv; R.qkv_stride = a.qkv_stride; R.q_rot = ws;
        R.k_cache = a.k_cache_table; R.v_cache = a.v_cache_table; R.seq_len = a.seq_len;
        R.cos_table = a.rotary_cos_table; R.sin_table = a.rotary_sin_table;
        R.batch = B; R.heads = H; R.head_dim = D; R.rotary_dim = a.rotary_embedding_dim; R.max_len = a.memory_max_len;
        R.num_layer = a.num_layer; R.idx_layer = a.idx_layer; R.dtype = a.dtype;
        if ((rc = launch_rope_append(R, stream))) return rc;
    }

    const int64_t layer_elems = (int64_t)a.memory_max_len * H * D;
    sfa_decode_args d = {};
    d.q = fused ? nullptr : ws;
    d.k_cache = reinterpret_cast<const uint8_t*>(a.k_cache_table) + 2 * layer_elems * a.idx_layer;
    d.v_cache = reinterpret_cast<const uint8_t*>(a.v_cache_table) + 2 * layer_elems * a.idx_layer;
    d.seqlens = a.seq_len; d.o = a.o; d.lse = a.lse;
    d.batch = B; d.seqlen_k = a.memory_max_len - 1; d.heads_q = H; d.heads_kv = H; d.head_dim = D;
    d.q_batch_stride = (int64_t)H * D; d.q_head_stride = D;
    d.k_batch_stride = layer_elems * a.num_layer; d.k_row_stride = (int64_t)H * D; d.k_head_stride = D;
    d.v_batch_stride = d.k_batch_stride; d.v_row_stride = d.k_row_stride; d.v_head_stride = D;
    d.o_batch_stride = (int64_t)H * D; d.o_head_stride = D;
    d.softmax_scale = a.head_dim_inv;
    d.num_splits = splits; d.dtype = a.dtype;
    d.workspace = part_bytes ? ws + q_bytes : nullptr; d.workspace_bytes = part_bytes;
    if (!fused) return decode_impl(d, /*seqlen_add=*/1, stream);
    if ((uintptr_t)a.qkv % 16) return fail(SFA_ERR_INVALID, "decoder step: qkv must be 16-byte aligned");
    const FusedStep f = {a.qkv, a.rotary_cos_table, a.rotary_sin_table, a.qkv_stride, a.rotary_embedding_dim};
    return decode_impl(d, /*seqlen_add=*/1, stream, &f);
}

int decoder_step_paged_impl(const sfa_decoder_step_paged_args& a, cudaStream_t stream) {
    int rc = check_dtype(a.dtype);
    if (rc) return rc;
    if (!a.qkv || !a.o || !a.seq_len || !a.k_pool || !a.v_pool || !a.block_table)
        return fail(SFA_ERR_INVALID, "paged decoder step: null tensor pointer");
    if (a.batch_size < 0 || a.max_len <= 0 || a.heads_q <= 0 || a.heads_kv <= 0 || a.num_pages <= 0)
        return fail(SFA_ERR_INVALID, "paged decoder step: bad dimension");
    if (a.page_size < 16 || (a.page_size & (a.page_size - 1)))
        return fail(SFA_ERR_INVALID, "paged decoder step: page_size must be a power of two >= 16");
    if (a.block_table_stride < ((int64_t)a.max_len + a.page_size - 1) / a.page_size)
        return fail(SFA_ERR_INVALID, "paged decoder step: block table rows are shorter than max_len / page_size");
    if (a.qkv_stride < (int64_t)(a.heads_q + 2 * a.heads_kv) * a.head_dim || a.qkv_stride % 8)
        return fail(SFA_ERR_INVALID, "paged decoder step: qkv stride must be >= (heads_q + 2 heads_kv)*head_dim and a multiple of 8");
    if ((uintptr_t)a.qkv % 16) return fail(SFA_ERR_INVALID, "paged decoder step: qkv must be 16-byte aligned");
    if (a.rotary_embedding_dim % 2 || a.rotary_embedding_dim > a.head_dim || a.rotary_embedding_dim < 0)
        return fail(SFA_ERR_INVALID, "rotary_embedding_dim must be even and <= head_dim");
    if ((a.rotary_cos_table == nullptr) != (a.rotary_sin_table == nullptr))
        return fail(SFA_ERR_INVALID, "rotary cos and sin tables must be given together");
    sfa_decode_args d = {};
    d.q = nullptr; d.k_cache = a.k_pool; d.v_cache = a.v_pool;
    d.seqlens = a.seq_len; d.o = a.o; d.lse = a.lse;
    d.batch = a.batch_size; d.seqlen_k = a.max_len - 1; d.heads_q = a.heads_q; d.heads_kv = a.heads_kv; d.head_dim = a.head_dim;
    d.q_batch_stride = (int64_t)a.heads_q * a.head_dim; d.q_head_stride = a.head_dim;
    d.k_batch_stride = a.page_stride; d.k_row_stride = a.row_stride; d.k_head_stride = a.head_stride;
    d.v_batch_stride = a.page_stride; d.v_row_stride = a.row_stride; d.v_head_stride = a.head_stride;
    d.o_batch_stride = (int64_t)a.heads_q * a.head_dim; d.o_head_stride = a.head_dim;
    d.softmax_scale = a.softmax_scale;
    d.num_splits = a.num_splits; d.dtype = a.dtype;
    const FusedStep f = {a.qkv, a.rotary_cos_table, a.rotary_sin_table, a.qkv_stride, a.rotary_embedding_dim};
    const PagedKV pk = {a.block_table, a.block_table_stride, a.page_size};
    return decode_impl(d, /*seqlen_add=*/1, stream, &f, &pk);
}

// Local split-KV pass of the sharded decode: always leaves (o_part, ml_part) in the workspace; fills the
// exchange launch for this rank.
int shard_local_pass(const sfa_decode_args& a, int rank, int world, ShardLaunch& S, cudaStream_t stream) {
    int rc = check_dtype(a.dtype);
    if (rc) return rc;
    if (!a.q || !a.k_cache || !a.v_cache || !a.o) return fail(SFA_ERR_INVALID, "sharded decode: null tensor pointer");
    if (world < 1 || world > 8 || rank < 0 || rank >= world)
        return fail(SFA_ERR_INVALID, "sharded decode: world must be 1..8 and 0 <= rank < world");
    if (a.heads_kv <= 0 || a.heads_q % a.heads_kv) return fail(SFA_ERR_INVALID, "sharded decode: heads_q must be a multiple of heads_kv");
    if (a.batch <= 0) return fail(SFA_ERR_INVALID, "sharded decode: batch must be positive");
    const DeviceInfo* dev = current_device();
    if (!dev) return SFA_ERR_NO_DEVICE;

    DecodeLaunch L;
    L.q = a.q; L.k = a.k_cache; L.v = a.v_cache; L.seqlens = a.seqlens; L.o = a.o; L.lse = a.lse;
    L.batch = a.batch; L.seqlen_k = a.seqlen_k; L.heads_q = a.heads_q; L.heads_kv = a.heads_kv; L.head_dim = a.head_dim;
    L.q_bs = a.q_batch_stride; L.q_hs = a.q_head_stride;
    L.k_bs = a.k_batch_stride; L.k_rs = a.k_row_stride; L.k_hs = a.k_head_stride;
    L.v_bs = a.v_batch_stride; L.v_rs = a.v_row_stride; L.v_hs = a.v_head_stride;
    L.o_bs = a.o_batch_stride; L.o_hs = a.o_head_stride;
    L.scale = a.softmax_scale > 0.f ? a.softmax_scale : 1.f / sqrtf((float)a.head_dim);
    L.seqlen_add = 0;
    L.num_splits = resolve_splits(a.num_splits, a.batch, a.heads_kv * decode_head_chunks(a.heads_q, a.heads_kv), a.seqlen_k, dev->sm_count);
    L.dtype = a.dtype;
    L.force_partials = true;
    const size_t rows = (size_t)a.batch * a.heads_q * L.num_splits;
    const size_t need = rows * a.head_dim * sizeof(float) + rows * sizeof(float2) + 256;
    uint8_t* ws = reinterpret_cast<uint8_t*>(a.workspace);
    if (ws) {
        if (a.workspace_bytes < need) return fail(SFA_ERR_INVALID, "sharded decode: workspace too small, need " + std::to_string(need) + " bytes");
    } else {
        ws = reinterpret_cast<uint8_t*>(get_workspace(need, stream));
        if (!ws) return SFA_ERR_CUDA;
    }
    L.o_part = reinterpret_cast<float*>(ws);
    L.ml_part = reinterpret_cast<float2*>(ws + align_up(rows * a.head_dim * sizeof(float), 16));
    if ((rc = launch_decode(L, stream))) return rc;

    S = ShardLaunch{};
    S.o_part = L.o_part; S.ml_part = L.ml_part; S.num_splits = L.num_splits;
    S.o = a.o; S.lse = a.lse; S.o_bs = a.o_batch_stride; S.o_hs = a.o_head_stride;
    S.batch = a.batch; S.heads_q = a.heads_q; S.head_dim = a.head_dim; S.dtype = a.dtype;
    S.rank = rank; S.world = world;
    return SFA_OK;
}

int decode_sharded_impl(const sfa_decode_args& a, const sfa_shard_comm& c, cudaStream_t stream) {
    if (c.mode != 0 && c.mode != 1) return fail(SFA_ERR_INVALID, "sharded decode: mode must be 0 (peer exchange) or 1 (publish)");
    if (c.world < 1 || c.world > 8 || c.rank < 0 || c.rank >= c.world)
        return fail(SFA_ERR_INVALID, "sharded decode: world must be 1..8 and 0 <= rank < world");
    if (c.mode == 0)
        for (int r = 0; r < c.world; ++r)
            if (!c.peer_ws[r]) return fail(SFA_ERR_INVALID, "sharded decode: missing peer workspace pointer");
    if (c.mode == 1 && !c.payload) return fail(SFA_ERR_INVALID, "sharded decode: publish mode needs a payload buffer");
    ShardLaunch S;
    int rc = shard_local_pass(a, c.rank, c.world, S, stream);
    if (rc) return rc;
    S.step = c.step; S.peer_ws = c.peer_ws; S.payload = c.payload; S.mode = c.mode;
    return launch_shard_exchange(S, stream);
}

// Test aid: `world` ranks on ONE GPU.  The local passes run one after the other; the exchange of all ranks is one
// cooperative launch.  Every rank's args must carry its own split workspace.
int decode_sharded_emulated_impl(const sfa_decode_args* args, int world, void* const* ws, uint32_t step, cudaStream_t stream) {
    if (world < 1 || world > 8) return fail(SFA_ERR_INVALID, "sharded decode: world must be 1..8");
    ShardLaunch S[8];
    for (int r = 0; r < world; ++r) {
        if (!args[r].workspace) return fail(SFA_ERR_INVALID, "emulated sharded decode: every rank needs its own workspace");
        if (!ws[r]) return fail(SFA_ERR_INVALID, "emulated sharded decode: missing exchange workspace");
        int rc = shard_local_pass(args[r], r, world, S[r], stream);
        if (rc) return rc;
        S[r].step = step; S[r].peer_ws = ws; S[r].mode = 0;
    }
    return launch_shard_exchange_emulated(S, world, stream);
}

}  // namespace
}  // namespace sfa

// ============================================================================ C ABI
extern "C" {

int sfa_fwd(const sfa_fwd_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_fwd: null args");
    return sfa::launch_fwd(*args, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_decode(const sfa_decode_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_decode: null args");
    return sfa::decode_impl(*args, 0, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_decode_paged(const sfa_decode_paged_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_paged: null args");
    if (!args->block_table) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_paged: null block table");
    const int ps = args->page_size;
    if (ps < 16 || (ps & (ps - 1))) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_paged: page_size must be a power of two >= 16");
    if (args->num_pages <= 0) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_paged: num_pages must be positive");
    const int64_t blocks = ((int64_t)args->base.seqlen_k + ps - 1) / ps;
    if (args->block_table_stride < blocks)
        return sfa::fail(SFA_ERR_INVALID, "sfa_decode_paged: block table rows are shorter than seqlen_k / page_size");
    const sfa::PagedKV pk = {args->block_table, args->block_table_stride, ps};
    return sfa::decode_impl(args->base, 0, reinterpret_cast<cudaStream_t>(stream), nullptr, &pk);
}

int sfa_fwd_host(const sfa_fwd_host_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_fwd_host: null args");
    return sfa::fwd_host_impl(args->t, args->pieces, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_decode_host(const sfa_decode_host_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_host: null args");
    return sfa::decode_host_impl(args->t, args->pieces, reinterpret_cast<cudaStream_t>(stream));
}

int32_t sfa_fwd_host_last_pieces(void) { return sfa::host_last_plan_pieces(); }
int32_t sfa_fwd_host_trace(float* ms, int32_t capacity) { return sfa::fwd_host_trace(ms, capacity); }

int32_t sfa_fwd_host_plan(const sfa_fwd_host_args* args, int32_t* pieces, int32_t capacity) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_fwd_host_plan: null args");
    const sfa_fwd_args& t = args->t;
    const std::vector<sfa::HostPiece> plan =
        sfa::plan_fwd_host(t.batch, t.seqlen_q, t.seqlen_k, t.heads_q, t.heads_kv, t.head_dim, t.causal != 0, args->pieces);
    for (size_t i = 0; i < plan.size() && pieces && (int32_t)i < capacity; ++i) {
        const sfa::HostPiece& pc = plan[i];
        const int32_t row[6] = {pc.b, pc.hk0, pc.hkn, pc.r0, pc.r1, pc.unit};
        for (int j = 0; j < 6; ++j) pieces[6 * i + j] = row[j];
    }
    return (int32_t)plan.size();
}

int sfa_copy_2d(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width_bytes, size_t height,
                int32_t kind, void* stream) {
    if (!dst || !src) return sfa::fail(SFA_ERR_INVALID, "sfa_copy_2d: null pointer");
    if (kind != 1 && kind != 2) return sfa::fail(SFA_ERR_INVALID, "sfa_copy_2d: kind must be 1 (host to device) or 2 (device to host)");
    if (width_bytes > dst_pitch || width_bytes > src_pitch) return sfa::fail(SFA_ERR_INVALID, "sfa_copy_2d: width exceeds a pitch");
    if (width_bytes == 0 || height == 0) return SFA_OK;
    if (!sfa::current_device()) return SFA_ERR_NO_DEVICE;
    SFA_CUDA_CHECK(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width_bytes, height,
                                     kind == 1 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost,
                                     reinterpret_cast<cudaStream_t>(stream)));
    return SFA_OK;
}

int sfa_decode_num_splits(const sfa_decode_args* args) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_num_splits: null args");
    const sfa::DeviceInfo* dev = sfa::current_device();
    if (!dev) return SFA_ERR_NO_DEVICE;
    return sfa::resolve_splits(args->num_splits, args->batch, args->heads_kv * sfa::decode_head_chunks(args->heads_q, args->heads_kv), args->seqlen_k,
                               dev->sm_count);
}

size_t sfa_decode_workspace_bytes(int32_t batch, int32_t heads_q, int32_t head_dim, int32_t num_splits) {
    return sfa::decode_workspace_bytes(batch, heads_q, head_dim, num_splits);
}

int sfa_decoder_step(const sfa_decoder_step_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_decoder_step: null args");
    return sfa::decoder_step_impl(*args, reinterpret_cast<cudaStream_t>(stream));
}

size_t sfa_decoder_step_workspace_bytes(int32_t batch_size, int32_t num_heads, int32_t head_dim) {
    // rotated queries of the three-launch form + partials of the largest split count the launcher can choose
    const size_t q_bytes = sfa::align_up((size_t)batch_size * num_heads * head_dim * 2, 256);
    return q_bytes + sfa::decode_workspace_bytes(batch_size, num_heads, head_dim, 256) + 256;
}

int sfa_decoder_step_paged(const sfa_decoder_step_paged_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_decoder_step_paged: null args");
    return sfa::decoder_step_paged_impl(*args, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_decode_sharded(const sfa_decode_args* local_args, const sfa_shard_comm* comm, void* stream) {
    if (!local_args || !comm) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_sharded: null args");
    return sfa::decode_sharded_impl(*local_args, *comm, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_decode_sharded_emulated(const sfa_decode_args* args, int32_t world, void* const* ws, uint32_t step, void* stream) {
    if (!args || !ws) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_sharded_emulated: null args");
    return sfa::decode_sharded_emulated_impl(args, world, ws, step, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_shard_merge(const float* payloads, int32_t world, void* o, float* lse, int32_t batch, int32_t heads_q,
                    int32_t head_dim, int64_t o_batch_stride, int64_t o_head_stride, int32_t dtype, void* stream) {
    int rc = sfa::check_dtype(dtype);
    if (rc) return rc;
    if (!payloads || !o) return sfa::fail(SFA_ERR_INVALID, "sfa_shard_merge: null pointer");
    if (batch <= 0) return SFA_OK;
    if (!sfa::current_device()) return SFA_ERR_NO_DEVICE;
    sfa::ShardLaunch S = {};
    S.o = o; S.lse = lse; S.o_bs = o_batch_stride; S.o_hs = o_head_stride;
    S.batch = batch; S.heads_q = heads_q; S.head_dim = head_dim; S.dtype = dtype;
    S.rank = 0; S.world = world; S.payload = const_cast<float*>(payloads); S.mode = 2;
    return sfa::launch_shard_exchange(S, reinterpret_cast<cudaStream_t>(stream));
}

size_t sfa_shard_workspace_bytes(int32_t world, int32_t batch, int32_t heads_q, int32_t head_dim) {
    return sfa::shard_workspace_bytes(world, batch * heads_q, head_dim);
}

int sfa_peer_alloc(size_t bytes, void** ptr, uint8_t handle[64]) {
    if (!ptr || !handle || bytes == 0) return sfa::fail(SFA_ERR_INVALID, "sfa_peer_alloc: bad arguments");
    if (!sfa::current_device()) return SFA_ERR_NO_DEVICE;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    SFA_CUDA_CHECK(cudaMalloc(ptr, bytes));
    SFA_CUDA_CHECK(cudaMemset(*ptr, 0, bytes));
    SFA_CUDA_CHECK(cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    SFA_CUDA_CHECK(cudaIpcGetMemHandle(&h, *ptr));
    memcpy(handle, &h, 64);
    return SFA_OK;
}

int sfa_peer_open(const uint8_t handle[64], void** ptr) {
    if (!ptr || !handle) return sfa::fail(SFA_ERR_INVALID, "sfa_peer_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    SFA_CUDA_CHECK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return SFA_OK;
}

int sfa_peer_close(void* ptr) {
    if (ptr) SFA_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
    return SFA_OK;
}

int sfa_peer_free(void* ptr) {
    if (ptr) SFA_CUDA_CHECK(cudaFree(ptr));
    return SFA_OK;
}

int sfa_shard_watchdog(int clear) {
    if (!sfa::current_device()) return SFA_ERR_NO_DEVICE;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return sfa::fail(SFA_ERR_CUDA, std::string("cudaDeviceSynchronize: ") + cudaGetErrorString(e));
    int v = 0;
    int rc = sfa::shard_error_flag_read(&v, clear);
    return rc ? rc : v;
}

int sfa_rotary_table(void* cos_table, void* sin_table, int32_t max_seq_len, int32_t rot_dim, int32_t dtype, void* stream) {
    int rc = sfa::check_dtype(dtype);
    if (rc) return rc;
    if (!sfa::current_device()) return SFA_ERR_NO_DEVICE;
    return sfa::launch_rotary_table(cos_table, sin_table, max_seq_len, rot_dim, dtype, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_fill(void* array, float value, int64_t n, int32_t dtype, void* stream) {
    int rc = sfa::check_dtype(dtype);
    if (rc) return rc;
    return sfa::launch_fill(array, value, n, dtype, reinterpret_cast<cudaStream_t>(stream));
}

int sfa_tuning_set(const char* name, int32_t value) {
    if (!name) return sfa::fail(SFA_ERR_INVALID, "sfa_tuning_set: null name");
    sfa::DecodeTuning& t = sfa::decode_tuning();
    const std::string k(name);
    if (k == "decode_swap") {
#ifndef SFA_PROBES
        if (value == 0) return sfa::fail(SFA_ERR_UNSUPPORTED, "decode_swap = 0 exists in the probe build only");
#endif
        t.swap = value;
    } else if (k == "decode_cluster") t.cluster = value;
    else if (k == "decode_pdl") t.pdl = value;
    else if (k == "decode_waves") t.waves = value;
    else if (k == "decode_ctas") t.ctas_per_sm = value;
    else if (k == "decoder_fused") t.fused_step = value;
    else if (k == "host_taper") sfa::host_tuning().taper = value;
    else if (k == "host_heavy") sfa::host_tuning().heavy = value;
    else if (k == "host_taper_bytes") sfa::host_tuning().taper_bytes = value > 0 ? value : 1;
    else if (k == "host_piece_bytes") sfa::host_tuning().piece_bytes = value > 0 ? value : 1;
    else if (k == "host_trace") sfa::host_tuning().trace = value;
    else if (k == "host_kernel_streams") sfa::host_tuning().kernel_streams = value;
    else if (k == "host_wave_items") sfa::host_tuning().wave_items = value > 0 ? value : 1;
    else return sfa::fail(SFA_ERR_INVALID, "sfa_tuning_set: unknown switch " + k);
    return SFA_OK;
}

int sfa_fwd_probe_read(uint64_t* out, int32_t n) {
    if (!out || n <= 0) return sfa::fail(SFA_ERR_INVALID, "sfa_fwd_probe_read: null buffer");
    if (cudaDeviceSynchronize() != cudaSuccess) return sfa::fail(SFA_ERR_CUDA, "device synchronize failed");
    return sfa::fwd_probe_read(reinterpret_cast<unsigned long long*>(out), n);
}

int sfa_fwd_watchdog(int clear) {
    if (!sfa::current_device()) return SFA_ERR_NO_DEVICE;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return sfa::fail(SFA_ERR_CUDA, std::string("cudaDeviceSynchronize: ") + cudaGetErrorString(e));
    int v = 0, v1 = 0;
    int rc = sfa::fwd_error_flag_read(&v);
    if (rc) return rc;
#ifdef SFA_PROBES
    if ((rc = sfa::fwd_error_flag_read_v1(&v1))) return rc;
#endif
    if (clear && v) {
        // an aborted persistent kernel may have left its schedule counters behind
        if ((rc = sfa::fwd_error_flag_clear())) return rc;
        rc = sfa::fwd_sched_reset();
    }
#ifdef SFA_PROBES
    if (clear && v1) rc = sfa::fwd_error_flag_clear_v1();
#endif
    return rc ? rc : (v ? v : v1);
}

}  // extern "C"

// ============================================================================ C++ drop-ins
namespace {
template <typename T> struct DtypeOf;
template <> struct DtypeOf<half> { static constexpr int value = SFA_FP16; };
template <> struct DtypeOf<__nv_bfloat16> { static constexpr int value = SFA_BF16; };

void report(const char* where, int rc) {
    if (rc) fprintf(stderr, "[StarFlashAttention] %s failed (%d): %s\n", where, rc, sfa_last_error());
}
}  // namespace

template <typename T>
void run_flash_decoder(Flash_decoder_input& input, Flash_decoder_params& params, cudaStream_t stream) {
    sfa_decoder_step_args a = {};
    a.qkv = input.qkv; a.o = input.o; a.seq_len = reinterpret_cast<const int32_t*>(input.seq_len);
    a.k_cache_table = input.k_cache_table; a.v_cache_table = input.v_cache_table;
    a.rotary_cos_table = input.rotary_cos_table; a.rotary_sin_table = input.rotary_sin_table;
    a.lse = input.lse;
    a.batch_size = input.batch_size; a.memory_max_len = input.memory_max_len; a.num_heads = input.num_heads;
    a.head_dim = input.head_dim;
    a.head_dim_inv = input.head_dim_inv > 0.f ? input.head_dim_inv : 1.f / sqrtf((float)input.head_dim);
    a.rotary_embedding_dim = input.rotary_embedding_dim;
    a.qkv_stride = input.stride > 0 ? input.stride : 3 * input.num_heads * input.head_dim;
    a.num_layer = input.num_layer > 0 ? input.num_layer : 1; a.idx_layer = input.idx_layer;
    a.num_splits = params.num_splits;
    a.dtype = DtypeOf<T>::value;
    report("run_flash_decoder", sfa_decoder_step(&a, stream));
}
template void run_flash_decoder<half>(Flash_decoder_input&, Flash_decoder_params&, cudaStream_t);
template void run_flash_decoder<__nv_bfloat16>(Flash_decoder_input&, Flash_decoder_params&, cudaStream_t);

template <typename T>
void compute_rotary_table(T* rotary_cos_table, T* rotary_sin_table, int max_seq_len, int rot_embed_dim) {
    report("compute_rotary_table",
           sfa_rotary_table(rotary_cos_table, rotary_sin_table, max_seq_len, rot_embed_dim, DtypeOf<T>::value, nullptr));
}
template void compute_rotary_table<half>(half*, half*, int, int);
template void compute_rotary_table<__nv_bfloat16>(__nv_bfloat16*, __nv_bfloat16*, int, int);

void init_half_array(half* array, half value, int n, int /*numBlocks*/, int /*blockSize*/) {
    report("init_half_array", sfa_fill(array, __half2float(value), n, SFA_FP16, nullptr));
}

template <typename T>
void run_flash_fwd(Flash_fwd_params& p, cudaStream_t stream) {
    sfa_fwd_args a = {};
    a.q = p.q; a.k = p.k; a.v = p.v; a.o = p.o; a.lse = p.lse;
    a.batch = p.batch; a.seqlen_q = p.seqlen_q; a.seqlen_k = p.seqlen_k; a.heads_q = p.num_heads;
    a.heads_kv = p.num_heads_k > 0 ? p.num_heads_k : p.num_heads; a.head_dim = p.head_dim;
    a.q_batch_stride = p.q_batch_stride; a.q_row_stride = p.q_row_stride; a.q_head_stride = p.q_head_stride;
    a.k_batch_stride = p.k_batch_stride; a.k_row_stride = p.k_row_stride; a.k_head_stride = p.k_head_stride;
    a.v_batch_stride = p.v_batch_stride; a.v_row_stride = p.v_row_stride; a.v_head_stride = p.v_head_stride;
    a.o_batch_stride = p.o_batch_stride; a.o_row_stride = p.o_row_stride; a.o_head_stride = p.o_head_stride;
    a.softmax_scale = p.softmax_scale; a.causal = p.is_causal; a.dtype = DtypeOf<T>::value;
    report("run_flash_fwd", sfa_fwd(&a, stream));
}
template void run_flash_fwd<half>(Flash_fwd_params&, cudaStream_t);
template void run_flash_fwd<__nv_bfloat16>(Flash_fwd_params&, cudaStream_t);

template <typename T>
void run_flash_decode(Flash_decode_params& p, cudaStream_t stream) {
    sfa_decode_args a = {};
    a.q = p.q; a.k_cache = p.k_cache; a.v_cache = p.v_cache; a.seqlens = p.seqlens; a.o = p.o; a.lse = p.lse;
    a.batch = p.batch; a.seqlen_k = p.seqlen_k; a.heads_q = p.num_heads;
    a.heads_kv = p.num_heads_k > 0 ? p.num_heads_k : p.num_heads; a.head_dim = p.head_dim;
    a.q_batch_stride = p.q_batch_stride; a.q_head_stride = p.q_head_stride;
    a.k_batch_stride = p.k_batch_stride; a.k_row_stride = p.k_row_stride; a.k_head_stride = p.k_head_stride;
    a.v_batch_stride = p.v_batch_stride; a.v_row_stride = p.v_row_stride; a.v_head_stride = p.v_head_stride;
    a.o_batch_stride = p.o_batch_stride; a.o_head_stride = p.o_head_stride;
    a.softmax_scale = p.softmax_scale; a.num_splits = p.num_splits; a.dtype = DtypeOf<T>::value;
    report("run_flash_decode", sfa_decode(&a, stream));
}
template void run_flash_decode<half>(Flash_decode_params&, cudaStream_t);
template void run_flash_decode<__nv_bfloat16>(Flash_decode_params&, cudaStream_t);
