// Internal host-side contracts between the translation units of libStarFlashAttention.so.
#pragma once
#include <vector>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stddef.h>
#include <string>

#include "../../include/star_flash_attn.h"

namespace sfa {

// Thread-local error string behind sfa_last_error().
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);
void count_launch(int n = 1);

struct DeviceInfo {
    int device = -1;
    int sm_count = 0;
    int cc_major = 0, cc_minor = 0;
};
// Cached per-device properties; returns nullptr (and sets the error) when no sm_100 device is current.
const DeviceInfo* current_device();

// Grow-only scratch of one (device, stream); outgrown buffers are retired, never freed under a caller's feet
// (freed by sfa_release_workspaces only); grown only outside stream capture.
void* get_workspace(size_t bytes, cudaStream_t stream);

#define SFA_CUDA_CHECK(expr)                                                                  \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
            return ::sfa::fail(SFA_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
    } while (0)

// Tuning / A-B switches of the decode path.  Initialised from the environment ONCE per process (SFA_DECODE_SWAP,
// SFA_DECODE_CLUSTER, SFA_DECODE_PDL, SFA_DECODE_WAVES, SFA_DECODE_CTAS, SFA_DECODER_FUSED); tests and tuning tools
// change them through sfa_tuning_set().  Nothing on a launch path calls getenv.
struct DecodeTuning {
    int swap = 1;        // transposed MMAs (the product form); 0 = untransposed, in the -DSFA_PROBES build only
    int cluster = 1;     // 2 .. 8 splits merge inside a thread-block cluster; 0 = always partials + combine launch
    int pdl = 2;         // programmatic dependent launch: 0 off, 1 combine kernel only, 2 every decode kernel
    int waves = 4;       // rounds of CTAs the two-launch split count aims at
    int ctas_per_sm = 3;
    int fused_step = 1;  // decoder step: RoPE + append inside the decode kernel; 0 = stand-alone kernel first
};
DecodeTuning& decode_tuning();

// ------------------------------------------------------------------ decode (decode_sm100.cu)
struct DecodeLaunch {
    const void* q; const void* k; const void* v; const int32_t* seqlens;
    void* o; float* lse;
    int batch, seqlen_k, heads_q, heads_kv, head_dim;
    int64_t q_bs, q_hs, k_bs, k_rs, k_hs, v_bs, v_rs, v_hs, o_bs, o_hs;
    float scale;
    int seqlen_add;          // added to every sequence length (1 for the decoder step: the appended token)
    int num_splits;          // resolved (>= 1)
    float* o_part;           // (B,Hq,S,D) fp32   (unused when num_splits == 1)
    float2* ml_part;         // (B,Hq,S)  (m in log2 domain, l)
    int dtype;
    // Fused decoder step (nullptr = plain decode): q/k/v of the new token are read from fused_qkv
    // ((batch, heads_q + 2 heads_kv, head_dim) rows of qkv_stride elements), q and k are rotated at position
    // seqlens[b] and k, v are appended to the cache row seqlens[b] by the decode kernel itself.
    const void* fused_qkv = nullptr;
    const void* cos_table = nullptr; const void* sin_table = nullptr;
    int64_t qkv_stride = 0;
    int rotary_dim = 0;
    bool force_partials = false;   // write (o_part, ml_part) even with one split (cross-GPU sharded decode)
    // Paged cache (nullptr = contiguous): k / v are page pools, k_bs / v_bs the page strides, block_table
    // (batch, bt_stride) int32 maps the i-th page_size-key block of a sequence to its pool page.
    const int32_t* block_table = nullptr;
    int64_t bt_stride = 0;
    int page_size = 0;
};
// CTA rows per kv head: 1 while the GQA group has at most 8 query heads, else the smallest count of equal chunks of <= 8.
int decode_head_chunks(int heads_q, int heads_kv);
// `heads_kv` counts head chunks (heads_kv * decode_head_chunks(...)).
int choose_decode_splits(int batch, int heads_kv, int max_seqlen, int sm_count);
size_t decode_workspace_bytes(int batch, int heads_q, int head_dim, int num_splits);
int launch_decode(const DecodeLaunch& p, cudaStream_t stream);

// ------------------------------------------------------------------ cross-GPU split-KV decode (shard_sm100.cu)
struct ShardLaunch {
    const float* o_part; const float2* ml_part; int num_splits;   // this rank's split partials (modes 0, 1)
    void* o; float* lse; int64_t o_bs, o_hs;
    int batch, heads_q, head_dim, dtype;
    int rank, world;
    uint32_t step;
    void* const* peer_ws;   // mode 0: world pointers, peer_ws[rank] = own workspace
    float* payload;         // mode 1: (rows, D + 1) out; mode 2: (world, rows, D + 1) in
    int mode;               // 0 fused peer exchange + merge, 1 publish only, 2 merge gathered payloads
};
size_t shard_workspace_bytes(int world, int rows, int head_dim);
int launch_shard_exchange(const ShardLaunch& a, cudaStream_t stream);
// All ranks' exchange blocks in one cooperative launch on one GPU (tests; synchronises the stream).
int launch_shard_exchange_emulated(const ShardLaunch* ranks, int world, cudaStream_t stream);
int shard_error_flag_read(int* value, int clear);

// ------------------------------------------------------------------ rope / cache (rope_cache.cu)
struct RopeAppendLaunch {
    const void* qkv; int qkv_stride;
    void* q_rot;                       // (B,H,D) rotated queries
    void* k_cache; void* v_cache;      // (B,L,M,H,D)
    const int32_t* seq_len;
    const void* cos_table; const void* sin_table;
    int batch, heads, head_dim, rotary_dim, max_len, num_layer, idx_layer;
    int dtype;
};
int launch_rope_append(const RopeAppendLaunch& p, cudaStream_t stream);
int launch_rotary_table(void* cos_t, void* sin_t, int max_seq_len, int rot_dim, int dtype, cudaStream_t stream);
int launch_fill(void* arr, float value, int64_t n, int dtype, cudaStream_t stream);

// ------------------------------------------------------------------ prefill (fwd_sm100.cu)
int launch_fwd(const sfa_fwd_args& a, cudaStream_t stream);
int launch_fwd_v1(const sfa_fwd_args& a, cudaStream_t stream);   // round-1 kernel (one CTA per query block), A/B only
int fwd_error_flag_read_v1(int* value);
int fwd_error_flag_clear_v1();
// Tuning overrides of the prefill launcher, read from the environment once per process.
struct FwdTuning {
    int poly = -1, softrel = -1, stagger = -1, nq = 0, v1 = 0, group = 0, delay = -1, defer = -1, poll = 0, tail = -1;
};
const FwdTuning& fwd_tuning();
// Two zeroed ints (next-item counter, CTAs done) for one launch of the persistent prefill kernel; the kernel
// leaves them zero again.  Slots rotate per launch, launches under stream capture get slots of their own.
int* fwd_sched_counters(cudaStream_t stream);
// Rezeroes every slot (after a watchdog abort left counters behind).
int fwd_sched_reset();
// Pipeline watchdog of the prefill kernel: non-zero after an mbarrier wait timed out.
int fwd_error_flag_read(int* value);
int fwd_error_flag_clear();
int fwd_probe_read(unsigned long long* out, int n);

// ------------------------------------------------------------------ host-buffer pipeline (host_pipeline.cu)
int fwd_host_impl(const sfa_fwd_args& a, int want_pieces, cudaStream_t stream);
int decode_host_impl(const sfa_decode_args& a, int want_pieces, cudaStream_t stream);
void release_host_pipelines();
// Plan switches of sfa_fwd_host (sfa_tuning_set "host_taper", "host_heavy", "host_taper_bytes", "host_piece_bytes",
// "host_trace"; tests lower the byte thresholds so that small shapes go through the tapered plan).
struct HostTuning {
    int taper = 1;                 // 0: uniform pieces only
    int heavy = -1;                // -1: by the estimated kernel / copy ratio; 0 / 1: plan as a light / heavy call
    int taper_bytes = 4 << 20;     // a refined piece keeps at least this much input
    int piece_bytes = 8 << 20;     // the default plan has at most input bytes / piece_bytes pieces (and at least one per batch entry)
    int wave_items = 128;          // heavy calls: work items (256-row blocks x query heads) per piece - one wave of the 148 CTAs
    int kernel_streams = 2;        // kernel streams the pieces take turns on (1 .. 4; the first is the caller's)
    int trace = 0;                 // 1: timing events around every piece's upload, kernel and download (fwd_host_trace)
};
HostTuning& host_tuning();
// One piece of the sfa_fwd_host plan: rows [r0, r1) of kv-heads [hk0, hk0 + hkn) of batch entry b; pieces of one
// unit (b, head group) are consecutive and share the unit's K/V staging.  Pure host logic (no device needed).
struct HostPiece { int b, hk0, hkn, r0, r1, unit; };
std::vector<HostPiece> plan_fwd_host(int B, int Sq, int Sk, int Hq, int Hkv, int D, bool causal, int want_pieces);
int host_last_plan_pieces();
int fwd_host_trace(float* out, int capacity);

}  // namespace sfa
