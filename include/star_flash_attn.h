/*
 * star_flash_attn.h — C ABI of the B200-native attention forward / split-KV decode library
 * (libStarFlashAttention.so).  Plain pointers and sizes only; no torch or C++ types.
 *
 * Every entry point is asynchronous on the CUDA stream it is given (a cudaStream_t passed as
 * void*), allocates nothing in steady state and never synchronises the device, so a call can
 * be captured in a CUDA graph.  Return value: 0 on success, a negative sfa_status otherwise;
 * sfa_last_error() returns a thread-local message for the last failure.
 *
 * Which reference interface each function replaces (paths relative to the reference tree):
 *   sfa_decoder_step   <- run_flash_decoder<half>(Flash_decoder_input&, Flash_decoder_params&,
 *                         cudaStream_t)                       src/flash_attn.h:7-8, src/flash_attn.cu:937-1018
 *                         and its Python face mha_fwd_cuda    src/flash_api.cpp:42-68
 *   sfa_rotary_table   <- compute_rotary_table<half>          src/flash_attn.h:9-10, src/flash_attn.cu:531-538
 *   sfa_fill           <- init_half_array                     src/flash_attn.h:11,   src/flash_attn.cu:502-510
 *   sfa_decode         <- flash_decoder_kernel + flash_combine_kernel without the RoPE/append
 *                         prologue, extended to GQA / bf16 / LSE        src/flash_attn.cu:554-935
 *   sfa_fwd            <- no reference counterpart (the reference has no prefill kernel);
 *                         softmax(QK^T*scale [+causal])V as in examples/python/testFlashDecoder.py:84-92
 *   sfa_fwd_host,      <- the launcher's own host<->device traffic (blocking cudaMemcpy + cudaDeviceSynchronize per
 *   sfa_decode_host       call)                                          src/flash_attn.cu:973-975,986
 *   sfa_decode_static  <- no reference counterpart: a decode entry that allocates
 *                         nothing
 * The C++-mangled drop-in symbols (run_flash_decoder<half>, compute_rotary_table<half>,
 * init_half_array) are exported from the same library; see include/src/flash_attn.h.
 */
#ifndef STAR_FLASH_ATTN_H_
#define STAR_FLASH_ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SFA_ABI_VERSION 2

typedef enum sfa_dtype { SFA_FP16 = 0, SFA_BF16 = 1 } sfa_dtype;

typedef enum sfa_status {
    SFA_OK = 0,
    SFA_ERR_INVALID = -1,     /* bad argument (shape, alignment, dtype, null pointer)        */
    SFA_ERR_UNSUPPORTED = -2, /* valid request this build has no kernel for (e.g. head_dim)  */
    SFA_ERR_CUDA = -3,        /* a CUDA runtime/driver call failed; see sfa_last_error()     */
    SFA_ERR_NO_DEVICE = -4    /* no sm_100 device: there is no CPU fallback                  */
} sfa_status;

/* ---------------------------------------------------------------- prefill forward ----------
 * O = softmax(Q K^T * softmax_scale [+ causal mask]) V,  LSE = logsumexp of the scaled scores.
 * q (batch, seqlen_q, heads_q, head_dim), k/v (batch, seqlen_k, heads_kv, head_dim), o like q;
 * strides are in ELEMENTS; the head_dim axis is contiguous; every stride must be a multiple of
 * 8 elements (16 bytes) and base pointers 16-byte aligned.  lse (batch, heads_q, seqlen_q)
 * fp32 contiguous, may be NULL.  heads_q % heads_kv == 0 (GQA).  head_dim in {64, 128}.
 * Causal mask is bottom-right aligned (query i sees keys j <= i + seqlen_k - seqlen_q).
 * softmax_scale <= 0 selects 1/sqrt(head_dim). */
typedef struct sfa_fwd_args {
    const void* q;
    const void* k;
    const void* v;
    void* o;
    float* lse;
    int32_t batch, seqlen_q, seqlen_k, heads_q, heads_kv, head_dim;
    int64_t q_batch_stride, q_row_stride, q_head_stride;
    int64_t k_batch_stride, k_row_stride, k_head_stride;
    int64_t v_batch_stride, v_row_stride, v_head_stride;
    int64_t o_batch_stride, o_row_stride, o_head_stride;
    float softmax_scale;
    int32_t causal;
    int32_t dtype; /* sfa_dtype */
} sfa_fwd_args;

int sfa_fwd(const sfa_fwd_args* args, void* stream);

/* ---------------------------------------------------------------- split-KV decode ----------
 * One query token per sequence.  q (batch, heads_q, head_dim); k_cache/v_cache
 * (batch, max_len, heads_kv, head_dim) addressed through element strides; seqlens int32
 * (batch) = number of visible keys per sequence (NULL => seqlen_k for all); o like q;
 * lse (batch, heads_q) fp32 or NULL.  heads_q % heads_kv == 0; a GQA group of up to 8 query heads
 * shares one pass over K/V, a larger group (MQA with 16, 32 ... heads) is served in equal chunks of <= 8.
 * num_splits <= 0 lets the launcher choose.  workspace: device scratch of at least
 * sfa_decode_workspace_bytes(...) bytes, or NULL to use the library's per-device grow-only
 * buffer (grown with cudaMalloc outside graph capture only). */
typedef struct sfa_decode_args {
    const void* q;
    const void* k_cache;
    const void* v_cache;
    const int32_t* seqlens;
    void* o;
    float* lse;
    int32_t batch, seqlen_k, heads_q, heads_kv, head_dim;
    int64_t q_batch_stride, q_head_stride;
    int64_t k_batch_stride, k_row_stride, k_head_stride;
    int64_t v_batch_stride, v_row_stride, v_head_stride;
    int64_t o_batch_stride, o_head_stride;
    float softmax_scale;
    int32_t num_splits;
    int32_t dtype; /* sfa_dtype */
    void* workspace;
    size_t workspace_bytes;
} sfa_decode_args;

int sfa_decode(const sfa_decode_args* args, void* stream);
/* Number of splits the launcher would use for this shape (args->num_splits honoured if > 0). */
int sfa_decode_num_splits(const sfa_decode_args* args);
size_t sfa_decode_workspace_bytes(int32_t batch, int32_t heads_q, int32_t head_dim, int32_t num_splits);

/* Split-KV decode that allocates nothing.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid allocation).  Same arguments, arithmetic and results as sfa_decode (args->workspace and
 * args->workspace_bytes are ignored); the per-split partials live in ONE static device array of
 * SFA_DECODE_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit, so the call makes no
 * cudaMalloc / cudaFree and calls no allocating library routine.  Limits: a launch whose partials
 * (sfa_decode_workspace_bytes(batch, heads_q, head_dim, sfa_decode_num_splits(args)) bytes) exceed the array returns
 * SFA_ERR_INVALID with the size it needs - lower num_splits or the batch, or use sfa_decode with a workspace.  Every
 * call on a device shares the array: calls that OVERLAP on the GPU (two streams, concurrently replayed graphs) race
 * on it; calls ordered on one stream do not.  The array keeps the last call's contents; nothing in it needs to start
 * at zero (every split writes its own partial before the combine kernel reads it). */
#define SFA_DECODE_STATIC_SCRATCH_BYTES (16u << 20)
int sfa_decode_static(const sfa_decode_args* args, void* stream);
size_t sfa_decode_static_scratch_bytes(void);

/* ---------------------------------------------------------------- split-KV decode over a paged cache ---
 * Row f1 of the scope table ("a paged/strided KV-cache layout"); the reference documents one
 * dense (batch, layer, max_len, heads, head_dim) cache (src/params.h:22-25) and has no paged form.
 * `base.k_cache` / `base.v_cache` are page POOLS (num_pages, page_size, heads_kv, head_dim) whose
 * `*_batch_stride` fields carry the PAGE stride in elements; block_table (batch, block_table_stride) int32 names
 * the pool page that holds keys [i * page_size, (i + 1) * page_size) of sequence b at [b][i]; page_size is a
 * power of two >= 16; `base.seqlen_k` is the largest addressable length (at most block_table columns * page_size)
 * and `base.seqlens` the visible length per sequence.  Same kernel, same split/merge arithmetic as sfa_decode:
 * with equal num_splits the result is bit-identical to sfa_decode over the gathered cache. */
typedef struct sfa_decode_paged_args {
    sfa_decode_args base;
    const int32_t* block_table;
    int64_t block_table_stride;
    int32_t page_size;
    int32_t num_pages; /* pool size; > 0.  Table entries are trusted to lie in [0, num_pages). */
} sfa_decode_paged_args;

int sfa_decode_paged(const sfa_decode_paged_args* args, void* stream);

/* ---------------------------------------------------------------- cross-GPU split-KV decode ---
 * One long sequence whose KV length is sharded over the GPUs of a box (one process per GPU).  Every rank
 * calls sfa_decode_sharded with ITS shard as k_cache/v_cache (seqlen_k = shard length, seqlens = number of
 * visible keys inside the shard) and the same q; the split-KV kernel runs over the shard, then one kernel
 * merges the rank's splits, stores (O_r, LSE_r) into every peer's workspace over NVLink (P2P stores into
 * CUDA-IPC mapped buffers), waits for the peers' rows and merges them: every rank ends with the full O, LSE.
 * No reference counterpart (the reference is single-GPU); the merge is flash_combine_kernel's
 * (src/flash_attn.cu:877-935) applied across devices.
 *   mode 0: fused peer exchange (needs peer_ws[0..world), peer_ws[rank] = own workspace from sfa_peer_alloc,
 *           the others from sfa_peer_open; `step` must increase by one per call, starting at 1).
 *   mode 1: publish only: (rows, head_dim + 1) fp32 to `payload` (rows = batch * heads_q; O rows, then the
 *           LSEs in log2 units) for an external all-gather, to be merged by sfa_shard_merge. */
typedef struct sfa_shard_comm {
    int32_t rank, world;
    uint32_t step;
    int32_t mode;
    void* peer_ws[8];
    float* payload;
} sfa_shard_comm;

int sfa_decode_sharded(const sfa_decode_args* local_args, const sfa_shard_comm* comm, void* stream);
/* Test aid: `world` ranks emulated on ONE GPU.  args[r] is rank r's call (each with its own split workspace in
 * args[r].workspace), ws[r] its exchange workspace (plain device memory of sfa_shard_workspace_bytes).  The
 * local passes run back to back; the exchange blocks of all ranks go out as one cooperative launch, because
 * separate launches that spin on each other's flags are not guaranteed to run at the same time on one GPU.
 * Synchronises the stream. */
int sfa_decode_sharded_emulated(const sfa_decode_args* args, int32_t world, void* const* ws, uint32_t step, void* stream);
/* Merges `world` gathered payloads laid out (world, rows, head_dim + 1) into o / lse. */
int sfa_shard_merge(const float* payloads, int32_t world, void* o, float* lse, int32_t batch, int32_t heads_q,
                    int32_t head_dim, int64_t o_batch_stride, int64_t o_head_stride, int32_t dtype, void* stream);
size_t sfa_shard_workspace_bytes(int32_t world, int32_t batch, int32_t heads_q, int32_t head_dim);
/* Peer-visible device buffer: cudaMalloc + zero fill + CUDA IPC handle (64 bytes) to pass to the other
 * processes, which map it with sfa_peer_open.  Close mappings before the owner frees the buffer. */
int sfa_peer_alloc(size_t bytes, void** ptr, uint8_t handle[64]);
int sfa_peer_open(const uint8_t handle[64], void** ptr);
int sfa_peer_close(void* ptr);
int sfa_peer_free(void* ptr);
/* Non-zero after a rank waited more than 2 s for a peer's rows (1 + the missing rank); synchronises. */
int sfa_shard_watchdog(int clear);

/* ---------------------------------------------------------------- drop-in decoder step -----
 * The reference's run_flash_decoder with plain arguments.  For every sequence b:
 * rotate q and k of the new token (interleaved pairs, position seq_len[b]) over the first
 * rotary_dim dims, append k,v to layer idx_layer of the caches at row seq_len[b], and attend
 * q over the seq_len[b]+1 cached keys with scale head_dim_inv.
 * qkv (batch, 3, heads, head_dim) contiguous per token with row stride `qkv_stride` elements;
 * caches (batch, num_layer, memory_max_len, heads, head_dim) contiguous; seq_len int32 (batch),
 * each < memory_max_len; o (batch, heads, head_dim).  rotary_cos/sin: (memory_max_len,
 * rotary_dim/2) tables in the data dtype, or both NULL to evaluate the angles in fp32.
 * q/k/v biases are accepted and ignored, as in the reference (src/flash_api.cpp:18-32). */
typedef struct sfa_decoder_step_args {
    const void* qkv;
    void* o;
    const int32_t* seq_len;
    void* k_cache_table;
    void* v_cache_table;
    const void* rotary_cos_table;
    const void* rotary_sin_table;
    float* lse; /* optional (batch, heads) */
    int32_t batch_size, memory_max_len, num_heads, head_dim;
    float head_dim_inv;
    int32_t rotary_embedding_dim, qkv_stride, num_layer, idx_layer;
    int32_t num_splits; /* <= 0: launcher's choice */
    int32_t dtype;      /* sfa_dtype; the reference is fp16-only */
    /* Optional caller-owned device scratch for the split partials (at least sfa_decoder_step_workspace_bytes);
     * NULL: the library's scratch of this (device, stream), grown outside graph capture only.  A CUDA graph that
     * captures the step should own its scratch. */
    void* workspace;
    size_t workspace_bytes;
} sfa_decoder_step_args;

int sfa_decoder_step(const sfa_decoder_step_args* args, void* stream);
/* Scratch the step may need for this shape (an upper bound over the split counts the launcher can choose). */
size_t sfa_decoder_step_workspace_bytes(int32_t batch_size, int32_t num_heads, int32_t head_dim);

/* ---------------------------------------------------------------- decoder step over a paged cache ---
 * The same fused step (RoPE of q and k at position seq_len[b], append of k, v at row seq_len[b], attention over
 * seq_len[b] + 1 keys; intended by src/flash_attn.cu:591-601,696-701) with GQA and with the cache held as page
 * pools (num_pages, page_size, heads_kv, head_dim) behind a block table (see sfa_decode_paged).  The page of row
 * seq_len[b], block_table[b][seq_len[b] / page_size], must already be assigned by the caller; a sequence with
 * seq_len[b] >= max_len appends nothing and attends over its max_len cached keys.
 * qkv: per token (heads_q + 2 * heads_kv, head_dim) - the q heads, then the k heads, then the v heads - with row
 * stride `qkv_stride` elements.  rotary tables (max_len, rotary_dim / 2) in the data dtype, or both NULL. */
typedef struct sfa_decoder_step_paged_args {
    const void* qkv;
    void* o;           /* (batch, heads_q, head_dim) contiguous */
    float* lse;        /* optional (batch, heads_q) */
    const int32_t* seq_len;
    void* k_pool;
    void* v_pool;
    const int32_t* block_table;
    const void* rotary_cos_table;
    const void* rotary_sin_table;
    int64_t page_stride, row_stride, head_stride; /* of both pools, in elements */
    int64_t block_table_stride;
    int64_t qkv_stride;
    int32_t batch_size, max_len, heads_q, heads_kv, head_dim, page_size, num_pages;
    int32_t rotary_embedding_dim;
    float softmax_scale; /* <= 0: 1 / sqrt(head_dim) */
    int32_t num_splits;  /* <= 0: launcher's choice */
    int32_t dtype;
} sfa_decoder_step_paged_args;

int sfa_decoder_step_paged(const sfa_decoder_step_paged_args* args, void* stream);

/* cos/sin tables (max_seq_len, rot_dim/2): angle = pos / 10000^(2i/rot_dim). */
int sfa_rotary_table(void* cos_table, void* sin_table, int32_t max_seq_len, int32_t rot_dim,
                     int32_t dtype, void* stream);
/* Fill n elements with a constant (the reference's test utility). */
int sfa_fill(void* array, float value, int64_t n, int32_t dtype, void* stream);

/* ---------------------------------------------------------------- host-buffer entry points ---
 * The same operations on PINNED HOST tensors (cudaHostAlloc / cudaHostRegister): inputs are uploaded, the
 * kernels run and the results are downloaded by one call, cut into pieces so that the upload of piece i+1, the
 * kernel of piece i and the download of piece i-1 overlap on three streams (two are owned by the library, the
 * kernels run on `stream`).  The results are complete once `stream` has been synchronised.  The reference's own
 * host path is blocking cudaMemcpy calls and a device synchronise per call (src/flash_attn.cu:973-975,986).
 * Every pointer of `t` is a host pointer; strides are in elements and describe the HOST tensors (a caller may pass a
 * slice of the heads of a larger tensor: head stride must equal head_dim, batch / row strides are free); lse is
 * (batch, heads_q, seqlen_q) resp. (batch, heads_q) fp32 contiguous or NULL.  pieces <= 0: library's choice
 * (prefill: about 8 large (batch entry, kv-head group, block of query rows) pieces, plus geometrically smaller ones
 * at the end of the call when the kernels are short beside the copies; 8 for decode: batch ranges); pieces > 0: that
 * many uniform pieces.  Device staging is owned by the library (grow-only rings, per device; released by
 * sfa_release_workspaces).  Not graph-capturable. */
typedef struct sfa_fwd_host_args {
    sfa_fwd_args t;
    int32_t pieces;
} sfa_fwd_host_args;
int sfa_fwd_host(const sfa_fwd_host_args* args, void* stream);

typedef struct sfa_decode_host_args {
    sfa_decode_args t; /* workspace fields are ignored */
    int32_t pieces;
} sfa_decode_host_args;
int sfa_decode_host(const sfa_decode_host_args* args, void* stream);
/* The piece plan sfa_fwd_host would follow for `args` (only the dimensions, `causal` and `pieces` are read; no device
 * is needed): writes up to `capacity` pieces of 6 ints each - batch entry, first kv-head, kv-heads, first row, end
 * row, unit (pieces of one unit share its K/V staging) - in issue order and returns the number of pieces of the plan. */
int32_t sfa_fwd_host_plan(const sfa_fwd_host_args* args, int32_t* pieces, int32_t capacity);
/* Diagnostics: number of pieces the most recent sfa_fwd_host call of this process was cut into. */
int32_t sfa_fwd_host_last_pieces(void);
/* Diagnostics (after sfa_tuning_set("host_trace", 1)): synchronises the device and writes, for up to `capacity`
 * pieces of the last sfa_fwd_host call on the current device, 5 floats - milliseconds from the first upload's start
 * to the piece's upload start, upload end, kernel start, kernel end, download end; returns the number of pieces. */
int32_t sfa_fwd_host_trace(float* ms, int32_t capacity);

/* ---------------------------------------------------------------- misc ---------------------- */
/* Pitched copy between pinned host memory and the device on `stream` (cudaMemcpy2DAsync): `height` rows of
 * `width_bytes`, row pitches in bytes; kind 1 = host to device, 2 = device to host.  Used by the host-buffer entry
 * points when a chunk is a slice of heads (rows of a (batch, seqlen, heads, head_dim) tensor are then not contiguous);
 * the reference's own host path is blocking cudaMemcpy calls (src/flash_attn.cu:973-975). */
int sfa_copy_2d(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width_bytes, size_t height,
                int32_t kind, void* stream);

const char* sfa_last_error(void);
int sfa_abi_version(void);
/* Number of kernels this library has launched in this process (all entry points). */
uint64_t sfa_launch_count(void);
/* Pipeline watchdog of the prefill kernel.  Its mbarrier waits are bounded (2 s); a wait that times
 * out ends the kernel and latches a non-zero code here.  Synchronises the device; returns the code
 * (0 = healthy) or a negative sfa_status, and clears the latch when `clear` is non-zero. */
int sfa_fwd_watchdog(int clear);
/* Tuning aid: cycle counters written by the instrumented prefill build (SFA_FWD_DBG=5, tools/fwd_probe.py);
 * copies min(n, 32) 64-bit counters after synchronising the device. */
int sfa_fwd_probe_read(uint64_t* out, int32_t n);
/* Test / tuning aid: sets one of the library's A-B switches for the rest of the process (they start from the
 * SFA_* environment variables, read once): "decode_cluster" (1: 2..8 splits merge in a thread-block cluster; 0:
 * partials + combine launch), "decode_pdl" (0..2), "decode_waves", "decode_ctas", "decoder_fused" (1: RoPE +
 * append inside the decode kernel; 0: stand-alone kernel first), "decode_swap" (0 in the probe build only),
 * "host_taper" (0: sfa_fwd_host cuts uniform pieces only), "host_heavy" (-1 auto / 0 / 1: plan the call as one whose
 * kernels are short / as long as its copies), "host_taper_bytes" (smallest refined piece, default 4 MiB),
 * "host_piece_bytes" (the default plan has at most input bytes / this many pieces, default 8 MiB), "host_trace",
 * "host_kernel_streams" (1 .. 4 kernel streams the pieces take turns on, the first being `stream`; default 2),
 * "host_wave_items" (work items per piece of a heavy call, default 128).
 * None of them changes the arithmetic contract; not thread-safe against concurrent launches. */
int sfa_tuning_set(const char* name, int32_t value);
/* Releases the per-device workspaces (synchronises the device). */
void sfa_release_workspaces(void);

#ifdef __cplusplus
}
#endif
#endif /* STAR_FLASH_ATTN_H_ */
