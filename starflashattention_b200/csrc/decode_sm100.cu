// Split-KV decode attention (one query token per sequence) for sm_100a.
//
// Replaces the reference's flash_decoder_kernel / flash_combine_kernel pair
// (reference src/flash_attn.cu:554-875 and :877-935), extended to GQA, bf16, ragged sequence lengths and
// an LSE output.  The FUSED instantiation is the reference's whole decoder step in one kernel, as its
// flash_decoder_kernel intends (:591-601 q/k RoPE at position seq_len[b], :696-701 cache append): the new
// token's q, k, v are read from the packed qkv row, q and k are rotated in registers, and the one warp whose
// 16-key step holds the appended position feeds its fragments from qkv and writes k, v to the cache.  HBM-bound: every K and V byte is read
// exactly once, straight from global memory into tensor-core fragments, so there is no shared
// memory staging on the streaming path at all.
//
// Work decomposition: grid (num_splits, heads_kv, batch), 4 warps per CTA.  A CTA owns one
// contiguous chunk of one kv-head's keys and serves all heads_q/heads_kv query heads of its GQA
// group, so K/V are read once per group (a group of more than 8 query heads - MQA with 16 or 32 heads - is cut
// into equal chunks of <= 8, one CTA row each: grid.y = heads_kv * chunks, K/V read once per chunk).
// Inside the CTA the warps take 16-key steps round robin;
// each warp keeps a private online-softmax state (m, l, O) and the four states are merged through
// shared memory at the end.
//
// Fragment trick: mma.sync.m16n8k16 fed straight from 16-byte global loads.  A dot product does not care about
// the order of its terms, so the head_dim axis of Q.K is permuted such that each thread's K fragment is one
// contiguous 16-byte load, and the head_dim axis of the P.V output is permuted such that each thread's V fragment
// comes from 16-byte loads too (byte permutes pair up the keys).  The MMAs run TRANSPOSED (default, SWAP):
// S^T = K Q^T puts 16 keys on the M rows and the (up to 8) query heads of the group on the 8 columns, and
// O^T += V^T P^T puts 16 head_dim rows on M - every accumulator register is live, against half of them being the
// zero padding of a 16-row query operand in the untransposed form (kept as SFA_DECODE_SWAP=0).  P changes from
// accumulator to B-operand layout with one movmatrix per 8 keys.  Half the MMAs, 32 instead of 64 output
// registers, 128 registers per thread, 4 CTAs per SM: C4 went from 5.77 to 6.49 TB/s.
#include "internal.h"
#include "merge.cuh"
#include "rope.cuh"

#include <cooperative_groups.h>
#include <math.h>
#include <type_traits>

namespace sfa {
namespace {

constexpr int kWarps = 4;
constexpr int kStepKeys = 16;
constexpr bool kDecodeSwapDefault = true;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

__device__ __forceinline__ uint4 ld_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

template <typename T>
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                          uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma_16816<__half>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                  uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma_16816<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                         uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <typename T>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ uint32_t u4_word(const uint4& v, int i) {
    return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w));
}

struct DecodeParams {
    const void* q;
    const void* k;
    const void* v;
    const int32_t* seqlens;
    void* o;
    float* lse;
    float* o_part;
    float2* ml_part;
    int64_t q_bs, q_hs, k_bs, k_rs, k_hs, v_bs, v_rs, v_hs, o_bs, o_hs;
    int seqlen_k, seqlen_add, heads_q, group, num_splits;
    float scale_log2;
    // fused decoder step (FUSED kernels only): the new token's q/k/v come from `qkv`, q and k are rotated at
    // position seqlens[b], and the CTA whose chunk ends at that position appends k, v to the cache
    const void* qkv;
    const void* cos_table;
    const void* sin_table;
    int64_t qkv_stride;
    int rotary_dim, heads_kv;
    int kv_chunks;        // CTAs along grid.y per kv head: a GQA group of more than 8 query heads is served in chunks of `group`
    int force_partials;
    // paged cache (PAGED kernels only): k / v are page pools (pages, page_size, heads_kv, D) with k_bs / v_bs the
    // page stride; block_table[b * bt_stride + i] is the pool page holding keys [i, i + 1) << page_shift of sequence b
    const int32_t* block_table;
    int64_t bt_stride;
    int page_shift;
};

// Kernel variant: decode_tuning().swap == 0 selects the untransposed MMAs, which exist in the -DSFA_PROBES build only
// (a rejected variant kept for A/B runs: 3 CTAs per SM, zero-padded 16-row query operand).
inline bool decode_swap_enabled() {
#ifdef SFA_PROBES
    return decode_tuning().swap != 0;
#else
    return true;
#endif
}

// Shared memory used only for the end-of-CTA merge of the four warp states.
template <int D>
struct MergeSmem {
    float o[kWarps][8][D];
    float m[kWarps][8];
    float l[kWarps][8];
    float M[8], L[8];   // CLUSTER: the CTA's merged (m, l) per head, read by the other CTAs of the cluster
};

__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t x) {   // transpose of an 8x8 b16 matrix held one row-pair per thread
    uint32_t y;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(y) : "r"(x));
    return y;
}

// SWAP: the MMAs run transposed - S^T = K Q^T (16 keys x 8 heads per MMA) and O^T += V^T P^T (16 head_dim rows x 8
// heads) - so that every accumulator register is a live (key|dim, head) entry instead of half of them being the
// zero padding of the 16-row query operand: half the MMAs, 32 instead of 64 output registers, 4 CTAs per SM.
// PAGED: the cache is a pool of pages found through a per-sequence block table.  A 16-key step never straddles a
// page (page_size is a power of two >= 16 and chunks start on multiples of 64), so the indirection is one
// warp-uniform table read per step; everything after the row pointer is the same streaming code.
// CLUSTER: the num_splits (2 .. 8) CTAs of one (batch, kv-head) form a thread-block cluster and merge their
// partials through distributed shared memory - no partials in global memory and no combine launch.  The combine
// kernel costs ~5 us on its own, a sixth of a decoder step over a few thousand keys and half of one over a few
// hundred (the reference's own example shapes).
template <typename T, int D, bool FUSED, bool SWAP, bool PAGED = false, bool CLUSTER = false>
__global__ void __launch_bounds__(kWarps * 32, SWAP ? 4 : 3) decode_splitkv_kernel(const DecodeParams p) {
    constexpr int NJ = D / 32;   // 32-wide head_dim blocks of the QK^T contraction (2 MMAs each)
    constexpr int NH = D / 64;   // 64-wide head_dim blocks of the P.V output (8 n-tiles each)
    __shared__ MergeSmem<D> sm;
    asm volatile("griddepcontrol.wait;" ::: "memory");   // the kernel before this one has completed (no-op without PDL)

    const int split = blockIdx.x, hk = blockIdx.y, b = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = lane >> 2, t = lane & 3;
    const int g = p.group;
    // grid.y counts chunks of <= 8 query heads; chunk hk serves query heads [hk * g, hk * g + g) and reads kv head hkv
    const int hkv = p.kv_chunks == 1 ? hk : hk / p.kv_chunks;

    const int pos = p.seqlens ? p.seqlens[b] : p.seqlen_k;   // FUSED: row the new token is appended at
    int len = pos + p.seqlen_add;
    len = max(0, min(len, p.seqlen_k + p.seqlen_add));   // never read past the cache
    // FUSED: the appended token exists only while the cache has room (as in the stand-alone append kernel)
    const bool has_new = FUSED && pos >= 0 && pos < p.seqlen_k + p.seqlen_add;
    // row of the rotary tables: they hold seqlen_k + seqlen_add (= memory_max_len) rows; a full cache (pos ==
    // max_len, nothing is appended) or a garbage length must not index past them
    const int pos_rope = min(max(pos, 0), p.seqlen_k + p.seqlen_add - 1);
    // Per-sequence chunking keeps ragged batches balanced across the splits.
    const int chunk = (((len + p.num_splits - 1) / p.num_splits) + 63) & ~63;
    const int k_begin = split * chunk;
    const int k_end = min(len, k_begin + chunk);

    // The new token of the fused step: (q | k | v) planes of heads_q, heads_kv, heads_kv heads.  Its 16-key
    // step is the last one of the last non-empty split; that step is peeled off the streaming loop so that
    // the loop itself is the plain decode loop (no extra live registers, no extra branch).
    const int new_base = k_begin + ((pos - k_begin) & ~(kStepKeys - 1));
    const bool cta_has_new = has_new && pos >= k_begin && pos < k_end;
    const int k_end_main = cta_has_new ? new_base : k_end;
    auto token_row = [&](int plane_head) {   // element pointer of head `plane_head` of the packed qkv row
        return reinterpret_cast<const T*>(p.qkv) + (int64_t)b * p.qkv_stride + (int64_t)plane_head * D;
    };

    // ---- query fragments (A operand), rows >= g are zero padding
    uint32_t qa[NJ][4];
    {
        const T* qrow = FUSED ? token_row(hk * g + r)
                              : reinterpret_cast<const T*>(p.q) + (int64_t)b * p.q_bs + (int64_t)(hk * g + r) * p.q_hs;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            uint4 w = make_uint4(0, 0, 0, 0);
            if (r < g) {
                w = *reinterpret_cast<const uint4*>(qrow + 32 * j + 8 * t);
                if constexpr (FUSED) w = rope_rotate8<T>(w, 32 * j + 8 * t, pos_rope, p.rotary_dim, p.cos_table, p.sin_table);
            }
            qa[j][0] = w.x; qa[j][1] = w.y; qa[j][2] = w.z; qa[j][3] = w.w;
        }
    }

    constexpr int NO = SWAP ? NH * 4 : NH * 8;
    float o_acc[NO][4];
#pragma unroll
    for (int i = 0; i < NO; ++i) { o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f; }
    float m_run = -INFINITY;   // running max (log2 domain), identical across the quad (SWAP: head 2t, across r)
    float l_run = 0.f;         // this thread's share of the denominator
    float m_run1 = -INFINITY, l_run1 = 0.f;   // SWAP only: head 2t + 1

    const T* kbase = reinterpret_cast<const T*>(p.k) + (PAGED ? 0 : (int64_t)b * p.k_bs) + (int64_t)hkv * p.k_hs;
    const T* vbase = reinterpret_cast<const T*>(p.v) + (PAGED ? 0 : (int64_t)b * p.v_bs) + (int64_t)hkv * p.v_hs;
    const int32_t* bt = PAGED ? p.block_table + (int64_t)b * p.bt_stride : nullptr;

    // One 16-key step of this warp.  NEW (fused step only): the step holds the appended token, whose k (rotated
    // here) and v come from `qkv` instead of the cache and are written to the cache row `pos` by the threads
    // whose fragment is that row.
    auto step = [&](const int base, auto new_tag) {
        constexpr bool NEW = decltype(new_tag)::value;
        const int last = k_end - 1;
        const T* tok_k = NEW ? token_row(p.heads_q + hkv) : nullptr;
        const T* tok_v = NEW ? token_row(p.heads_q + p.heads_kv + hkv) : nullptr;
        // rows of this step: cache row `key` is kb + (key - key0) * k_rs (paged: inside the step's page)
        const T* kb = kbase;
        const T* vb = vbase;
        int key0 = 0;
        if constexpr (PAGED) {   // (reading the entry one step ahead was measured: no difference)
            const int blk = base >> p.page_shift;
            const int64_t page = bt[blk];
            key0 = blk << p.page_shift;
            kb = kbase + page * p.k_bs;
            vb = vbase + page * p.v_bs;
        }
        T* kcache = const_cast<T*>(kb);   // FUSED: the cache is appended to
        T* vcache = const_cast<T*>(vb);
        // ---- issue every load of the step up front (16 x 16 B per thread at D=128)
        uint4 kf[2][NJ];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            const int key = min(base + 8 * nt + r, last);
            const bool from_tok = NEW && key == pos;
            const T* row = (from_tok ? tok_k : kb + (int64_t)(key - key0) * p.k_rs) + 8 * t;
#pragma unroll
            for (int j = 0; j < NJ; ++j) kf[nt][j] = ld_stream(row + 32 * j);
            if constexpr (NEW) {
                if (from_tok) {
#pragma unroll
                    for (int j = 0; j < NJ; ++j) {
                        kf[nt][j] = rope_rotate8<T>(kf[nt][j], 32 * j + 8 * t, pos_rope, p.rotary_dim, p.cos_table, p.sin_table);
                        if (base + 8 * nt + r == pos)   // the unclamped owner of the row appends it
                            *reinterpret_cast<uint4*>(kcache + (int64_t)(pos - key0) * p.k_rs + 32 * j + 8 * t) = kf[nt][j];
                    }
                }
            }
        }
        uint4 vf[4][NH];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const int key_raw = base + 2 * t + (kk & 1) + 8 * (kk >> 1);
            const int key = min(key_raw, last);
            const bool from_tok = NEW && key == pos;
            const T* row = (from_tok ? tok_v : vb + (int64_t)(key - key0) * p.v_rs) + 8 * r;
#pragma unroll
            for (int h = 0; h < NH; ++h) {
                vf[kk][h] = ld_stream(row + 64 * h);
                if constexpr (NEW) {
                    if (from_tok && key_raw == pos)
                        *reinterpret_cast<uint4*>(vcache + (int64_t)(pos - key0) * p.v_rs + 8 * r + 64 * h) = vf[kk][h];
                }
            }
        }

        if constexpr (SWAP) {
            // ---- S^T = K Q^T: rows = keys base + {r, r + 8}, columns = heads {2t, 2t + 1}
            float st[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                mma_16816<T>(st, kf[0][j].x, kf[1][j].x, kf[0][j].y, kf[1][j].y, qa[j][0], qa[j][1]);
                mma_16816<T>(st, kf[0][j].z, kf[1][j].z, kf[0][j].w, kf[1][j].w, qa[j][2], qa[j][3]);
            }
            const bool dead0 = base + r >= k_end, dead1 = base + r + 8 >= k_end;
            const float s0 = dead0 ? -INFINITY : st[0] * p.scale_log2, s1 = dead0 ? -INFINITY : st[1] * p.scale_log2;
            const float s2 = dead1 ? -INFINITY : st[2] * p.scale_log2, s3 = dead1 ? -INFINITY : st[3] * p.scale_log2;
            float mxa = fmaxf(s0, s2), mxb = fmaxf(s1, s3);
#pragma unroll
            for (int o = 4; o <= 16; o <<= 1) {
                mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, o));
                mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, o));
            }
            const float ma = fmaxf(m_run, mxa), mb = fmaxf(m_run1, mxb);   // finite: the step holds a live key
            const float alpha_a = exp2f(m_run - ma), alpha_b = exp2f(m_run1 - mb);
            const float p0 = exp2f(s0 - ma), p2 = exp2f(s2 - ma), p1 = exp2f(s1 - mb), p3 = exp2f(s3 - mb);
            l_run = l_run * alpha_a + (p0 + p2);
            l_run1 = l_run1 * alpha_b + (p1 + p3);
            m_run = ma; m_run1 = mb;
            if (__any_sync(0xffffffffu, alpha_a != 1.f || alpha_b != 1.f)) {
#pragma unroll
                for (int i = 0; i < NO; ++i) {
                    o_acc[i][0] *= alpha_a; o_acc[i][2] *= alpha_a;
                    o_acc[i][1] *= alpha_b; o_acc[i][3] *= alpha_b;
                }
            }
            // ---- O^T += V^T P^T: P^T fragments are the transposed 8x8 blocks of the packed scores
            const uint32_t pb0 = movmatrix_trans(pack2<T>(p0, p1));   // head r, keys 2t, 2t+1
            const uint32_t pb1 = movmatrix_trans(pack2<T>(p2, p3));   // head r, keys 2t+8, 2t+9
#pragma unroll
            for (int h = 0; h < NH; ++h)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t w0 = u4_word(vf[0][h], i), w1 = u4_word(vf[1][h], i);
                    const uint32_t w2 = u4_word(vf[2][h], i), w3 = u4_word(vf[3][h], i);
                    mma_16816<T>(o_acc[h * 4 + i], __byte_perm(w0, w1, 0x5410u), __byte_perm(w0, w1, 0x7632u),
                                 __byte_perm(w2, w3, 0x5410u), __byte_perm(w2, w3, 0x7632u), pb0, pb1);
                }
            return;
        }
        // ---- S = Q K^T for two 8-key tiles
        float s[2][4];
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                mma_16816<T>(s[nt], qa[j][0], 0u, qa[j][1], 0u, kf[nt][j].x, kf[nt][j].y);
                mma_16816<T>(s[nt], qa[j][2], 0u, qa[j][3], 0u, kf[nt][j].z, kf[nt][j].w);
            }
        }
        // ---- scale, mask the tail, online softmax (this thread: head r, keys base+8nt+2t+{0,1})
        float mx = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                float x = s[nt][e] * p.scale_log2;
                if (base + 8 * nt + 2 * t + e >= k_end) x = -INFINITY;
                s[nt][e] = x;
                mx = fmaxf(mx, x);
            }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float m_new = fmaxf(m_run, mx);           // finite: the step holds at least one live key
        const float alpha = exp2f(m_run - m_new);       // 0 on the first step (m_run = -inf)
        float psum = 0.f;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                s[nt][e] = exp2f(s[nt][e] - m_new);
                psum += s[nt][e];
            }
        l_run = l_run * alpha + psum;
        m_run = m_new;
        if (__any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
            for (int i = 0; i < NH * 8; ++i) { o_acc[i][0] *= alpha; o_acc[i][1] *= alpha; }
        }
        // ---- O += P V
        const uint32_t pa0 = pack2<T>(s[0][0], s[0][1]);   // keys 2t, 2t+1
        const uint32_t pa2 = pack2<T>(s[1][0], s[1][1]);   // keys 2t+8, 2t+9
#pragma unroll
        for (int h = 0; h < NH; ++h)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t sel = (i & 1) ? 0x7632u : 0x5410u;
                const uint32_t b0 = __byte_perm(u4_word(vf[0][h], i >> 1), u4_word(vf[1][h], i >> 1), sel);
                const uint32_t b1 = __byte_perm(u4_word(vf[2][h], i >> 1), u4_word(vf[3][h], i >> 1), sel);
                mma_16816<T>(o_acc[h * 8 + i], pa0, 0u, pa2, 0u, b0, b1);
            }
    };
    for (int base = k_begin + warp * kStepKeys; base < k_end_main; base += kWarps * kStepKeys)
        step(base, std::false_type{});
    if constexpr (FUSED) {
        if (cta_has_new && warp == ((new_base - k_begin) / kStepKeys) % kWarps) step(new_base, std::true_type{});
    }

    // The next kernel on the stream (the combine kernel of the two-launch form, or the next decode call) may be
    // scheduled from here on (programmatic dependent launch, see decode_pdl_level); its griddepcontrol.wait still
    // holds it until this whole grid has completed and its writes are visible.
    asm volatile("griddepcontrol.launch_dependents;");
    // ---- merge the four warp states
    if constexpr (SWAP) {
#pragma unroll
        for (int o = 4; o <= 16; o <<= 1) {
            l_run += __shfl_xor_sync(0xffffffffu, l_run, o);
            l_run1 += __shfl_xor_sync(0xffffffffu, l_run1, o);
        }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int hh = 2 * t + e;
            if (hh < g) {
                if (r == 0) { sm.m[warp][hh] = e ? m_run1 : m_run; sm.l[warp][hh] = e ? l_run1 : l_run; }
#pragma unroll
                for (int h = 0; h < NH; ++h)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        sm.o[warp][hh][64 * h + 8 * r + 2 * i] = o_acc[h * 4 + i][e];
                        sm.o[warp][hh][64 * h + 8 * r + 2 * i + 1] = o_acc[h * 4 + i][2 + e];
                    }
            }
        }
    } else {
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    if (r < g) {
        if (t == 0) { sm.m[warp][r] = m_run; sm.l[warp][r] = l_run; }
#pragma unroll
        for (int h = 0; h < NH; ++h)
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int e = 0; e < 2; ++e) sm.o[warp][r][64 * h + 16 * t + 8 * e + i] = o_acc[h * 8 + i][e];
    }
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < g * D; idx += kWarps * 32) {
        const int hh = idx / D, d = idx - hh * D;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) M = fmaxf(M, sm.m[w][hh]);
        float L = 0.f, acc = 0.f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const float mw = sm.m[w][hh];
            const float sc = (mw == -INFINITY) ? 0.f : exp2f(mw - M);
            L += sc * sm.l[w][hh];
            acc += sc * sm.o[w][hh][d];
        }
        const int hq = hk * g + hh;
        if constexpr (CLUSTER) {
            sm.o[0][hh][d] = acc;   // this thread alone reads and writes column (hh, d)
            if (d == 0) { sm.M[hh] = M; sm.L[hh] = L; }
        } else if (p.num_splits == 1 && !p.force_partials) {
            const float inv = L > 0.f ? 1.f / L : 0.f;
            T* orow = reinterpret_cast<T*>(p.o) + (int64_t)b * p.o_bs + (int64_t)hq * p.o_hs;
            orow[d] = (T)(acc * inv);
            if (d == 0 && p.lse) p.lse[(int64_t)b * p.heads_q + hq] = L > 0.f ? (M + log2f(L)) * kLn2 : -INFINITY;
        } else {
            const int64_t row = ((int64_t)b * p.heads_q + hq) * p.num_splits + split;
            p.o_part[row * D + d] = acc;
            if (d == 0) p.ml_part[row] = make_float2(M, L);
        }
    }
    if constexpr (CLUSTER) {
        // The splits of this (batch, kv-head) are the CTAs of the cluster (rank = split).  Every CTA merges a
        // 128-column slice of the group's g x D outputs out of all the CTAs' shared memory.
        namespace cg = cooperative_groups;
        cg::cluster_group cluster = cg::this_cluster();
        cluster.sync();
        const int S = p.num_splits, rank = (int)cluster.block_rank();
        for (int idx = rank * (kWarps * 32) + threadIdx.x; idx < g * D; idx += S * (kWarps * 32)) {
            const int hh = idx / D, d = idx - hh * D;
            float ms[8], M = -INFINITY;
#pragma unroll
            for (int s2 = 0; s2 < 8; ++s2) {
                ms[s2] = s2 < S ? *cluster.map_shared_rank(&sm.M[hh], s2) : -INFINITY;
                M = fmaxf(M, ms[s2]);
            }
            float L = 0.f, acc = 0.f;
#pragma unroll
            for (int s2 = 0; s2 < 8; ++s2) {
                if (s2 < S) {
                    const float w = (ms[s2] == -INFINITY) ? 0.f : exp2f(ms[s2] - M);
                    L += w * *cluster.map_shared_rank(&sm.L[hh], s2);
                    acc += w * *cluster.map_shared_rank(&sm.o[0][hh][d], s2);
                }
            }
            const int hq = hk * g + hh;
            const float inv = L > 0.f ? 1.f / L : 0.f;
            T* orow = reinterpret_cast<T*>(p.o) + (int64_t)b * p.o_bs + (int64_t)hq * p.o_hs;
            orow[d] = (T)(acc * inv);
            if (d == 0 && p.lse) p.lse[(int64_t)b * p.heads_q + hq] = L > 0.f ? (M + log2f(L)) * kLn2 : -INFINITY;
        }
        cluster.sync();   // nobody leaves while its shared memory is still being read
    }
}

// Merge of the per-split partials (reference flash_combine_kernel, src/flash_attn.cu:877-935, in fp32).
template <typename T, int D>
__global__ void __launch_bounds__(D * kMergeGroups) decode_combine_kernel(const DecodeParams p) {
    extern __shared__ float w_sm[];   // num_splits weights
    __shared__ MergeScratch<D> sc;
    const int hq = blockIdx.x, b = blockIdx.y;
    const int64_t row0 = ((int64_t)b * p.heads_q + hq) * p.num_splits;
    asm volatile("griddepcontrol.wait;" ::: "memory");   // the split kernel has completed (no-op without PDL)
    asm volatile("griddepcontrol.launch_dependents;");    // the next decode kernel may line up behind this one
    float acc, L, M;
    merge_splits<D>(p.o_part, p.ml_part, row0, p.num_splits, w_sm, sc, acc, L, M);
    if (threadIdx.x < D) {
        const float inv = L > 0.f ? 1.f / L : 0.f;
        T* orow = reinterpret_cast<T*>(p.o) + (int64_t)b * p.o_bs + (int64_t)hq * p.o_hs;
        orow[threadIdx.x] = (T)(acc * inv);
        if (threadIdx.x == 0 && p.lse) p.lse[(int64_t)b * p.heads_q + hq] = L > 0.f ? (M + log2f(L)) * kLn2 : -INFINITY;
    }
}

// Portable cluster size.  Non-portable sizes (9 .. 16, cudaFuncAttributeNonPortableClusterSizeAllowed) were measured
// and rejected: with the machine full the scheduler cannot pack them (C4: 9 cluster splits 257 us against 179 us for
// 9 two-launch splits, 16: 232 against 181), and with few units they are erratic (B=4, 4096 keys: 12 splits 25.5 us,
// 16 splits 35.8 us); profiles/r1g_split_sweep16.txt.
constexpr int kMaxClusterSplits = 8;

// In-cluster merge (see the CLUSTER note above the kernel): SFA_DECODE_CLUSTER=0 keeps the two-launch form.
inline bool decode_cluster_enabled() { return decode_tuning().cluster != 0; }

// Programmatic dependent launch (SFA_DECODE_PDL: 0 off, 1 the combine kernel only, 2 = default every decode kernel):
// a kernel launched this way may be scheduled once the kernel before it on the stream has executed
// griddepcontrol.launch_dependents in every CTA (ours do so on leaving the streaming loop; a foreign kernel never does,
// which means "at completion"), and its own griddepcontrol.wait - the first instruction of both kernels - holds it
// until that kernel has completed and flushed.  Launch latency and CTA start-up move under the predecessor's tail.
inline int decode_pdl_level() { return decode_tuning().pdl; }

template <typename K>
int launch_ex(K kern, dim3 grid, int threads, size_t smem, int cluster, bool pdl, const DecodeParams& p, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[2];
    int n = 0;
    if (cluster > 0) {
        at[n].id = cudaLaunchAttributeClusterDimension;
        at[n].val.clusterDim.x = cluster; at[n].val.clusterDim.y = 1; at[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl) {
        at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = at; cfg.numAttrs = n;
    SFA_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, p));
    count_launch();
    return SFA_OK;
}

template <typename T, int D>
int launch_typed(const DecodeLaunch& a, const DecodeParams& p, cudaStream_t stream) {
    const dim3 grid(a.num_splits, a.heads_kv * p.kv_chunks, a.batch);
    const int threads = kWarps * 32;
    const bool pdl_all = decode_pdl_level() >= 2;
    if (a.num_splits >= 2 && a.num_splits <= kMaxClusterSplits && !a.force_partials && decode_cluster_enabled() &&
        (a.block_table || decode_swap_enabled())) {
        const int cs = a.num_splits;
        if (a.block_table)
            return a.fused_qkv ? launch_ex(decode_splitkv_kernel<T, D, true, true, true, true>, grid, threads, 0, cs, pdl_all, p, stream)
                               : launch_ex(decode_splitkv_kernel<T, D, false, true, true, true>, grid, threads, 0, cs, pdl_all, p, stream);
        return a.fused_qkv ? launch_ex(decode_splitkv_kernel<T, D, true, true, false, true>, grid, threads, 0, cs, pdl_all, p, stream)
                           : launch_ex(decode_splitkv_kernel<T, D, false, true, false, true>, grid, threads, 0, cs, pdl_all, p, stream);
    }
    int rc;
    if (a.block_table) {
        rc = a.fused_qkv ? launch_ex(decode_splitkv_kernel<T, D, true, true, true>, grid, threads, 0, 0, pdl_all, p, stream)
                         : launch_ex(decode_splitkv_kernel<T, D, false, true, true>, grid, threads, 0, 0, pdl_all, p, stream);
    } else if (decode_swap_enabled()) {
        rc = a.fused_qkv ? launch_ex(decode_splitkv_kernel<T, D, true, true>, grid, threads, 0, 0, pdl_all, p, stream)
                         : launch_ex(decode_splitkv_kernel<T, D, false, true>, grid, threads, 0, 0, pdl_all, p, stream);
    } else {
#ifdef SFA_PROBES
        rc = a.fused_qkv ? launch_ex(decode_splitkv_kernel<T, D, true, false>, grid, threads, 0, 0, pdl_all, p, stream)
                         : launch_ex(decode_splitkv_kernel<T, D, false, false>, grid, threads, 0, 0, pdl_all, p, stream);
#else
        rc = fail(SFA_ERR_UNSUPPORTED, "the untransposed decode variant exists in the probe build only");
#endif
    }
    if (rc) return rc;
    if (a.num_splits > 1 && !a.force_partials)
        return launch_ex(decode_combine_kernel<T, D>, dim3(a.heads_q, a.batch), D * kMergeGroups, a.num_splits * sizeof(float), 0,
                         decode_pdl_level() >= 1, p, stream);
    return SFA_OK;
}

}  // namespace

// Splits heuristic.  (1) When 2 .. 8 splits per (batch, kv-head) unit fill ONE wave of resident CTAs (4 per SM), take
// that: the splits of a unit then form a thread-block cluster that merges in shared memory (one launch, no
// tail wave).  Measured (tools/decode_split_sweep.py, profiles/r1g_split_sweep.txt): 64 units x 8 splits beat the
// older 16 .. 28 splits + combine launch by 8 - 16 % from 4k to 32k keys; 256 units x 2 by 24 %.  (2) Few units and
// a short cache (launch-latency bound): up to 8 cluster splits of at least 128 keys.  (3) Otherwise (few units, long
// cache) the two-launch form: one wave when that is at most 24 splits, else `waves` rounds of 3 CTAs per SM, chunks
// of at least 256 keys, at most 256 splits.
DecodeTuning& decode_tuning() {
    static DecodeTuning t = [] {
        DecodeTuning v;
        auto geti = [](const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; };
        v.swap = geti("SFA_DECODE_SWAP", v.swap);
        v.cluster = geti("SFA_DECODE_CLUSTER", v.cluster);
        v.pdl = geti("SFA_DECODE_PDL", v.pdl);
        v.waves = geti("SFA_DECODE_WAVES", v.waves);
        v.ctas_per_sm = geti("SFA_DECODE_CTAS", v.ctas_per_sm);
        v.fused_step = geti("SFA_DECODER_FUSED", v.fused_step);
        return v;
    }();
    return t;
}

int decode_head_chunks(int heads_q, int heads_kv) {
    if (heads_kv <= 0 || heads_q < heads_kv) return 1;
    const int g = heads_q / heads_kv;
    int chunks = (g + 7) / 8;
    while (chunks < g && g % chunks != 0) ++chunks;
    return chunks;
}

int choose_decode_splits(int batch, int heads_kv, int max_seqlen, int sm_count) {
    const int waves = decode_tuning().waves > 0 ? decode_tuning().waves : 4;
    const long units = (long)batch * heads_kv;
    const int ctas_per_sm = decode_tuning().ctas_per_sm > 0 ? decode_tuning().ctas_per_sm : 3;
    const long slots = (long)sm_count * ctas_per_sm * waves;
    long s = (slots + units - 1) / units;
    const long max_by_len = max_seqlen / 256 > 0 ? max_seqlen / 256 : 1;
    if (s > max_by_len) s = max_by_len;
    if (s > 256) s = 256;   // one long sequence (batch 1): 8 kv-heads x 222 chunks fill 4 waves of 3 CTAs/SM
    if (s < 1) s = 1;
    if (decode_cluster_enabled() && decode_swap_enabled()) {
        const long one_wave = (long)sm_count * 4 / units;   // splits per unit that fill the resident CTA slots once
        if (one_wave >= 2) {
            long sc = one_wave < kMaxClusterSplits ? one_wave : kMaxClusterSplits;
            const long by_len = max_seqlen / 128 > 0 ? max_seqlen / 128 : 1;
            if (sc > by_len) sc = by_len;
            // 8 splits still count as the one-wave choice while they fill 80 % of the slots (64 units: 512 of 592)
            if (one_wave <= kMaxClusterSplits + 2) return (int)sc;
            if (s <= kMaxClusterSplits) return (int)(s > sc ? s : sc);
            // one wave is also the best two-launch count while it means at most ~24 splits (32 units: 16 - 18 splits
            // 58 us at 16k keys and 99 us at 32k, against 62 and 101 us for the 56 of the rule below)
            if (one_wave <= 24 && s > one_wave) return (int)(one_wave < by_len ? one_wave : by_len);
            // ... and for very few units with a long cache (one sequence: 8 units, 74 splits each fill the 592 CTA slots
            // once): under ncu the split kernel takes 25.3 us with 72 splits against 28.1 - 28.5 with 48 or 128 and 32.9
            // with 96 (1.3 waves); the combine kernel costs 6 us either way (profiles/r2_fwd_tuning.md section 7)
            // Which count near one wave: back-to-back calls over rotating caches (tools/decode_few_units.py,
            // profiles/r2_decode_few_units.txt) put an EVEN cut first - 32768 keys: 64 splits 27.3 us, 72 27.7, 74 27.9;
            // 65536 keys: 64 splits (1024-key chunks) 46.6 us, 74 (896) 47.0, 72 (960, 69 busy CTAs) 51.2 - so take the
            // largest count in the upper quarter of the wave that cuts the cache into equal 64-key multiples, else the
            // CTAs that a whole-8s count leaves busy (chunks are rounded up to 64 keys, the rest would start and idle)
            if (one_wave <= 96 && s > one_wave) {
                long w = 0;
                for (long c = one_wave; c >= one_wave * 3 / 4 && c >= 2; --c)
                    if (max_seqlen % (c * 64) == 0) { w = c; break; }
                if (w == 0) {
                    const long w8 = one_wave / 8 * 8 > 0 ? one_wave / 8 * 8 : one_wave;
                    const long chunk = ((max_seqlen + w8 - 1) / w8 + 63) / 64 * 64;
                    w = (max_seqlen + chunk - 1) / chunk;
                }
                return (int)(w < by_len ? w : by_len);
            }
        }
    }
    return (int)s;
}

size_t decode_workspace_bytes(int batch, int heads_q, int head_dim, int num_splits) {
    if (num_splits <= 1) return 0;
    const size_t rows = (size_t)batch * heads_q * num_splits;
    return rows * head_dim * sizeof(float) + rows * sizeof(float2) + 256;
}

int launch_decode(const DecodeLaunch& a, cudaStream_t stream) {
    if (a.heads_kv <= 0 || a.heads_q % a.heads_kv != 0) return fail(SFA_ERR_INVALID, "heads_q must be a multiple of heads_kv");
    // One CTA serves up to 8 query heads (the 8 columns of its MMAs).  A larger GQA / MQA group is cut into equal chunks
    // of <= 8 heads, each a CTA row of its own over the same kv head: K/V are then read once per chunk.
    const int chunks = decode_head_chunks(a.heads_q, a.heads_kv);
    const int g = a.heads_q / a.heads_kv / chunks;
    if ((long)a.heads_kv * chunks > 65535) return fail(SFA_ERR_UNSUPPORTED, "decode: more than 65535 head chunks");
    if (a.head_dim != 64 && a.head_dim != 128) return fail(SFA_ERR_UNSUPPORTED, "head_dim must be 64 or 128");
    if (a.batch == 0) return SFA_OK;
    if (a.block_table) {
        if (a.page_size < kStepKeys || (a.page_size & (a.page_size - 1)))
            return fail(SFA_ERR_INVALID, "page_size must be a power of two >= 16");
    }
    DecodeParams p;
    p.q = a.q; p.k = a.k; p.v = a.v; p.seqlens = a.seqlens; p.o = a.o; p.lse = a.lse;
    p.o_part = a.o_part; p.ml_part = a.ml_part;
    p.q_bs = a.q_bs; p.q_hs = a.q_hs; p.k_bs = a.k_bs; p.k_rs = a.k_rs; p.k_hs = a.k_hs;
    p.v_bs = a.v_bs; p.v_rs = a.v_rs; p.v_hs = a.v_hs; p.o_bs = a.o_bs; p.o_hs = a.o_hs;
    p.seqlen_k = a.seqlen_k; p.seqlen_add = a.seqlen_add; p.heads_q = a.heads_q; p.group = g;
    p.num_splits = a.num_splits;
    p.scale_log2 = a.scale * kLog2e;
    p.qkv = a.fused_qkv; p.cos_table = a.cos_table; p.sin_table = a.sin_table; p.qkv_stride = a.qkv_stride;
    p.rotary_dim = a.rotary_dim; p.heads_kv = a.heads_kv; p.kv_chunks = chunks; p.force_partials = a.force_partials ? 1 : 0;
    p.block_table = a.block_table; p.bt_stride = a.bt_stride; p.page_shift = 0;
    while (a.block_table && (1 << p.page_shift) < a.page_size) ++p.page_shift;
    if (a.dtype == SFA_FP16) {
        return a.head_dim == 128 ? launch_typed<__half, 128>(a, p, stream) : launch_typed<__half, 64>(a, p, stream);
    }
    return a.head_dim == 128 ? launch_typed<__nv_bfloat16, 128>(a, p, stream)
                             : launch_typed<__nv_bfloat16, 64>(a, p, stream);
}

}  // namespace sfa
