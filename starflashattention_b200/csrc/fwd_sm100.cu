// Prefill attention forward for sm_100a: O = softmax(Q K^T * scale [+ causal mask]) V, LSE.
//
// The reference has no prefill kernel; the arithmetic is the naive formula its
// Python example runs (examples/python/testFlashDecoder.py:84 scores, :88 softmax, :92 probs @ V).  The loop this
// kernel replaces is the reference's single-buffered tile loop (src/flash_attn.cu:769-843, "TODO: DoubleBuffer" :720).
//
// Persistent kernel: one CTA per SM, each working through WORK ITEMS = (batch, head, block of NQ x 128 query rows)
// handed out by an atomic counter in heavy-first order (see launch_fwd).  Roles inside a CTA:
//   warps 0 .. 4*NQ-1   softmax warpgroups, one per 128-row query tile; thread i of a warpgroup owns
//                       row i (TMEM lane i): tcgen05.ld S -> score buffer handed back -> online softmax
//                       in fp32 (exp2 on log2e-prescaled scores, fused causal/tail masking) ->
//                       tcgen05.st P as bf16/fp16 into the tile's own P columns; at the end of an item
//                       O / l -> swizzled shared memory -> TMA store, LSE -> global
//   warp 4*NQ           scheduler + TMA producer: fetches the next item (atomicAdd) one item ahead and
//                       publishes it to the other warps, loads the item's Q tiles, then K_0 V_0 K_1 V_1 ...
//                       through a ring of 128-key tiles (128-byte swizzle, mbarrier full/empty) that
//                       runs on ACROSS items: the next item's Q and first tiles arrive under the current
//                       item's last P.V and epilogue
//   warps 4*NQ+1 ..     tcgen05 issuers, one warp per query tile (one elected lane issues):
//                       S_q = Q_q K_j^T (SS, both operands K-major) and O_q += P_q V_j (TS: A = P in
//                       TMEM, B = V in smem, MN-major); the first of them owns the TMEM allocation
// TMEM: P0 P1 | S | O0 O1 = 2 x 64 + 128 + 2 x 128 columns at head_dim 128.  P does not alias S: a
// softmax warpgroup frees the score buffer as soon as its tcgen05.ld has landed in registers, so
// the tensor pipe computes S(q, j+1) while the warpgroup is still exponentiating tile j.  Issue order of tile q's
// MMA warp: Q.K(q, j) then P.V(q, j-1), for j = 0, 1, ...  The single score buffer alternates between
// the two query tiles (head_dim 64 and the one-tile variant have room for two buffers).
// Every mbarrier is followed by running counters (tiles, ring loads, items seen so far), not by the tile index
// inside an item, so the pipeline never drains between items.
// O is rescaled lazily: the reference maximum of a row only moves when the true maximum has grown
// by more than 2^8, so the TMEM read-modify-write of O is rare after the first tiles.
#include "internal.h"
#include "sm100_ptx.cuh"

#include <cuda.h>
#include <limits.h>
#include <math.h>
#include <mutex>
#include <type_traits>

namespace sfa {
namespace {

using namespace ptx;

constexpr int kBlockM = 128;        // query rows per tile (one TMEM lane each)
constexpr int kBlockN = 128;        // keys per tile
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kRescaleThreshold = 8.0f;   // log2 units
// Head start of softmax warpgroup 0 at head_dim 64 (cycles): with a score buffer each, the two warpgroups otherwise
// run in lockstep (both on the MUFU, then both off it).  Measured on five shapes (profiles/r1_fwd_tuning.md):
// 1700 cycles is +0.4 % .. +4.2 %, never a loss.
constexpr int kStaggerCycles64 = 1700;

struct FwdParams {
    float* lse;
    int* error_flag;
    int* sched;           // [0] next-item counter, [1] CTAs done (both zero between launches)
    int seqlen_q, seqlen_k, heads_q, group;
    int causal;
    float scale_log2;
    int stagger_cycles;   // head start of warpgroup 0 over warpgroup 1 (see launch_fwd)
    // work-item order: groups of `group_heads` (batch, head) pairs - the last `tail_heads` heads form one group of
    // their own; inside a group query blocks from the last (heaviest under the causal mask) to the first, heads innermost
    int m_blocks, heads_total, group_heads, tail_heads, total_items;
};

__device__ int g_fwd_error;   // set when an mbarrier wait times out (pipeline bug guard)
#ifdef SFA_PROBES
// event trace of the CTA with blockIdx.x == 0 (tools/fwd_trace.py): slot 32 + (n * 2 + q) * 16 + e holds clock64()
// of event e of the n-th key tile this CTA processed for query tile q; slots 24.. are per-CTA stamps
__device__ unsigned long long g_fwd_probe[32 + 4096];
#define SFA_TRACE(cond, n, q, e) do { if ((cond) && (n) < 128) g_fwd_probe[32 + ((n) * 2 + (q)) * 16 + (e)] = clock64(); } while (0)
#define SFA_STAMP(cond, i) do { if (cond) g_fwd_probe[i] = clock64(); } while (0)
#else
#define SFA_TRACE(cond, n, q, e) do { } while (0)
#define SFA_STAMP(cond, i) do { } while (0)
#endif

// Bounded mbarrier wait: a mis-programmed pipeline must end the kernel, not hang the GPU.  The bound is a count of
// failed try_waits (each suspends the thread for a hardware time slice), not a clock: this code is inlined at every
// wait of the kernel, and the %globaltimer version made up a third of its instructions (instruction cache).
constexpr uint32_t kMaxWaitSpins = 1u << 19;   // measured: 2^24 failed try_waits take about 60 s, so this is about 2 s
__device__ __forceinline__ bool wait_barrier(uint32_t bar, uint32_t parity, int* error_flag, int code) {
    if (mbar_try_wait(bar, parity)) return true;
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins >= kMaxWaitSpins || ((spins & 0xFFF) == 0 && *(volatile int*)error_flag != 0)) {
            atomicCAS(error_flag, 0, code);
            return false;
        }
    }
    return true;
}

// ---- packed fp32x2 arithmetic (FFMA2 / FADD2 / FMUL2 on sm_100): halves the issue slots of the softmax
__device__ __forceinline__ uint64_t pack_f2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack_f2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
// Order-pinned variants (asm volatile keeps their relative program order): the exp2 loop below is
// software-pipelined by hand so that the MUFU, the softmax's critical unit, is fed at an even rate.
__device__ __forceinline__ uint64_t ffma2_v(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ uint64_t fadd2_v(uint64_t a, uint64_t b) {
    uint64_t r;
    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ float ex2_v(float x) {
    float y;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t fmul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// 2^x for two values on the FMA/ALU pipes instead of the MUFU: x = n + f with n = round(x) taken from the
// mantissa of x + 1.5*2^23, f in [-0.5, 0.5], 2^f by a degree-3 minimax polynomial (max relative error
// 7.5e-5, below the bf16/fp16 rounding of P), and n added to the exponent field.  x is clamped at -125
// so the exponent cannot wrap; callers use this only on tiles without -inf entries.
__device__ __forceinline__ uint64_t exp2_poly2(float x0, float x1) {
    constexpr float kMagic = 12582912.f;   // 1.5 * 2^23
    x0 = fmaxf(x0, -125.f);
    x1 = fmaxf(x1, -125.f);
    const uint64_t x = pack_f2(x0, x1);
    const uint64_t magic = pack_f2(kMagic, kMagic);
    const uint64_t t = fadd2(x, magic);
    const uint64_t u = fadd2(t, pack_f2(-kMagic, -kMagic));
    const uint64_t f = fadd2(x, u ^ 0x8000000080000000ull);   // x - round(x)
    uint64_t r = ffma2(pack_f2(0.055171459913253784f, 0.055171459913253784f), f,
                       pack_f2(0.2426108568906784f, 0.2426108568906784f));
    r = ffma2(r, f, pack_f2(0.6932609677314758f, 0.6932609677314758f));
    r = ffma2(r, f, pack_f2(0.9999281167984009f, 0.9999281167984009f));
    float t0, t1, r0, r1;
    unpack_f2(t, t0, t1);
    unpack_f2(r, r0, r1);
    r0 = __int_as_float(__float_as_int(r0) + (__float_as_int(t0) << 23));
    r1 = __int_as_float(__float_as_int(r1) + (__float_as_int(t1) << 23));
    return pack_f2(r0, r1);
}
// The same polynomial with scalar arithmetic: the packed FADD2 / FFMA2 issue every second cycle, so they save nothing
// over two scalar instructions, and the scalar form needs no pack / unpack moves (tools/micro/exp_poly.cu: 14.2 against
// 14.8 cycles per pair and sub-partition with two warps, a quarter of the pairs on the polynomial; 16.6 without it).
__device__ __forceinline__ float exp2_poly1(float x) {
    constexpr float kMagic = 12582912.f;   // 1.5 * 2^23
    x = fmaxf(x, -125.f);
    const float t = x + kMagic;
    const float f = x - (t - kMagic);      // x - round(x)
    float r = fmaf(0.055171459913253784f, f, 0.2426108568906784f);
    r = fmaf(r, f, 0.6932609677314758f);
    r = fmaf(r, f, 0.9999281167984009f);
    return __int_as_float(__float_as_int(r) + (__float_as_int(t) << 23));
}
template <bool BF16>
__device__ __forceinline__ uint32_t pack_pair(float lo, float hi) {
    if constexpr (BF16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&v);
    } else {
        __half2 v = __floats2half2_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&v);
    }
}

// conversion pinned in program order like the other members of the exp2 loop
template <bool BF16>
__device__ __forceinline__ uint32_t pack_pair_v(float lo, float hi) {
    uint32_t r;
    if constexpr (BF16) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    else asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// Register split between the softmax warpgroups and the warpgroup holding the producer / issuer warps.
// setmaxnreg.inc blocks until the CTA's own pool (registers per thread at launch x threads) can cover the
// request, so the split must fit in that pool exactly - an over-subscribed split hangs the kernel.
template <int NQ>
struct FwdRegs {
    static constexpr int kThreads = NQ * 128 + 128;   // softmax warpgroups + one warpgroup of four single-purpose warps
    static constexpr bool kSet = NQ == 2;
    // registers ptxas may use at launch: floor(65536 / threads) rounded down to a multiple of 8
    static constexpr int kLaunch = (65536 / kThreads) / 8 * 8;
    static constexpr int kSoftmax = 216;
    static constexpr int kOther = 72;
    static_assert(!kSet || NQ * 128 * kSoftmax + 128 * kOther <= kThreads * kLaunch,
                  "register split exceeds the CTA pool");
    static_assert(kSoftmax % 8 == 0 && kOther % 8 == 0, "setmaxnreg takes multiples of 8");
};

constexpr int next_pow2_cols(int n) { return n <= 32 ? 32 : n <= 64 ? 64 : n <= 128 ? 128 : n <= 256 ? 256 : 512; }

// TMEM columns: P tiles (64 columns each: 128 keys as packed 16-bit pairs), score buffers, O accumulators.
template <int D, int NQ>
struct FwdTmem {
    static constexpr int kPOff = 0;
    static constexpr int kSOff = NQ * (kBlockN / 2);
    // two score buffers wherever 512 columns allow it (everything but two query tiles at head_dim 128)
    static constexpr int kSBufs = (kSOff + 2 * kBlockN + NQ * D <= 512) ? 2 : 1;
    static constexpr int kOOff = kSOff + kSBufs * kBlockN;
    static constexpr int kUsed = kOOff + NQ * D;
    static constexpr int kCols = next_pow2_cols(kUsed);
    static_assert(kUsed <= 512, "TMEM budget exceeded");
};

template <int D, int NQ, int NSTAGE>
struct FwdSmem {
    static_assert(NSTAGE % 2 == 0, "K tiles live in the even ring slots, V tiles in the odd ones");
    static constexpr int kSubTileBytes = kBlockM * 128;              // 128 rows x 64 elements x 2 B
    static constexpr int kTileBytes = (D / 64) * kSubTileBytes;      // one 128 x D tile
    static constexpr int kQOff = 0;
    static constexpr int kRingOff = NQ * kTileBytes;
    // O staging for the TMA store: per query tile one 128-row x 64-column sub-tile (used D / 64 times per item)
    static constexpr int kStageOff = kRingOff + NSTAGE * kTileBytes;
    static constexpr int kBarOff = kStageOff + NQ * kSubTileBytes;
    // barriers: q_full[NQ], q_empty[NQ], kv_full[NSTAGE], kv_empty[NSTAGE], p_ready[NQ][2], pv_done[NQ],
    //           s_free[2], sched_full[2], sched_empty[2], s_full[2]
    static constexpr int kNumBars = 5 * NQ + 2 * NSTAGE + 8;
    static constexpr int kSchedOff = (kBarOff + kNumBars * 8 + 15) / 16 * 16;   // two published work items, 32 B each
    static constexpr int kTmemPtrOff = kSchedOff + 64;
    static constexpr int kBytes = kTmemPtrOff + 16;
    static constexpr int kAlloc = kBytes + 1024;                     // slack for the 1024 B alignment
    static_assert(kAlloc <= 232448, "shared memory budget exceeded");
};

// One work item as the scheduler warp publishes it: two query tiles (n1, and everything about query tile 1, is
// unused with one query tile per CTA) of one (batch, head).
struct Item {
    int m0, h, b, hk;   // first query row, query head, batch entry, kv head
    int n0, n1;         // key tiles query tile 0 / 1 needs (causal: up to the diagonal of its last row)
    int n_max;
    int index;          // >= total_items: the end of the schedule
};

template <int NQ>
__device__ __forceinline__ Item decode_item(const FwdParams& p, int idx) {
    Item it;
    it.index = idx;
    const int bulk_heads = p.heads_total - p.tail_heads;
    const int bulk_items = bulk_heads * p.m_blocks;
    int head0, gh, r;
    if (idx < bulk_items) {
        const int per_group = p.group_heads * p.m_blocks;
        const int g = idx / per_group;
        r = idx - g * per_group;
        head0 = g * p.group_heads;
        gh = min(p.group_heads, bulk_heads - head0);
    } else {
        r = idx - bulk_items;
        head0 = bulk_heads;
        gh = p.tail_heads;
    }
    const int m_desc = r / gh;
    const int head = head0 + (r - m_desc * gh);
    const int m_block = p.m_blocks - 1 - m_desc;
    it.b = head / p.heads_q;
    it.h = head - it.b * p.heads_q;
    it.hk = it.h / p.group;
    it.m0 = m_block * (NQ * kBlockM);
    const int kv_tiles_total = (p.seqlen_k + kBlockN - 1) / kBlockN;
    const int causal_shift = p.seqlen_k - p.seqlen_q;   // query i sees keys <= i + shift
    int n[2] = {0, 0};
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        int t = kv_tiles_total;
        if (p.causal) {
            const int last_row = min(it.m0 + (q + 1) * kBlockM, p.seqlen_q) - 1;
            const int last_key = last_row + causal_shift;
            t = last_key < 0 ? 0 : min(kv_tiles_total, last_key / kBlockN + 1);
        }
        if (it.m0 + q * kBlockM >= p.seqlen_q) t = 0;   // tile entirely past the end of Q
        n[q] = t;
    }
    it.n0 = n[0];
    it.n1 = n[1];
    it.n_max = max(n[0], n[1]);
    return it;
}

// BF16: data type (false = fp16).  POLY: of every 8 column pairs, how many take the polynomial exp2.
template <int D, int NQ, int NSTAGE, bool BF16, int POLY, int EXPDELAY, bool DEFER, int POLL, int POLYSCALAR>
__global__ void __launch_bounds__(FwdRegs<NQ>::kThreads, 1)
fwd_kernel(const __grid_constant__ CUtensorMap q_map, const __grid_constant__ CUtensorMap k_map,
           const __grid_constant__ CUtensorMap v_map, const __grid_constant__ CUtensorMap o_map, const FwdParams p) {
    using L = FwdSmem<D, NQ, NSTAGE>;
    using R = FwdRegs<NQ>;
    using T = FwdTmem<D, NQ>;
    constexpr bool kSetRegs = R::kSet;
    constexpr int kSoftmaxWarps = 4 * NQ;
    constexpr int kWarpKV = kSoftmaxWarps, kWarpQK = kSoftmaxWarps + 1, kWarpPV = kSoftmaxWarps + 2, kWarpSched = kSoftmaxWarps + 3;
    constexpr int kSBufs = T::kSBufs;
    constexpr int kSub = D / 64;
    constexpr int kSchedConsumers = 4 * NQ + 3;   // softmax warps + K/V producer + the two issuer warps

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + L::kBarOff;
    auto bar_q_full = [&](int q) { return bar_base + 8u * q; };
    // every Q.K of the item that reads Q tile q has completed: the tile may be overwritten by the next item's
    auto bar_q_empty = [&](int q) { return bar_base + 8u * (NQ + q); };
    auto bar_kv_full = [&](uint32_t s) { return bar_base + 8u * (2 * NQ + s); };
    auto bar_kv_empty = [&](uint32_t s) { return bar_base + 8u * (2 * NQ + NSTAGE + s); };
    // P is handed to the P.V warp in two halves of 64 keys, so P.V of the first half runs under the second's exp2
    auto bar_p_ready = [&](int q, int half) { return bar_base + 8u * (2 * NQ + 2 * NSTAGE + 2 * q + half); };
    // every P.V of query tile q has completed: P(q) may be overwritten, O(q) may be rescaled / read
    auto bar_pv_done = [&](int q) { return bar_base + 8u * (4 * NQ + 2 * NSTAGE + q); };
    // warpgroup q has loaded its scores into registers: the next Q.K may overwrite that buffer
    auto bar_s_free = [&](int i) { return bar_base + 8u * (5 * NQ + 2 * NSTAGE + i); };
    auto bar_sched_full = [&](int i) { return bar_base + 8u * (5 * NQ + 2 * NSTAGE + 2 + i); };
    auto bar_sched_empty = [&](int i) { return bar_base + 8u * (5 * NQ + 2 * NSTAGE + 4 + i); };
    // Scores of a tile are in TMEM.  One barrier per query tile (two tiles), or per score buffer (one tile, two
    // alternating buffers): with ONE barrier for two buffers the Q.K of tiles t and t + 1 can both complete before the
    // warpgroup looks for tile t, and a parity wait cannot tell "two phases on" from "not yet" - it would hang.
    auto bar_s_full = [&](int i) { return bar_base + 8u * (5 * NQ + 2 * NSTAGE + 6 + i); };
    auto s_full_index = [&](int q, uint32_t t) { return NQ == 2 ? q : (int)(t & 1); };
    auto s_full_parity = [&](uint32_t t) { return NQ == 2 ? (t & 1) : ((t >> 1) & 1); };
    volatile int4* sched_slot = reinterpret_cast<volatile int4*>(smem_gen + L::kSchedOff);   // [2 * entry], [2 * entry + 1]
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem_gen + L::kTmemPtrOff);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int causal_shift = p.seqlen_k - p.seqlen_q;
#ifdef SFA_PROBES
    const bool trace_cta = blockIdx.x == 0;
#endif
    SFA_STAMP(trace_cta && threadIdx.x == 0, 24);

    if (warp == kWarpKV && lane == 0) {
        prefetch_tensormap(&q_map);
        prefetch_tensormap(&k_map);
        prefetch_tensormap(&v_map);
        prefetch_tensormap(&o_map);
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            mbar_init(bar_q_full(q), 1);
            mbar_init(bar_q_empty(q), 1);
            mbar_init(bar_p_ready(q, 0), 4);   // one elected arrival per softmax warp of the tile's warpgroup
            mbar_init(bar_p_ready(q, 1), 4);
            mbar_init(bar_pv_done(q), 1);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_s_free(i), 4);
            mbar_init(bar_s_full(i), 1);
            mbar_init(bar_sched_full(i), 1);
            mbar_init(bar_sched_empty(i), kSchedConsumers);
        }
#pragma unroll
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(bar_kv_full(s), 1);
            mbar_init(bar_kv_empty(s), 1);   // one commit behind the last Q.K (K slots) / P.V (V slots) that reads the tile
        }
        fence_barrier_init();
    }
    if (warp == kWarpQK) {
        tmem_alloc(smem_base + L::kTmemPtrOff, T::kCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    SFA_STAMP(trace_cta && threadIdx.x == 0, 25);

    // Consumer side of the schedule: the k-th item of this CTA (index >= total_items ends the loop).
    auto next_item = [&](int k, bool& ok) {
        ok = __all_sync(0xffffffffu, wait_barrier(bar_sched_full(k & 1), (k >> 1) & 1, p.error_flag, 12)) != 0;
        const int4 a = const_cast<const int4&>(reinterpret_cast<const volatile int4&>(sched_slot[2 * (k & 1)]));
        const int4 b = const_cast<const int4&>(reinterpret_cast<const volatile int4&>(sched_slot[2 * (k & 1) + 1]));
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_sched_empty(k & 1));
        Item it;
        it.m0 = a.x; it.h = a.y; it.b = a.z; it.hk = a.w;
        it.n0 = b.x; it.n1 = b.y; it.n_max = b.z; it.index = ok ? b.w : INT_MAX;
        return it;
    };
    auto wait_all = [&](uint32_t bar, uint32_t parity, int code) {
        return __all_sync(0xffffffffu, wait_barrier(bar, parity, p.error_flag, code)) != 0;
    };
    // the same wait as a spin on test_wait: no suspension, so no wake-up latency - for the one wait that sits on the
    // kernel's critical cycle (POLL bit 0: the Q.K warp on the score buffer; bit 1: the P.V warp on P)
    auto poll_all = [&](uint32_t bar, uint32_t parity, int code) {
        uint32_t spins = 0;
        bool ok_ = true;
        while (!mbar_test_wait(bar, parity)) {
            if (++spins >= (1u << 27)) {
                atomicCAS(p.error_flag, 0, code);
                ok_ = false;
                break;
            }
        }
        return __all_sync(0xffffffffu, ok_) != 0;
    };

    // Four single-purpose warps beside the softmax warpgroups; with two query tiles they form a third warpgroup so the
    // register file can be re-split: the softmax warpgroups take what these give up.
    if (warp == kWarpSched) {
        // =============================================================== scheduler + Q producer
        // Fetches the next item one item ahead (atomicAdd; the first one is static), decodes it and publishes it
        // to every other warp; loads the item's Q tiles as soon as the previous item's last Q.K has released them -
        // a warp of its own, so a Q load never queues behind the K/V ring's waits.
        if constexpr (kSetRegs) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R::kOther));
        int item = blockIdx.x;
        const uint64_t pol_q = l2_policy_evict_first();   // a Q tile is read once
        int q_loads0 = 0, q_loads1 = 0;   // Q tiles loaded so far, per query tile
        bool ok = true;
        for (int k = 0; ok; ++k) {
            if (k >= 2) ok = wait_all(bar_sched_empty(k & 1), ((k >> 1) - 1) & 1, 13);
            if (!ok) break;
            Item it;
            it.index = item;
            if (item < p.total_items) it = decode_item<NQ>(p, item);
            if (lane == 0) {
                if (item < p.total_items) {
                    const_cast<int4&>(reinterpret_cast<volatile int4&>(sched_slot[2 * (k & 1)])) = make_int4(it.m0, it.h, it.b, it.hk);
                    const_cast<int4&>(reinterpret_cast<volatile int4&>(sched_slot[2 * (k & 1) + 1])) = make_int4(it.n0, it.n1, it.n_max, item);
                } else {
                    const_cast<int4&>(reinterpret_cast<volatile int4&>(sched_slot[2 * (k & 1) + 1])) = make_int4(0, 0, 0, INT_MAX);
                }
                mbar_arrive(bar_sched_full(k & 1));
            }
            __syncwarp();
            if (item >= p.total_items) break;
            int nxt = 0;
            if (lane == 0) nxt = (int)gridDim.x + atomicAdd(p.sched, 1);   // used one item later: its latency is hidden
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int n_q = q == 0 ? it.n0 : it.n1;
                if (n_q == 0) continue;
                const int loads = q == 0 ? q_loads0 : q_loads1;
                if (loads >= 1) ok = ok && wait_all(bar_q_empty(q), (loads - 1) & 1, 11);
                if (ok && elect_one()) {
                    mbar_arrive_expect_tx(bar_q_full(q), L::kTileBytes);
#pragma unroll
                    for (int s = 0; s < kSub; ++s)
                        tma_load_4d_hint(smem_base + L::kQOff + q * L::kTileBytes + s * L::kSubTileBytes, &q_map, bar_q_full(q),
                                         64 * s, it.h, it.m0 + q * kBlockM, it.b, pol_q);
                }
                __syncwarp();
                if (q == 0) ++q_loads0; else ++q_loads1;
            }
            item = __shfl_sync(0xffffffffu, nxt, 0);
        }
        // Last CTA out rezeroes the schedule counters for the next launch that uses this slot.
        if (lane == 0) {
            __threadfence();
            const unsigned old = atomicInc(reinterpret_cast<unsigned*>(p.sched + 1), gridDim.x - 1);
            if (old == gridDim.x - 1) p.sched[0] = 0;
        }
    } else if (warp == kWarpKV) {
        // =============================================================== K/V producer
        // K_0 V_0 K_1 V_1 ... of every item through one ring that runs on across items (K in the even slots, V in
        // the odd ones): the next item's first tiles arrive under the current item's last P.V and epilogue.
        if constexpr (kSetRegs) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R::kOther));
        uint32_t ring = 0;          // K/V tiles loaded so far (all items)
        const uint64_t pol_kv = l2_policy_evict_last();   // every CTA working on the head reads these tiles again
        bool ok = true;
        for (int k = 0; ok; ++k) {
            const Item it = next_item(k, ok);
            if (it.index >= p.total_items) break;
            for (int i = 0; i < 2 * it.n_max && ok; ++i) {
                const uint32_t gi = ring + (uint32_t)i, slot = gi % NSTAGE, use = gi / NSTAGE;
                if (use >= 1) ok = wait_all(bar_kv_empty(slot), (use - 1) & 1, 1);
                if (ok && elect_one()) {
                    mbar_arrive_expect_tx(bar_kv_full(slot), L::kTileBytes);
                    const CUtensorMap* map = (i & 1) ? &v_map : &k_map;
#pragma unroll
                    for (int s = 0; s < kSub; ++s)
                        tma_load_4d_hint(smem_base + L::kRingOff + slot * L::kTileBytes + s * L::kSubTileBytes, map,
                                         bar_kv_full(slot), 64 * s, it.hk, (i >> 1) * kBlockN, it.b, pol_kv);
                }
                __syncwarp();
            }
            ring += 2u * (uint32_t)it.n_max;
        }
    } else if (warp == kWarpQK) {
        // =============================================================== Q.K issuer: S(q, j) = Q_q K_j^T for both query tiles
        // Order: key tile major, query tile minor - the order in which the softmax warpgroups hand the score buffer
        // back.  This warp does nothing else, so it is parked on the buffer's barrier when the hand-back comes: the
        // turnaround of the single score buffer (head_dim 128) is the critical cycle of the kernel.
        if constexpr (kSetRegs) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R::kOther));
        // The whole warp runs the loop (uniform control flow keeps descriptors in uniform registers); one elected
        // lane issues.  Descriptors are built once per operand tile and advanced by compile-time constants.
        constexpr uint32_t idesc_qk = mma_idesc(BF16, 0, 0, kBlockM, kBlockN);
        // Q, K: K-major, rows of 128 B, 8-row groups 1024 B apart.
        constexpr uint64_t desc_kmajor = smem_desc_sw128(16, 1024);
        const uint64_t ring_k_desc = smem_desc_at(desc_kmajor, smem_base + L::kRingOff);
        const uint64_t q_desc0 = smem_desc_at(desc_kmajor, smem_base + L::kQOff);
        bool ok = true;
        uint32_t ring = 0;                  // K/V tiles of earlier items
        uint32_t tiles0 = 0, tiles1 = 0;    // key tiles each query tile has been through in earlier items
        int q_uses0 = 0, q_uses1 = 0;       // items that loaded Q tile 0 / 1
        // the Q.K issued last (any item): its warpgroup hands the shared score buffer on
        bool has_prev = false;
        int prev_q = 0;
        uint32_t prev_t = 0;
#ifdef SFA_PROBES
        const bool tr = trace_cta && lane == 0;
#endif
        for (int k = 0; ok; ++k) {
            const Item it = next_item(k, ok);
            if (it.index >= p.total_items) break;
            for (int j = 0; j < it.n_max && ok; ++j) {
                const uint32_t gi = ring + 2u * (uint32_t)j, kslot = gi % NSTAGE;
                ok = wait_all(bar_kv_full(kslot), (gi / NSTAGE) & 1, 2);
                const uint64_t k_desc = ring_k_desc + (uint64_t)(kslot * (L::kTileBytes >> 4));
#pragma unroll
                for (int q = 0; q < NQ; ++q) {
                    const int n_q = q == 0 ? it.n0 : it.n1;
                    if (j >= n_q) continue;
                    const uint32_t t = (q == 0 ? tiles0 : tiles1) + (uint32_t)j;   // running tile count of this query tile
                    if (j == 0) {      // the item's Q tile
                        ok = ok && wait_all(bar_q_full(q), (q == 0 ? q_uses0 : q_uses1) & 1, 3);
                        if (q == 0) ++q_uses0; else ++q_uses1;
                    }
                    SFA_TRACE(tr, t, q, 0);
                    // the buffer's previous scores must have been loaded into registers by their warpgroup
                    int buf = 0;
                    if constexpr (NQ == 2 && kSBufs == 1) {
                        if (has_prev) ok = ok && ((POLL & 1) ? poll_all(bar_s_free(prev_q), prev_t & 1, 9) : wait_all(bar_s_free(prev_q), prev_t & 1, 9));
                    } else if constexpr (NQ == 2) {
                        buf = q;
                        if (t >= 1) ok = ok && ((POLL & 1) ? poll_all(bar_s_free(q), (t - 1) & 1, 9) : wait_all(bar_s_free(q), (t - 1) & 1, 9));
                    } else {
                        buf = t & 1;
                        if (t >= 2) ok = ok && wait_all(bar_s_free(t & 1), ((t >> 1) - 1) & 1, 9);
                    }
                    tc_fence_after();
                    SFA_TRACE(tr, t, q, 1);
                    if (ok && elect_one()) {
                        const uint64_t q_desc = q_desc0 + (uint64_t)(q * (L::kTileBytes >> 4));
                        const uint32_t s_tmem = tmem_base + T::kSOff + buf * kBlockN;
                        // start-address field is in 16-byte units and never carries out of its 14 bits (smem < 256 KB)
#pragma unroll
                        for (int kc = 0; kc < D / 16; ++kc) {
                            const uint32_t off16 = ((kc >> 2) * L::kSubTileBytes + (kc & 3) * 32) >> 4;
                            mma_ss(s_tmem, q_desc + off16, k_desc + off16, idesc_qk, kc > 0);
                        }
                        tc_commit(bar_s_full(s_full_index(q, t)));
                        if (j == n_q - 1) tc_commit(bar_q_empty(q));
                    }
                    __syncwarp();
                    SFA_TRACE(tr, t, q, 2);
                    has_prev = true; prev_q = q; prev_t = t;
                }
                if (ok && elect_one()) tc_commit(bar_kv_empty(kslot));   // every Q.K that reads K_j has been issued
                __syncwarp();
            }
            tiles0 += (uint32_t)it.n0;
            tiles1 += (uint32_t)it.n1;
            ring += 2u * (uint32_t)it.n_max;
        }
    } else if (warp == kWarpPV) {
        // =============================================================== P.V issuer: O(q) += P(q, j) V_j for both query tiles
        // (TS: A = P in TMEM, B = V in smem, MN-major), in two halves of 64 keys each, in the order the softmax
        // warpgroups produce P.  The first P.V of an item overwrites O: its P arrives only after the warpgroup has
        // read the previous item's O (see the deferred epilogue below).
        if constexpr (kSetRegs) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R::kOther));
        constexpr uint32_t idesc_pv = mma_idesc(BF16, 0, 1, kBlockM, D);
        // V: MN-major (head_dim contiguous), 8-key groups 1024 B apart, next 64 head_dim columns one sub-tile further.
        constexpr uint64_t desc_v = smem_desc_sw128(L::kSubTileBytes, 1024);
        const uint64_t ring_v_desc = smem_desc_at(desc_v, smem_base + L::kRingOff);
        bool ok = true;
        uint32_t ring = 0, tiles0 = 0, tiles1 = 0;
#ifdef SFA_PROBES
        const bool tr = trace_cta && lane == 0;
#endif
        for (int k = 0; ok; ++k) {
            const Item it = next_item(k, ok);
            if (it.index >= p.total_items) break;
            for (int j = 0; j < it.n_max && ok; ++j) {
                const uint32_t gi = ring + 2u * (uint32_t)j + 1u, vslot = gi % NSTAGE;
                ok = wait_all(bar_kv_full(vslot), (gi / NSTAGE) & 1, 4);
                const uint64_t v_desc = ring_v_desc + (uint64_t)(vslot * (L::kTileBytes >> 4));
#pragma unroll
                for (int q = 0; q < NQ; ++q) {
                    const int n_q = q == 0 ? it.n0 : it.n1;
                    if (j >= n_q) continue;
                    const uint32_t t = (q == 0 ? tiles0 : tiles1) + (uint32_t)j;
                    const uint32_t p_tmem = tmem_base + T::kPOff + q * (kBlockN / 2);
                    const uint32_t o_tmem = tmem_base + T::kOOff + q * D;
#pragma unroll
                    for (int half = 0; half < 2; ++half) {      // keys [64 * half, 64 * half + 64)
                        ok = ok && ((POLL & 2) ? poll_all(bar_p_ready(q, half), t & 1, 5) : wait_all(bar_p_ready(q, half), t & 1, 5));
                        tc_fence_after();
                        SFA_TRACE(tr, t, q, half == 0 ? 3 : 5);
                        if (ok && elect_one()) {
#pragma unroll
                            for (int kk = 0; kk < kBlockN / 32; ++kk) {
                                const int kc = half * (kBlockN / 32) + kk;
                                mma_ts(o_tmem, p_tmem + kc * 8, v_desc + (uint64_t)(kc * (2048 >> 4)), idesc_pv, j > 0 || kc > 0);
                            }
                            if (half == 1) tc_commit(bar_pv_done(q));
                        }
                        __syncwarp();
                    }
                    SFA_TRACE(tr, t, q, 6);
                }
                if (ok && elect_one()) tc_commit(bar_kv_empty(vslot));   // every P.V that reads V_j has been issued
                __syncwarp();
            }
            tiles0 += (uint32_t)it.n0;
            tiles1 += (uint32_t)it.n1;
            ring += 2u * (uint32_t)it.n_max;
        }
    } else {
        // =============================================================== softmax warpgroups
        // Warpgroup q serves query tile q; thread i owns row i of the tile (TMEM lane i).
        if constexpr (kSetRegs) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(R::kSoftmax));
        constexpr int kCols = kBlockN;
        const int q = warp >> 2;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const bool wg_leader = (warp & 3) == 0 && lane == 0;   // issues the warpgroup's TMA stores
        const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t p_tmem = lane_base + T::kPOff + q * (kBlockN / 2);   // my packed P columns
        const uint32_t o_tmem = lane_base + T::kOOff + q * D;
        const uint32_t stage = smem_base + L::kStageOff + q * L::kSubTileBytes;
        const uint64_t pol_o = l2_policy_evict_first();   // O is written once and not read again
        const float sl2 = p.scale_log2;
        const uint64_t sl2_2 = pack_f2(sl2, sl2);
        bool ok = true;
        uint32_t tiles = 0;         // key tiles this warpgroup has been through in earlier items
#ifdef SFA_PROBES
        const bool tr = trace_cta && lane == 0 && (warp & 3) == 0;
#endif
        // The epilogue of an item is DEFERRED into the next item's first tile: once that tile's P sits in TMEM (its
        // hand-over to the P.V warp held back), the warpgroup writes the finished item's O - by then the item's
        // last P.V is long done, so nothing is waited for, and the score buffer of the new item was drained first.
        // What the deferred epilogue needs of the finished item:
        bool pending = false;
        float pend_inv_l = 0.f, pend_lse = 0.f;
        int pend_h = 0, pend_b = 0, pend_row0 = 0;
        // O / l -> swizzled staging tile -> TMA store (rows past seqlen_q are clipped by the tensor map), LSE -> global
        auto write_item = [&](bool have_o, float inv_l, float lse, int h_, int b_, int row0) {
#pragma unroll 1
            for (int c = 0; c < kSub; ++c) {
                uint32_t o[64];
                SFA_STAMP(tr, 4 * c + 0);
                if (have_o) {
                    tmem_ld32(o_tmem + 64 * c, o);
                    tmem_ld32(o_tmem + 64 * c + 32, o + 32);
                    tmem_wait_ld();
                } else {
#pragma unroll
                    for (int i = 0; i < 64; ++i) o[i] = 0u;
                }
                uint32_t packed[32];
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    packed[i] = pack_pair<BF16>(__uint_as_float(o[2 * i]) * inv_l, __uint_as_float(o[2 * i + 1]) * inv_l);
                // the staging tile is free once the previous TMA store has read it
                SFA_STAMP(tr, 4 * c + 1);
                if (wg_leader) bulk_wait_read0();
                named_bar_sync(1 + q, 128);
                SFA_STAMP(tr, 4 * c + 2);
                // row r of the tile is 128 B; its 16-byte chunk i sits at chunk i ^ (r & 7) (the 128-byte swizzle
                // of the tensor map): eight consecutive rows hit eight different bank groups
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    st_shared_v4(stage + row_in_tile * 128 + ((i ^ (row_in_tile & 7)) << 4), packed[4 * i],
                                 packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
                fence_proxy_async_smem();
                named_bar_sync(1 + q, 128);
                if (wg_leader) {
                    tma_store_4d_hint(&o_map, stage, 64 * c, h_, row0, b_, pol_o);
                    bulk_commit();
                }
                SFA_STAMP(tr, 4 * c + 3);
            }
            SFA_STAMP(tr, 8);
            const int row = row0 + row_in_tile;
            if (row < p.seqlen_q && p.lse) p.lse[((int64_t)b_ * p.heads_q + h_) * p.seqlen_q + row] = lse;
        };
        // With a score buffer of their own (head_dim 64) nothing couples the two warpgroups, and started together
        // they stay in lockstep: both in the exp2 loop (MUFU contended), then both in the load + max phase (MUFU
        // idle).  A head start for warpgroup 0 puts one's exp2 loop under the other's max phase.
        if (NQ == 2 && q == 1 && p.stagger_cycles > 0) {
            const long long t0 = clock64();
            while (clock64() - t0 < p.stagger_cycles) __nanosleep(64);
        }
        for (int k = 0; ok; ++k) {
            const Item it = next_item(k, ok);
            const bool end = it.index >= p.total_items;
            const int n_q = end ? 0 : (q == 0 ? it.n0 : it.n1);   // (no runtime-indexed array: it would live in local memory)
            const int row0 = it.m0 + q * kBlockM;
            if (n_q == 0) {
                // nothing to hide the pending epilogue behind: write it now
                if (pending && ok) {
                    ok = wait_barrier(bar_pv_done(q), (tiles - 1u) & 1, p.error_flag, 10);
                    tc_fence_after();
                    write_item(true, pend_inv_l, pend_lse, pend_h, pend_b, pend_row0);
                    pending = false;
                }
                if (end) break;
                // rows that see no key at all (causal with more queries than keys): O = 0, LSE = -inf
                if (row0 < p.seqlen_q) write_item(false, 0.f, -INFINITY, it.h, it.b, row0);
                continue;
            }
            SFA_TRACE(tr, tiles, q, 15);
            const int row = row0 + row_in_tile;
            const int key_limit = p.causal ? row + causal_shift : INT_MAX;   // last visible key of this row
            float m_used = -INFINITY;   // reference maximum (raw score units) the exponentials are taken against
            uint64_t l_acc[2] = {0ull, 0ull};   // four partial row sums as two f32x2 accumulators

            for (int j = 0; j < n_q && ok; ++j) {
                const uint32_t t = tiles + (uint32_t)j;
                const int buf = kSBufs == 1 ? 0 : (NQ == 2 ? q : (int)(t & 1));
                const int free_idx = NQ == 2 ? q : (int)(t & 1);
                const uint32_t s_tmem = lane_base + T::kSOff + buf * kBlockN;
                ok = wait_barrier(bar_s_full(s_full_index(q, t)), s_full_parity(t), p.error_flag, 7);
                tc_fence_after();
                SFA_TRACE(tr, t, q, 8);
                uint32_t su[kCols];
#pragma unroll
                for (int c = 0; c < kCols; c += 32) tmem_ld32(s_tmem + c, &su[c]);
                // The buffer goes back the moment the loads have landed: its turnaround sits on the critical
                // cycle of the pipeline (two query tiles share one score buffer at head_dim 128).
                tmem_wait_ld();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_s_free(free_idx));
                SFA_TRACE(tr, t, q, 9);
                const int key0 = j * kBlockN;
                const bool need_mask = (key0 + kBlockN - 1 > row0 + causal_shift && p.causal) || (key0 + kBlockN > p.seqlen_k);
                const int lim = min(key_limit, p.seqlen_k - 1) - key0;   // columns <= lim stay (masked tiles only)
                // ---- row maximum: four independent chains
                float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
                auto scan_half = [&](int base) {
                    if (need_mask) {
#pragma unroll
                        for (int c = base; c < base + kCols / 2; ++c)
                            if (c > lim) su[c] = 0xff800000u;   // -inf
                    }
#pragma unroll
                    for (int c = base; c < base + kCols / 2; c += 2)
                        mx[(c >> 1) & 3] = fmaxf(mx[(c >> 1) & 3], fmaxf(__uint_as_float(su[c]), __uint_as_float(su[c + 1])));
                };
                scan_half(0);
                scan_half(kCols / 2);
                const float mx_all = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));

                // P(q) of the previous tile (the previous ITEM's last one when j == 0) is still being read by the
                // tensor pipe, and O(q) still accumulated into, until pv_done
                bool pv_waited = t == 0;
                auto wait_pv = [&]() {
                    if (!pv_waited) {
                        ok = wait_barrier(bar_pv_done(q), (t - 1) & 1, p.error_flag, 8);
                        tc_fence_after();
                        pv_waited = true;
                    }
                };
                const float m_new = fmaxf(m_used, mx_all);
                if (j == 0) {
                    m_used = m_new;
                } else {
                    const bool grow = (m_new - m_used) * sl2 > kRescaleThreshold;   // false for -inf - -inf (NaN)
                    if (__any_sync(0xffffffffu, grow)) {
                        wait_pv();
                        const float alpha = grow ? ex2((m_used - m_new) * sl2) : 1.f;
                        const uint64_t alpha2 = pack_f2(alpha, alpha);
                        l_acc[0] = fmul2(l_acc[0], alpha2);
                        l_acc[1] = fmul2(l_acc[1], alpha2);
                        if (grow) m_used = m_new;
#pragma unroll
                        for (int c = 0; c < D / 16; ++c) {   // 16-column chunks: the scores stay in registers
                            uint32_t o[16];
                            tmem_ld16(o_tmem + 16 * c, o);
                            tmem_wait_ld();
#pragma unroll
                            for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                            tmem_st16(o_tmem + 16 * c, o);
                        }
                    }
                }
                const float neg_m = (m_used == -INFINITY) ? 0.f : -m_used * sl2;
                const uint64_t neg_m2 = pack_f2(neg_m, neg_m);
                const bool hold = DEFER && j == 0 && pending;   // the hand-over of this P waits for the deferred epilogue
                SFA_TRACE(tr, t, q, 10);
                // ---- P = exp2(S * scale_log2 - m): MUFU for most pairs, an FMA-pipe polynomial for POLY of every 8
                // pairs on tiles without masked (-inf) entries; P leaves for TMEM in blocks of 64 keys.
                auto exp_and_store = [&](auto poly_tag) {
                    constexpr int kPoly = decltype(poly_tag)::value;
                    constexpr int kPairs = kCols / 2;
                    constexpr int kAhead = 3;            // pairs between the scale FFMA2 and its MUFUs
                    // pairs between a pair's two MUFUs and the instructions that consume their results (row sum,
                    // conversion): the warp issues in order, so a consumer right behind its MUFU stalls the MUFUs
                    // of the following pairs for one MUFU latency per pair (measured: 24 cycles per pair against 16
                    // of MUFU time, profiles/r1_fwd_tuning.md third pass)
                    constexpr int kDelay = EXPDELAY;
                    uint64_t xq[kAhead];                 // scaled scores waiting for their exp2
                    uint64_t done[kDelay > 0 ? kDelay : 1];   // exponentials waiting for their consumers
                    auto scale = [&](int i) {
                        return ffma2_v(pack_f2(__uint_as_float(su[2 * i]), __uint_as_float(su[2 * i + 1])), sl2_2, neg_m2);
                    };
#pragma unroll
                    for (int i = 0; i < kAhead; ++i) xq[i] = scale(i);
                    uint32_t pk[32];
                    auto consume = [&](int i, uint64_t e) {      // pair i: row sum, conversion, store and hand-off of P
                        float p0, p1;
                        unpack_f2(e, p0, p1);
                        l_acc[i & 1] = fadd2_v(l_acc[i & 1], e);
                        pk[i & 31] = pack_pair_v<BF16>(p0, p1);
                        if ((i & 31) == 31) {
                            if (i == 31) wait_pv();      // the previous tile's P.V must have drained before P is overwritten
                            tmem_st32(p_tmem + (i >> 5) * 32, pk);
                        }
                        // the hand-off of a 64-key block trails its store by a few pairs, so that the MUFU keeps
                        // running while the store drains (the last block has nothing left to overlap with)
                        constexpr int kTrail = 6;
                        if (((i & 31) == ((31 + kTrail) & 31) && i > 31) || i == kPairs - 1) {
                            tmem_wait_st();
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0 && !hold) {
                                if (i == kPairs - 1) mbar_arrive(bar_p_ready(q, kPairs / 32 - 1));
                                else mbar_arrive(bar_p_ready(q, (i >> 5) - 1));
                            }
                            SFA_TRACE(tr, t, q, i == kPairs - 1 ? 12 : 11);
                        }
                    };
#pragma unroll
                    for (int i = 0; i < kPairs; ++i) {
                        float x0, x1, p0, p1;
                        unpack_f2(xq[i % kAhead], x0, x1);
                        if (i + kAhead < kPairs) xq[i % kAhead] = scale(i + kAhead);
                        uint64_t e;
                        // which pairs take the polynomial: POLYSCALAR 0 / 1: the last kPoly of every 8 (a block);
                        // 2 / 3: spread, every (8 / kPoly)-th pair; odd values: scalar arithmetic
                        const bool on_poly = kPoly > 0 && ((POLYSCALAR < 2) ? ((i & 7) >= 8 - kPoly)
                                                                           : ((i % (8 / (kPoly > 0 ? kPoly : 1))) == (8 / (kPoly > 0 ? kPoly : 1)) - 1));
                        if (on_poly) {
                            if constexpr ((POLYSCALAR & 1) != 0) e = pack_f2(exp2_poly1(x0), exp2_poly1(x1));
                            else e = exp2_poly2(x0, x1);
                        } else {
                            p0 = ex2_v(x0); p1 = ex2_v(x1);
                            e = pack_f2(p0, p1);
                        }
                        if constexpr (kDelay == 0) {
                            consume(i, e);
                        } else {
                            if (i >= kDelay) consume(i - kDelay, done[i % kDelay]);
                            done[i % kDelay] = e;
                        }
                    }
                    if constexpr (kDelay > 0) {
#pragma unroll
                        for (int i = kPairs; i < kPairs + kDelay; ++i) consume(i - kDelay, done[i % kDelay]);
                    }
                };
                if (POLY > 0 && !need_mask) exp_and_store(std::integral_constant<int, POLY>{});
                else exp_and_store(std::integral_constant<int, 0>{});
                if (hold) {
                    // the finished item's last P.V was waited for above (wait_pv at the first P store); O is final
                    SFA_TRACE(tr, t - 1u, q, 13);
                    write_item(true, pend_inv_l, pend_lse, pend_h, pend_b, pend_row0);
                    pending = false;
                    SFA_TRACE(tr, t - 1u, q, 14);
                    // O has been read (tcgen05.wait::ld inside): the new item's first P.V may overwrite it
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(bar_p_ready(q, 0));
                        mbar_arrive(bar_p_ready(q, 1));
                    }
                }
            }
            // ---- end of the item: its O is written later (deferred epilogue)
            float l0, l1, l2, l3;
            unpack_f2(l_acc[0], l0, l1);
            unpack_f2(l_acc[1], l2, l3);
            const float l_sum = (l0 + l1) + (l2 + l3);
            pend_inv_l = l_sum > 0.f ? 1.f / l_sum : 0.f;
            pend_lse = l_sum > 0.f ? (m_used * sl2 + log2f(l_sum)) * kLn2 : -INFINITY;
            pend_h = it.h; pend_b = it.b; pend_row0 = row0;
            pending = true;
            tiles += (uint32_t)n_q;
            if constexpr (!DEFER) {     // the epilogue at the end of its own item
                SFA_STAMP(tr, 9);
                if (ok) ok = wait_barrier(bar_pv_done(q), (tiles - 1u) & 1, p.error_flag, 10);
                tc_fence_after();
                SFA_TRACE(tr, tiles - 1u, q, 13);
                write_item(true, pend_inv_l, pend_lse, pend_h, pend_b, pend_row0);
                pending = false;
                SFA_TRACE(tr, tiles - 1u, q, 14);
                SFA_STAMP(tr, 10);
            }
        }
        if (wg_leader) bulk_wait_all0();   // the stores read shared memory: it must outlive them
    }

    // ---- teardown: every tcgen05 operation of this CTA has completed once the softmax warps are here
    tc_fence_before();
    __syncthreads();
    if (warp == kWarpQK) {
        tc_fence_after();
        tmem_dealloc(tmem_base, T::kCols);
    }
    SFA_STAMP(trace_cta && threadIdx.x == 0, 28);
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            ptr = nullptr;
        return reinterpret_cast<EncodeTiledFn>(ptr);
    }();
    return fn;
}

// (head_dim, heads, rows, batch) view of a (batch, rows, heads, head_dim) strided tensor, boxes of
// 64 x 1 x 128 x 1 elements with the 128-byte swizzle the UMMA descriptors expect.
int make_map(CUtensorMap* map, const void* base, int dtype, int head_dim, int heads, int rows, int batch,
             int64_t head_stride, int64_t row_stride, int64_t batch_stride) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return fail(SFA_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[4] = {(cuuint64_t)head_dim, (cuuint64_t)heads, (cuuint64_t)rows, (cuuint64_t)batch};
    cuuint64_t strides[3] = {(cuuint64_t)head_stride * 2, (cuuint64_t)row_stride * 2, (cuuint64_t)batch_stride * 2};
    // A dimension of extent 1 may carry any stride; keep the encoder's 16-byte rule satisfied.
    for (int i = 0; i < 3; ++i)
        if (strides[i] == 0) strides[i] = 16;
    cuuint32_t box[4] = {64, 1, (cuuint32_t)kBlockM, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, dtype == SFA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                     const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SFA_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return SFA_OK;
}

struct FwdMaps {
    alignas(64) CUtensorMap q, k, v, o;
};

template <int D, int NQ, int NSTAGE, bool BF16, int POLY, int EXPDELAY, bool DEFER, int POLL, int POLYSCALAR>
int launch_instance(const FwdMaps& m, FwdParams p, int batch, int sm_count, cudaStream_t stream) {
    using L = FwdSmem<D, NQ, NSTAGE>;
    auto kern = fwd_kernel<D, NQ, NSTAGE, BF16, POLY, EXPDELAY, DEFER, POLL, POLYSCALAR>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] {
        attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kAlloc);
    });
    SFA_CUDA_CHECK(attr_err);
    // Work-item order (items are handed out in this order by the atomic counter).  Two forces: (1) items that run at
    // the same time should share K/V in L2 - few heads at a time, as a grid launched head by head would do; with 8 / 32
    // heads of C3 interleaved the ncu capture shows 2.7 / 4.3 GB of DRAM reads against 1.61 GB of Q + K + V
    // (profiles/r2_fwd_tuning.md); (2) under the causal mask the END of the schedule should be a long descending
    // sweep over many heads, so that the last items handed out are short: the spread of the CTAs' finishing times is
    // about the length of the items among the last grid's worth (one rank's eighth of C3: 4 / 32 heads per sweep
    // 0.510 / 0.430 ms; its share of C5: 1 / 4 heads 3.63 / 3.14 ms).  So: the bulk in groups of one grid's worth of
    // items (query blocks from the last, heaviest, to the first; heads innermost), and the last `tail_heads` heads as ONE
    // group, sized so that its last grid's worth of items is under 2 % of a CTA's total work, within 128 MiB of K/V.
    p.m_blocks = (p.seqlen_q + NQ * kBlockM - 1) / (NQ * kBlockM);
    p.heads_total = batch * p.heads_q;
    p.group_heads = (sm_count + p.m_blocks - 1) / p.m_blocks;
    p.tail_heads = 0;
    if (p.causal) {
        const long long kv_bytes_per_head = 4ll * (p.seqlen_k > 0 ? p.seqlen_k : 1) * D / p.group;   // K and V, per query head
        const long long l2_heads = (128ll << 20) / kv_bytes_per_head;
        // long sequences (more query blocks than SMs): still four heads per sweep (C5: 1 / 2 / 4 / 8 heads 28.98 / 28.53 /
        // 28.22 / 28.46 ms)
        const int few = (int)(l2_heads < 4 ? l2_heads : 4);
        if (p.group_heads < few) p.group_heads = few;
        const double per_cta = (double)p.heads_total * p.m_blocks * (p.m_blocks + 1) / sm_count;   // key-tile pairs per CTA
        long long want = (long long)(2.0 * sm_count / (0.02 * per_cta)) + 1;
        if (want > l2_heads) want = l2_heads;
        if (want < p.group_heads) want = p.group_heads;
        if (want > p.heads_total) want = p.heads_total;
        p.tail_heads = (int)want;
    }
#ifdef SFA_PROBES
    if (fwd_tuning().group > 0) { p.group_heads = fwd_tuning().group; p.tail_heads = 0; }
    if (fwd_tuning().tail >= 0) p.tail_heads = fwd_tuning().tail < p.heads_total ? fwd_tuning().tail : p.heads_total;
#endif
    if (p.group_heads < 1) p.group_heads = 1;
    // keep the query heads of one kv head in one group (GQA)
    p.group_heads = (p.group_heads + p.group - 1) / p.group * p.group;
    p.tail_heads = p.tail_heads / p.group * p.group;
    const long long total = (long long)p.m_blocks * p.heads_total;
    if (total > INT_MAX / 2) return fail(SFA_ERR_UNSUPPORTED, "fwd: more than 2^30 work items");
    p.total_items = (int)total;
    const int grid = (int)(total < sm_count ? total : sm_count);
    if ((p.sched = fwd_sched_counters(stream)) == nullptr) return SFA_ERR_CUDA;
    kern<<<grid, FwdRegs<NQ>::kThreads, L::kAlloc, stream>>>(m.q, m.k, m.v, m.o, p);
    SFA_CUDA_CHECK(cudaGetLastError());
    count_launch();
    return SFA_OK;
}

// Host launcher: the tile configuration per (head_dim, rows, dtype) - what the reference's Traits<> template
// (src/traits.h:7-26) selects at compile time.  Ring depth: as many 128-key tiles as fit beside the Q tiles and
// the O staging tiles.
template <bool BF16>
int launch_dtype(const FwdMaps& m, const FwdParams& p, int batch, int head_dim, int nq, int sm_count, cudaStream_t stream) {
    const FwdTuning& tune = fwd_tuning();
    (void)tune;
#define SFA_LAUNCH(D_, NQ_, NS_, POLY_, DELAY_, DEFER_, POLL_, PS_) \
    return launch_instance<D_, NQ_, NS_, BF16, POLY_, DELAY_, DEFER_, POLL_, PS_>(m, p, batch, sm_count, stream)
    // DELAY (pairs between a pair's MUFUs and its consumers in the exp2 loop) and DEFER (epilogue inside the next
    // item's first tile) as measured on C3 / C2 (profiles/r2_fwd_tuning.md): 2 / 0 pairs, not deferred.
    if (head_dim == 128) {
        if (nq == 1) SFA_LAUNCH(128, 1, 4, 0, 2, false, 0, 0);
#ifdef SFA_PROBES
        if (tune.poly == 2) SFA_LAUNCH(128, 2, 4, 2, 2, false, 0, 0);
        if (tune.poly == 12) SFA_LAUNCH(128, 2, 4, 2, 2, false, 0, 1);
        if (tune.poly == 11) SFA_LAUNCH(128, 2, 4, 1, 2, false, 0, 1);
        if (tune.defer == 1) SFA_LAUNCH(128, 2, 4, 0, 2, true, 0, 0);
        if (tune.poly == 32) SFA_LAUNCH(128, 2, 4, 2, 2, false, 0, 3);
        if (tune.poly == 31) SFA_LAUNCH(128, 2, 4, 1, 2, false, 0, 3);
        if (tune.delay == 0) SFA_LAUNCH(128, 2, 4, 0, 0, false, 0, 0);
        if (tune.poll == 1) SFA_LAUNCH(128, 2, 4, 0, 2, false, 1, 0);
#endif
        SFA_LAUNCH(128, 2, 4, 0, 2, false, 0, 0);
    }
    if (nq == 1) SFA_LAUNCH(64, 1, 6, 0, 0, false, 0, 0);
#ifdef SFA_PROBES
    if (tune.poly == 2) SFA_LAUNCH(64, 2, 6, 2, 0, false, 0, 0);
    if (tune.poly == 12) SFA_LAUNCH(64, 2, 6, 2, 0, false, 0, 1);
    if (tune.poly == 13) SFA_LAUNCH(64, 2, 6, 3, 0, false, 0, 1);
    if (tune.defer == 1) SFA_LAUNCH(64, 2, 6, 2, 0, true, 0, 0);
    if (tune.poly == 22) SFA_LAUNCH(64, 2, 6, 2, 0, false, 0, 2);   // packed, spread
    if (tune.poly == 32) SFA_LAUNCH(64, 2, 6, 2, 0, false, 0, 3);   // scalar, spread
    if (tune.poly == 34) SFA_LAUNCH(64, 2, 6, 4, 0, false, 0, 3);   // scalar, spread, every second pair
    if (tune.poly == 31) SFA_LAUNCH(64, 2, 6, 1, 0, false, 0, 3);   // scalar, one pair of 8
#endif
    // head_dim 64 is MUFU-bound (a tile's exp2 takes twice its MMA time).  In the round-1 kernel a quarter of the
    // exponentials on the FMA-pipe polynomial paid (0.404 -> 0.390 ms on C2); in the persistent kernel no share and no
    // arrangement of it does (none 0.399, 2/8 packed 0.413, 2/8 scalar spread 0.408, 1/8 0.423, 4/8 0.537 ms), although
    // the loop in isolation gains 14 % (tools/micro/exp_poly.cu) - profiles/r2_fwd_tuning.md.  MUFU only.
    SFA_LAUNCH(64, 2, 6, 0, 0, false, 0, 0);
#undef SFA_LAUNCH
}

}  // namespace

// Tuning overrides, read from the environment ONCE per process (test / tuning builds only act on most of them).
const FwdTuning& fwd_tuning() {
    static const FwdTuning t = [] {
        FwdTuning v;
        auto geti = [](const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; };
        v.poly = geti("SFA_FWD_POLY", -1);
        v.softrel = geti("SFA_FWD_SOFTREL", -1);
        v.stagger = geti("SFA_FWD_STAGGER", -1);
        v.nq = geti("SFA_FWD_NQ", 0);
        v.v1 = geti("SFA_FWD_V1", 0);
        v.group = geti("SFA_FWD_GROUP", 0);
        v.delay = geti("SFA_FWD_DELAY", -1);
        v.defer = geti("SFA_FWD_DEFER", -1);
        v.poll = geti("SFA_FWD_POLL", 0);
        v.tail = geti("SFA_FWD_TAIL", -1);
        return v;
    }();
    return t;
}

int fwd_error_flag_read(int* value) {
    SFA_CUDA_CHECK(cudaMemcpyFromSymbol(value, g_fwd_error, sizeof(int)));
    return SFA_OK;
}
int fwd_probe_read(unsigned long long* out, int n) {
#ifdef SFA_PROBES
    if (n > 32 + 4096) n = 32 + 4096;
    SFA_CUDA_CHECK(cudaMemcpyFromSymbol(out, g_fwd_probe, n * sizeof(unsigned long long)));
    return SFA_OK;
#else
    (void)out; (void)n;
    return fail(SFA_ERR_UNSUPPORTED, "the prefill trace exists in the probe build only (libStarFlashAttention_probes.so)");
#endif
}
int fwd_error_flag_clear() {
    const int zero = 0;
    SFA_CUDA_CHECK(cudaMemcpyToSymbol(g_fwd_error, &zero, sizeof(int)));
    return SFA_OK;
}

int launch_fwd(const sfa_fwd_args& a, cudaStream_t stream) {
    if (!a.q || !a.k || !a.v || !a.o) return fail(SFA_ERR_INVALID, "fwd: null tensor pointer");
    if (a.batch < 0 || a.seqlen_q < 0 || a.seqlen_k < 0 || a.heads_q <= 0 || a.heads_kv <= 0)
        return fail(SFA_ERR_INVALID, "fwd: negative or zero dimension");
    if (a.heads_q % a.heads_kv) return fail(SFA_ERR_INVALID, "fwd: heads_q must be a multiple of heads_kv");
    if (a.head_dim != 64 && a.head_dim != 128) return fail(SFA_ERR_UNSUPPORTED, "fwd: head_dim must be 64 or 128");
    if (a.dtype != SFA_FP16 && a.dtype != SFA_BF16) return fail(SFA_ERR_INVALID, "fwd: dtype must be fp16 or bf16");
    const int64_t strides[] = {a.q_batch_stride, a.q_row_stride, a.q_head_stride, a.k_batch_stride, a.k_row_stride,
                               a.k_head_stride, a.v_batch_stride, a.v_row_stride, a.v_head_stride, a.o_batch_stride,
                               a.o_row_stride, a.o_head_stride};
    for (int64_t s : strides)
        if (s % 8 != 0 || s < 0) return fail(SFA_ERR_INVALID, "fwd: every stride must be a non-negative multiple of 8 elements");
    const uintptr_t ptrs[] = {(uintptr_t)a.q, (uintptr_t)a.k, (uintptr_t)a.v, (uintptr_t)a.o};
    for (uintptr_t x : ptrs)
        if (x % 16) return fail(SFA_ERR_INVALID, "fwd: tensor base pointers must be 16-byte aligned");
    if (a.batch == 0 || a.seqlen_q == 0) return SFA_OK;
    const FwdTuning& tune = fwd_tuning();
#ifdef SFA_PROBES
    if (tune.v1) return launch_fwd_v1(a, stream);   // the round-1 kernel (one CTA per query block), A/B runs only
#endif
    const DeviceInfo* dev = current_device();
    if (!dev) return SFA_ERR_NO_DEVICE;

    FwdMaps m;
    int rc;
    // seqlen_k == 0 still needs valid maps; a one-row extent keeps the encoder happy and is never read.
    const int rows_k = a.seqlen_k > 0 ? a.seqlen_k : 1;
    if ((rc = make_map(&m.q, a.q, a.dtype, a.head_dim, a.heads_q, a.seqlen_q, a.batch, a.q_head_stride, a.q_row_stride, a.q_batch_stride))) return rc;
    if ((rc = make_map(&m.k, a.k, a.dtype, a.head_dim, a.heads_kv, rows_k, a.batch, a.k_head_stride, a.k_row_stride, a.k_batch_stride))) return rc;
    if ((rc = make_map(&m.v, a.v, a.dtype, a.head_dim, a.heads_kv, rows_k, a.batch, a.v_head_stride, a.v_row_stride, a.v_batch_stride))) return rc;
    if ((rc = make_map(&m.o, a.o, a.dtype, a.head_dim, a.heads_q, a.seqlen_q, a.batch, a.o_head_stride, a.o_row_stride, a.o_batch_stride))) return rc;

    FwdParams p = {};
    p.lse = a.lse;
    p.seqlen_q = a.seqlen_q; p.seqlen_k = a.seqlen_k; p.heads_q = a.heads_q; p.group = a.heads_q / a.heads_kv;
    p.causal = a.causal ? 1 : 0;
    const float scale = a.softmax_scale > 0.f ? a.softmax_scale : 1.f / sqrtf((float)a.head_dim);
    p.scale_log2 = scale * kLog2e;
    p.stagger_cycles = tune.stagger >= 0 ? tune.stagger : (a.head_dim == 64 ? kStaggerCycles64 : 0);
    void* flag_addr = nullptr;
    SFA_CUDA_CHECK(cudaGetSymbolAddress(&flag_addr, g_fwd_error));
    p.error_flag = reinterpret_cast<int*>(flag_addr);

    const int nq = tune.nq == 1 || tune.nq == 2 ? tune.nq : (a.seqlen_q > kBlockM ? 2 : 1);
    return a.dtype == SFA_BF16 ? launch_dtype<true>(m, p, a.batch, a.head_dim, nq, dev->sm_count, stream)
                               : launch_dtype<false>(m, p, a.batch, a.head_dim, nq, dev->sm_count, stream);
}

}  // namespace sfa
