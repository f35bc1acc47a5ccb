// Prefill attention forward for sm_100a: O = softmax(Q K^T * scale [+ causal mask]) V, LSE.
//
// The reference has no prefill kernel; the arithmetic is the naive formula its
// Python example runs (examples/python/testFlashDecoder.py:84 scores, :88 softmax, :92 probs @ V).
//
// Structure (one CTA per NQ x 128 query rows of one (batch, head)):
//   warps 0 .. 4*NQ-1   softmax warpgroups, one per 128-row query tile; thread i of a warpgroup owns
//                       row i (TMEM lane i): tcgen05.ld S -> score buffer handed back -> online softmax
//                       in fp32 (exp2 on log2e-prescaled scores, fused causal/tail masking) ->
//                       tcgen05.st P as bf16/fp16 into the tile's own P columns -> epilogue O / l -> global
//   warp 4*NQ           TMA producer: Q tiles once, then K_0 V_0 K_1 V_1 ... through a ring of
//                       128-key tiles in shared memory (128-byte swizzle), mbarrier full/empty
//   warps 4*NQ+1 ..     tcgen05 issuers, one warp per query tile (one elected lane issues):
//                       S_q = Q_q K_j^T (SS, both operands K-major) and O_q += P_q V_j (TS: A = P in
//                       TMEM, B = V in smem, MN-major); the first of them owns the TMEM allocation
// TMEM: P0 P1 | S | O0 O1 = 2 x 64 + 128 + 2 x 128 columns at head_dim 128.  P does not alias S: a
// softmax warpgroup frees the score buffer as soon as its tcgen05.ld has landed in registers, so
// the tensor pipe computes S(q, j+1) while the warpgroup is still exponentiating tile j.  The
// softmax warpgroups therefore run back to back (throughput bound) instead of sitting in the
// latency chain softmax -> P.V -> Q.K(next) of an aliased S/P layout.  Issue order of tile q's MMA warp:
// Q.K(q, j) then P.V(q, j-1), for j = 0, 1, ...  The single score buffer alternates between
// the two query tiles (head_dim 64 and the one-tile variant have room for two buffers).
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
    void* o;
    float* lse;
    int* error_flag;
    int64_t o_bs, o_rs, o_hs;
    int seqlen_q, seqlen_k, heads_q, group;
    int causal, is_bf16;
    float scale_log2;
    int stagger_cycles;   // head start of warpgroup 0 over warpgroup 1 (see launch_fwd)
    int load_order;       // 1: prologue loads go out Q0, K0, Q1, V0, ... (default); 0: Q0, Q1, K0, V0, ... (SFA_FWD_LOADORDER)
};

__device__ int g_fwd_error;
__device__ unsigned long long g_fwd_probe[32 + 2048];   // DBG == 5: cycle counters of CTA (0,0,0), see tools/fwd_probe.py   // set when an mbarrier wait times out (pipeline bug guard)

// Bounded mbarrier wait: a mis-programmed pipeline must end the kernel, not hang the GPU.
__device__ __forceinline__ bool wait_barrier(uint32_t bar, uint32_t parity, int* error_flag, int code) {
    if (mbar_try_wait(bar, parity)) return true;
    const uint64_t t0 = global_timer_ns();
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 0xFF) == 0) {
            if (*(volatile int*)error_flag != 0) return false;
            if (global_timer_ns() - t0 > 2000000000ull) {
                atomicCAS(error_flag, 0, code);
                return false;
            }
        }
    }
    return true;
}

// DBG == 6: event trace of CTA (0,0,0): slot 32 + (j * 2 + q) * 16 + e holds clock64() of event e of tile (q, j).
// Events 0-6 are written by the MMA warp, 8-12 by the first softmax warp of the tile's warpgroup.
__device__ __forceinline__ void trace_event(bool on, int q, int j, int e) {
    if (on && j < 64) g_fwd_probe[32 + (j * 2 + q) * 16 + e] = clock64();
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

// Register split between the softmax warpgroups and the warpgroup holding the TMA / MMA warps (NQ == 2 only).
// setmaxnreg.inc blocks until the CTA's own pool (registers per thread at launch x threads) can cover the
// request, so the split must fit in that pool exactly - an over-subscribed split hangs the kernel.
template <int NQ>
struct FwdRegs {
    static constexpr int kThreads = NQ * 128 + (NQ == 2 ? 128 : 64);
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
    static constexpr int kSubTileBytes = kBlockM * 128;              // 128 rows x 64 elements x 2 B
    static constexpr int kTileBytes = (D / 64) * kSubTileBytes;      // one 128 x D tile
    static constexpr int kQOff = 0;
    static constexpr int kRingOff = NQ * kTileBytes;
    static constexpr int kBarOff = kRingOff + NSTAGE * kTileBytes;
    // barriers: q_full[NQ], kv_full[NSTAGE], kv_empty[NSTAGE], s_full[NQ], p_ready[NQ][2], pv_done[NQ], s_free[2]
    static constexpr int kNumBars = 5 * NQ + 2 * NSTAGE + 2;
    static constexpr int kTmemPtrOff = kBarOff + kNumBars * 8;
    static constexpr int kBytes = kTmemPtrOff + 16;
    static constexpr int kAlloc = kBytes + 1024;                     // slack for the 1024 B alignment
    static_assert(kAlloc <= 232448, "shared memory budget exceeded");
};

// BF16: data type (false = fp16).  POLY: of every 8 column pairs, how many take the polynomial exp2.
// DBG: 0 = product; 1, 2, 4 = timing probes that give wrong results, 5 = cycle counters (tools/fwd_time.py).
template <int D, int NQ, int NSTAGE, bool BF16, int POLY, int DBG>
__global__ void __launch_bounds__(FwdRegs<NQ>::kThreads, 1)
fwd_kernel(const __grid_constant__ CUtensorMap q_map, const __grid_constant__ CUtensorMap k_map,
           const __grid_constant__ CUtensorMap v_map, const FwdParams p) {
    using L = FwdSmem<D, NQ, NSTAGE>;
    using R = FwdRegs<NQ>;
    using T = FwdTmem<D, NQ>;
    constexpr bool kSetRegs = R::kSet;
    constexpr int kSoftmaxWarps = 4 * NQ;
    constexpr int kSBufs = T::kSBufs;
    constexpr int kSub = D / 64;
    constexpr bool kFreeRun = DBG == 4;   // probe: the MMA warp never waits for the softmax warps
    constexpr bool kProbe = DBG == 4 || DBG == 5;
    constexpr bool kTrace = DBG == 6;

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + L::kBarOff;
    auto bar_q_full = [&](int q) { return bar_base + 8u * q; };
    auto bar_kv_full = [&](int s) { return bar_base + 8u * (NQ + s); };
    auto bar_kv_empty = [&](int s) { return bar_base + 8u * (NQ + NSTAGE + s); };
    auto bar_s_full = [&](int q) { return bar_base + 8u * (NQ + 2 * NSTAGE + q); };
    // P is handed to the MMA warp in two halves of 64 keys, so P.V of the first half runs under the second's exp2
    auto bar_p_ready = [&](int q, int half) { return bar_base + 8u * (2 * NQ + 2 * NSTAGE + 2 * q + half); };
    // every P.V of query tile q has completed: P(q) may be overwritten, O(q) may be rescaled / read
    auto bar_pv_done = [&](int q) { return bar_base + 8u * (4 * NQ + 2 * NSTAGE + q); };
    // warpgroup q has loaded its scores into registers: the next Q.K may overwrite that buffer
    auto bar_s_free = [&](int q) { return bar_base + 8u * (5 * NQ + 2 * NSTAGE + q); };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem_gen + L::kTmemPtrOff);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // DBG == 6: prologue / epilogue stamps of CTA (0,0,0): 24 kernel entry, 25 set-up done, 26 last P.V seen by
    // warpgroup 0, 27 its O rows stored, 28 teardown
    const bool stamp = kTrace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0;
    if (stamp) g_fwd_probe[24] = clock64();
    // Heavy (late) query blocks first: under the causal mask they own the most key tiles.
    const int m_block = gridDim.x - 1 - blockIdx.x;
    const int h = blockIdx.y, b = blockIdx.z;
    const int hk = h / p.group;
    const int m0 = m_block * (NQ * kBlockM);
    const int kv_tiles_total = (p.seqlen_k + kBlockN - 1) / kBlockN;
    const int causal_shift = p.seqlen_k - p.seqlen_q;   // query i sees keys <= i + shift

    // Key tiles each query tile needs (causal: up to the diagonal of its last row).
    int n_tiles[NQ];
    int n_max = 0;
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        int n = kv_tiles_total;
        if (p.causal) {
            const int last_row = min(m0 + (q + 1) * kBlockM, p.seqlen_q) - 1;
            const int last_key = last_row + causal_shift;
            n = last_key < 0 ? 0 : min(kv_tiles_total, last_key / kBlockN + 1);
        }
        if (m0 + q * kBlockM >= p.seqlen_q) n = 0;   // tile entirely past the end of Q
        n_tiles[q] = n;
        n_max = max(n_max, n);
    }
    // Q.K products are ordered j-major, q-minor; the k-th one writes score buffer k % kSBufs.
    auto qk_ordinal = [&](int q, int j) {
        int k = 0;
#pragma unroll
        for (int qq = 0; qq < NQ; ++qq) k += min(j, n_tiles[qq]) + ((qq < q && j < n_tiles[qq]) ? 1 : 0);
        return k;
    };
    // inverse of qk_ordinal: (query tile << 16) | key tile
    auto tile_of_ordinal = [&](int ord) {
        const int n0 = n_tiles[0], n1 = n_tiles[NQ - 1], m = min(n0, n1);
        if (ord < 2 * m) return ((ord & 1) << 16) | (ord >> 1);
        return ((n0 > n1 ? 0 : 1) << 16) | (m + ord - 2 * m);
    };
    // Score buffer of tile (q, j): the single shared buffer (two query tiles at head_dim 128), else one buffer
    // per query tile, else (one query tile) alternating buffers.
    auto s_buffer = [&](int q, int j) { return kSBufs == 1 ? 0 : (NQ == 2 ? q : (j & 1)); };
    // Barrier on which the warpgroup of tile (q, j) reports "scores loaded, buffer free".  Every barrier has
    // exactly one waiter at a time and that waiter follows each of its phases in order, which is what a parity
    // wait needs: per warpgroup when the buffer is shared or dedicated, per buffer when one tile alternates.
    auto s_free_index = [&](int q, int j) { return NQ == 2 ? q : (j & 1); };

    if (warp == kSoftmaxWarps && lane == 0) {
        prefetch_tensormap(&q_map);
        prefetch_tensormap(&k_map);
        prefetch_tensormap(&v_map);
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            mbar_init(bar_q_full(q), 1);
            mbar_init(bar_s_full(q), 1);
            mbar_init(bar_p_ready(q, 0), 4);   // one elected arrival per softmax warp of the tile's warpgroup
            mbar_init(bar_p_ready(q, 1), 4);
            mbar_init(bar_pv_done(q), 1);
        }
        mbar_init(bar_s_free(0), 4);
        mbar_init(bar_s_free(1), 4);
#pragma unroll
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(bar_kv_full(s), 1);
            mbar_init(bar_kv_empty(s), NQ);   // one arrival per query tile's MMA warp
        }
        fence_barrier_init();
        // The first loads need nobody's consent (the ring starts empty): issue them before the CTA-wide sync so
        // that their latency overlaps the TMEM allocation and the rest of the prologue.
        // Issue order Q0, K0, then the rest: the first Q.K needs exactly those two tiles (measured on CTA (0,0,0):
        // 5250 cycles from set-up to the first Q.K issue, all of it waiting for these loads).
        auto load_q = [&](int q) {
            if (n_tiles[q] == 0) return;
            mbar_arrive_expect_tx(bar_q_full(q), L::kTileBytes);
#pragma unroll
            for (int s = 0; s < kSub; ++s)
                tma_load_4d(smem_base + L::kQOff + q * L::kTileBytes + s * L::kSubTileBytes, &q_map, bar_q_full(q),
                            64 * s, h, m0 + q * kBlockM, b);
        };
        auto load_kv = [&](int i) {
            mbar_arrive_expect_tx(bar_kv_full(i), L::kTileBytes);
            const CUtensorMap* map = (i & 1) ? &v_map : &k_map;
#pragma unroll
            for (int s = 0; s < kSub; ++s)
                tma_load_4d(smem_base + L::kRingOff + i * L::kTileBytes + s * L::kSubTileBytes, map, bar_kv_full(i),
                            64 * s, hk, (i >> 1) * kBlockN, b);
        };
        const int n_pre = min(NSTAGE, 2 * n_max);
        load_q(0);
        if (p.load_order && n_pre > 0) load_kv(0);
#pragma unroll
        for (int q = 1; q < NQ; ++q) load_q(q);
        for (int i = p.load_order ? 1 : 0; i < n_pre; ++i) load_kv(i);
    }
    if (warp == kSoftmaxWarps + 1) {
        tmem_alloc(smem_base + L::kTmemPtrOff, T::kCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    if (stamp) g_fwd_probe[25] = clock64();

    // NQ == 2 runs three full warpgroups (the third holds the TMA and MMA warps plus two idle
    // warps) so the register file can be re-split: the softmax warpgroups take what the others give up.
    if (warp == kSoftmaxWarps) {
        // =============================================================== TMA producer
        if constexpr (kSetRegs) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R::kOther));
        // whole warp in the loop (uniform control flow), one elected lane issues; Q and the first NSTAGE
        // tiles were issued in the prologue
        for (int i = NSTAGE; i < 2 * n_max; ++i) {
            const int slot = i % NSTAGE, use = i / NSTAGE;
            if (!__all_sync(0xffffffffu, wait_barrier(bar_kv_empty(slot), (use & 1) ^ 1, p.error_flag, 1))) break;
            if (elect_one()) {
                mbar_arrive_expect_tx(bar_kv_full(slot), L::kTileBytes);
                const CUtensorMap* map = (i & 1) ? &v_map : &k_map;
#pragma unroll
                for (int s = 0; s < kSub; ++s)
                    tma_load_4d(smem_base + L::kRingOff + slot * L::kTileBytes + s * L::kSubTileBytes, map,
                                bar_kv_full(slot), 64 * s, hk, (i >> 1) * kBlockN, b);
            }
            __syncwarp();
        }
    } else if (warp > kSoftmaxWarps && warp <= kSoftmaxWarps + NQ) {
        // =============================================================== tcgen05 issuers, one warp per query tile
        if constexpr (kSetRegs) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R::kOther));
        // The whole warp runs the loop (uniform control flow keeps descriptors in uniform registers); one elected
        // lane issues.  Descriptors are built once per operand tile and advanced by compile-time constants.
        constexpr uint32_t idesc_qk = mma_idesc(BF16, 0, 0, kBlockM, kBlockN);
        constexpr uint32_t idesc_pv = mma_idesc(BF16, 0, 1, kBlockM, D);
        // Q, K: K-major, rows of 128 B, 8-row groups 1024 B apart.  V: MN-major (head_dim
        // contiguous), 8-key groups 1024 B apart, next 64 head_dim columns one sub-tile further.
        constexpr uint64_t desc_kmajor = smem_desc_sw128(16, 1024);
        constexpr uint64_t desc_v = smem_desc_sw128(L::kSubTileBytes, 1024);
        const uint64_t ring_k_desc = smem_desc_at(desc_kmajor, smem_base + L::kRingOff);
        const uint64_t ring_v_desc = smem_desc_at(desc_v, smem_base + L::kRingOff);
        bool ok = true;
        const int q = NQ == 1 ? 0 : warp - (kSoftmaxWarps + 1);   // the query tile this warp issues for
        const int n_q = q == 0 ? n_tiles[0] : n_tiles[NQ - 1];
        const uint64_t q_desc = smem_desc_at(desc_kmajor, smem_base + L::kQOff + q * L::kTileBytes);
        // start-address field is in 16-byte units and never carries out of its 14 bits (smem < 256 KB)
        auto issue_qk = [&](int slot, int buf) {
            const uint64_t k_desc = ring_k_desc + (uint64_t)(slot * (L::kTileBytes >> 4));
            const uint32_t s_tmem = tmem_base + T::kSOff + buf * kBlockN;
#pragma unroll
            for (int kc = 0; kc < D / 16; ++kc) {
                const uint32_t off16 = ((kc >> 2) * L::kSubTileBytes + (kc & 3) * 32) >> 4;
                mma_ss(s_tmem, q_desc + off16, k_desc + off16, idesc_qk, kc > 0);
            }
        };
        auto issue_pv = [&](int slot, int half, bool accumulate) {   // keys [64 * half, 64 * half + 64)
            const uint64_t v_desc = ring_v_desc + (uint64_t)(slot * (L::kTileBytes >> 4));
            const uint32_t p_tmem = tmem_base + T::kPOff + q * (kBlockN / 2);
            const uint32_t o_tmem = tmem_base + T::kOOff + q * D;
#pragma unroll
            for (int kk = 0; kk < kBlockN / 32; ++kk) {
                const int kc = half * (kBlockN / 32) + kk;
                mma_ts(o_tmem, p_tmem + kc * 8, v_desc + (uint64_t)(kc * (2048 >> 4)), idesc_pv, accumulate || kc > 0);
            }
        };
        auto wait_all = [&](uint32_t bar, uint32_t parity, int code) {
            return __all_sync(0xffffffffu, wait_barrier(bar, parity, p.error_flag, code)) != 0;
        };
        // A ring slot is handed back with one arrival per query tile; where the other tile does not use key tile
        // j (causal tail, or a query tile past the end of Q) this warp arrives for it as well.
        auto users = [&](int j) {
            int u = 0;
#pragma unroll
            for (int qq = 0; qq < NQ; ++qq) u += j < n_tiles[qq] ? 1 : 0;
            return u;
        };
        long long mt_p = 0, mt_kv = 0, mt_sfree = 0, mt_total = clock64(), c0 = 0;
        const bool tre = kTrace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
        const bool tr = tre && lane == 0;
        if (n_q > 0) ok = wait_all(bar_q_full(q), 0, 3);
        for (int j = 0; j <= n_q && ok; ++j) {
            // ---- S(q, j) = Q_q K_j^T
            if (j < n_q) {
                const int ki = 2 * j, kslot = ki % NSTAGE;
                const int ord = qk_ordinal(q, j), buf = s_buffer(q, j);
                trace_event(tr, q, j, 7);
                if constexpr (kProbe) c0 = clock64();
                ok = wait_all(bar_kv_full(kslot), (ki / NSTAGE) & 1, 2);
                if constexpr (kProbe) { mt_kv += clock64() - c0; c0 = clock64(); }
                trace_event(tr, q, j, 0);
                if constexpr (!kFreeRun) {
                    // the buffer's previous scores must have been loaded into registers by their warpgroup
                    if constexpr (NQ == 2 && kSBufs == 1) {
                        if (ord >= 1) {
                            const int prev = tile_of_ordinal(ord - 1);
                            ok = ok && wait_all(bar_s_free(prev >> 16), prev & 1, 9);
                        }
                    } else if constexpr (NQ == 2) {
                        if (j >= 1) ok = ok && wait_all(bar_s_free(q), (j - 1) & 1, 9);
                    } else {
                        if (j >= 2) ok = ok && wait_all(bar_s_free(j & 1), ((j >> 1) - 1) & 1, 9);
                    }
                }
                if constexpr (kProbe) mt_sfree += clock64() - c0;
                tc_fence_after();
                trace_event(tr, q, j, 1);
                if (ok && elect_one()) {
                    issue_qk(kslot, buf);
                    trace_event(tre, q, j, 13);
                    tc_commit(bar_s_full(q));
                    for (int u = users(j); u <= NQ; ++u) tc_commit(bar_kv_empty(kslot));
                }
                __syncwarp();
                trace_event(tr, q, j, 2);
            }
            // ---- O(q) += P(q, j-1) V_{j-1}, in two halves of 64 keys
            if (ok && j > 0) {
                const int vi = 2 * j - 1, vslot = vi % NSTAGE;
                if constexpr (kProbe) c0 = clock64();
                ok = wait_all(bar_kv_full(vslot), (vi / NSTAGE) & 1, 4);
                if constexpr (kProbe) { mt_kv += clock64() - c0; c0 = clock64(); }
                trace_event(tr, q, j - 1, 14);
                if constexpr (!kFreeRun) ok = ok && wait_all(bar_p_ready(q, 0), (j - 1) & 1, 5);
                tc_fence_after();
                trace_event(tr, q, j - 1, 3);
                if (ok && elect_one()) issue_pv(vslot, 0, j > 1);
                __syncwarp();
                trace_event(tr, q, j - 1, 4);
                if constexpr (!kFreeRun) ok = ok && wait_all(bar_p_ready(q, 1), (j - 1) & 1, 5);
                if constexpr (kProbe) mt_p += clock64() - c0;
                tc_fence_after();
                trace_event(tr, q, j - 1, 5);
                if (ok && elect_one()) {
                    issue_pv(vslot, 1, true);
                    trace_event(tre, q, j - 1, 15);
                    tc_commit(bar_pv_done(q));
                    for (int u = users(j - 1); u <= NQ; ++u) tc_commit(bar_kv_empty(vslot));
                }
                __syncwarp();
                trace_event(tr, q, j - 1, 6);
            }
        }
        if (kProbe && q == 0 && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
            g_fwd_probe[16] = mt_p; g_fwd_probe[17] = mt_sfree; g_fwd_probe[18] = mt_kv;
            g_fwd_probe[19] = clock64() - mt_total; g_fwd_probe[20] = n_q;
        }
    } else if (warp < kSoftmaxWarps) {
        // =============================================================== softmax warpgroups
        // Warpgroup q serves query tile q; thread i owns row i of the tile (TMEM lane i).
        if constexpr (kSetRegs) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(R::kSoftmax));
        constexpr int kCols = kBlockN;
        const int q = warp >> 2;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const int row = m0 + q * kBlockM + row_in_tile;
        const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t p_tmem = lane_base + T::kPOff + q * (kBlockN / 2);   // my packed P columns
        const uint32_t o_tmem = lane_base + T::kOOff + q * D;
        const int n_q = q == 0 ? n_tiles[0] : n_tiles[NQ - 1];
        const int key_limit = p.causal ? row + causal_shift : INT_MAX;   // last visible key of this row
        const float sl2 = p.scale_log2;
        const uint64_t sl2_2 = pack_f2(sl2, sl2);

        float m_used = -INFINITY;   // reference maximum (raw score units) the exponentials are taken against
        uint64_t l_acc[2] = {0ull, 0ull};   // four partial row sums as two f32x2 accumulators
        bool ok = true;

        const bool probe = DBG == 5 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (warp & 3) == 0;
        long long pt_wait = 0, pt_max = 0, pt_exp = 0, pt_pv = 0, pt0 = 0, pt1 = 0, pt2 = 0;
        const bool tr = kTrace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (warp & 3) == 0;
        // With a score buffer of their own (head_dim 64) nothing couples the two warpgroups, and started together
        // they stay in lockstep: both in the exp2 loop (MUFU contended), then both in the load + max phase (MUFU
        // idle).  A head start for warpgroup 0 puts one's exp2 loop under the other's max phase.
        if (NQ == 2 && q == 1 && p.stagger_cycles > 0) {
            const long long t0 = clock64();
            while (clock64() - t0 < p.stagger_cycles) __nanosleep(64);
        }
        for (int j = 0; j < (kFreeRun ? 0 : n_q) && ok; ++j) {
            if constexpr (DBG == 5) pt0 = clock64();
            const int buf = s_buffer(q, j);
            const uint32_t s_tmem = lane_base + T::kSOff + buf * kBlockN;
            ok = wait_barrier(bar_s_full(q), j & 1, p.error_flag, 7);
            tc_fence_after();
            if constexpr (DBG == 5) pt1 = clock64();
            trace_event(tr, q, j, 8);
            uint32_t su[kCols];
            if constexpr (DBG != 2) {
#pragma unroll
                for (int c = 0; c < kCols; c += 32) tmem_ld32(s_tmem + c, &su[c]);
                // The buffer goes back the moment the loads have landed: its turnaround sits on the critical
                // cycle of the pipeline (two query tiles share one score buffer at head_dim 128).
                tmem_wait_ld();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_s_free(s_free_index(q, j)));
                trace_event(tr, q, j, 9);
            }
            const int key0 = j * kBlockN;
            const bool need_mask = (key0 + kBlockN - 1 > m0 + q * kBlockM + causal_shift && p.causal) ||
                                   (key0 + kBlockN > p.seqlen_k);
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
            if constexpr (DBG != 2) scan_half(0);
            if constexpr (DBG == 2) {      // pipeline probe: no softmax work at all
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_s_free(s_free_index(q, j)));
                if (j > 0) ok = wait_barrier(bar_pv_done(q), (j - 1) & 1, p.error_flag, 8);
                if (lane == 0) {
                    mbar_arrive(bar_p_ready(q, 0));
                    mbar_arrive(bar_p_ready(q, 1));
                }
                continue;
            }
            scan_half(kCols / 2);
            const float mx_all = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));

            // P(q, j-1) is still being read by the tensor pipe, and O(q) still accumulated into, until pv_done
            bool pv_waited = j == 0;
            auto wait_pv = [&]() {
                if (!pv_waited) {
                    long long c0 = 0;
                    if constexpr (DBG == 5) c0 = clock64();
                    ok = wait_barrier(bar_pv_done(q), (j - 1) & 1, p.error_flag, 8);
                    tc_fence_after();
                    pv_waited = true;
                    if constexpr (DBG == 5) pt_pv += clock64() - c0;
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
            if constexpr (DBG == 5) pt2 = clock64();
            trace_event(tr, q, j, 10);
            // ---- P = exp2(S * scale_log2 - m): MUFU for most pairs, an FMA-pipe polynomial for POLY of every 8
            // pairs on tiles without masked (-inf) entries; P leaves for TMEM in blocks of 64 keys.
            auto exp_and_store = [&](auto poly_tag) {
                constexpr int kPoly = decltype(poly_tag)::value;
                constexpr int kPairs = kCols / 2;
                constexpr int kAhead = 3;            // pairs between the scale FFMA2 and its MUFUs
                uint64_t xq[kAhead];                 // scaled scores waiting for their exp2
                auto scale = [&](int i) {
                    return ffma2_v(pack_f2(__uint_as_float(su[2 * i]), __uint_as_float(su[2 * i + 1])), sl2_2, neg_m2);
                };
#pragma unroll
                for (int i = 0; i < kAhead; ++i) xq[i] = scale(i);
                uint32_t pk[32];
#pragma unroll
                for (int i = 0; i < kPairs; ++i) {
                    float x0, x1, p0, p1;
                    unpack_f2(xq[i % kAhead], x0, x1);
                    if (i + kAhead < kPairs) xq[i % kAhead] = scale(i + kAhead);
                    if constexpr (DBG == 1) {
                        p0 = x0; p1 = x1;
                    } else if ((i & 7) >= 8 - kPoly) {
                        unpack_f2(exp2_poly2(x0, x1), p0, p1);
                    } else {
                        p0 = ex2_v(x0); p1 = ex2_v(x1);
                    }
                    l_acc[i & 1] = fadd2_v(l_acc[i & 1], pack_f2(p0, p1));
                    pk[i & 31] = pack_pair<BF16>(p0, p1);
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
                        if (lane == 0) {
                            if (i == kPairs - 1) mbar_arrive(bar_p_ready(q, kPairs / 32 - 1));
                            else mbar_arrive(bar_p_ready(q, (i >> 5) - 1));
                        }
                        trace_event(tr, q, j, i == kPairs - 1 ? 12 : 11);
                    }
                }
            };
            if (POLY > 0 && !need_mask) exp_and_store(std::integral_constant<int, POLY>{});
            else exp_and_store(std::integral_constant<int, 0>{});
            if constexpr (DBG == 5) {
                const long long pt3 = clock64();
                pt_wait += pt1 - pt0; pt_max += pt2 - pt1; pt_exp += pt3 - pt2;
            }
        }
        if (probe) {
            g_fwd_probe[8 * q + 0] = pt_wait; g_fwd_probe[8 * q + 1] = pt_max; g_fwd_probe[8 * q + 2] = pt_exp;
            g_fwd_probe[8 * q + 3] = n_q; g_fwd_probe[8 * q + 4] = pt_pv;
        }

        // ---- epilogue: O / l -> global, LSE
        float l0, l1, l2, l3;
        unpack_f2(l_acc[0], l0, l1);
        unpack_f2(l_acc[1], l2, l3);
        const float l_sum = (l0 + l1) + (l2 + l3);
        const bool row_ok = row < p.seqlen_q;
        const float inv_l = l_sum > 0.f ? 1.f / l_sum : 0.f;
        uint8_t* o_row = reinterpret_cast<uint8_t*>(p.o) +
                         2 * ((int64_t)b * p.o_bs + (int64_t)row * p.o_rs + (int64_t)h * p.o_hs);
        if constexpr (kFreeRun) {   // probe: nobody followed the phases of pv_done, walk through all of them
            for (int j = 0; j + 1 < n_q && ok; ++j) ok = wait_barrier(bar_pv_done(q), j & 1, p.error_flag, 10);
        }
        if (n_q > 0 && ok) ok = wait_barrier(bar_pv_done(q), (n_q - 1) & 1, p.error_flag, 10);
        tc_fence_after();
        if (stamp) g_fwd_probe[26] = clock64();
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            if (n_q > 0) {
                tmem_ld32(o_tmem + 32 * c, o);
                tmem_wait_ld();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) o[i] = 0u;
            }
            uint32_t packed[16];
#pragma unroll
            for (int i = 0; i < 16; ++i)
                packed[i] = pack_pair<BF16>(__uint_as_float(o[2 * i]) * inv_l, __uint_as_float(o[2 * i + 1]) * inv_l);
            if (row_ok) {
                uint4* dst = reinterpret_cast<uint4*>(o_row + 64 * c);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    dst[i] = make_uint4(packed[4 * i], packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
            }
        }
        if (row_ok && p.lse) {
            const float lse = l_sum > 0.f ? (m_used * sl2 + log2f(l_sum)) * kLn2 : -INFINITY;
            p.lse[((int64_t)b * p.heads_q + h) * p.seqlen_q + row] = lse;
        }
        if (stamp) g_fwd_probe[27] = clock64();
    } else {
        if constexpr (kSetRegs) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(R::kOther));   // idle warps of the third warpgroup
    }

    // ---- teardown: every tcgen05 operation of this CTA has completed once the softmax warps are here
    tc_fence_before();
    __syncthreads();
    if (warp == kSoftmaxWarps + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, T::kCols);
    }
    if (stamp) g_fwd_probe[28] = clock64();
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

template <int D, int NQ, int NSTAGE, bool BF16, int POLY, int DBG>
int launch_instance(const CUtensorMap& qm, const CUtensorMap& km, const CUtensorMap& vm, const FwdParams& p,
                    int batch, cudaStream_t stream) {
    using L = FwdSmem<D, NQ, NSTAGE>;
    auto kern = fwd_kernel<D, NQ, NSTAGE, BF16, POLY, DBG>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] {
        attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kAlloc);
    });
    SFA_CUDA_CHECK(attr_err);
    const int m_blocks = (p.seqlen_q + NQ * kBlockM - 1) / (NQ * kBlockM);
    dim3 grid(m_blocks, p.heads_q, batch);
    kern<<<grid, FwdRegs<NQ>::kThreads, L::kAlloc, stream>>>(qm, km, vm, p);
    SFA_CUDA_CHECK(cudaGetLastError());
    count_launch();
    return SFA_OK;
}

// Host launcher: the tile configuration per (head_dim, rows, dtype) - what the reference's Traits<> template
// (src/traits.h:7-26) selects at compile time.  Ring depth: as many 128-key tiles as fit beside the Q tiles.
template <bool BF16>
int launch_dtype(const CUtensorMap& qm, const CUtensorMap& km, const CUtensorMap& vm, const FwdParams& p, int batch,
                 int head_dim, int nq, cudaStream_t stream) {
    // Tuning / probe overrides (tools/fwd_time.py): SFA_FWD_POLY in {0,2,3}, SFA_FWD_DBG probes (bf16, head_dim 128
    // and fp16, head_dim 64, two query tiles only), SFA_FWD_STAGES in {4,5} at head_dim 128.
    static const int poly = [] { const char* e = getenv("SFA_FWD_POLY"); return e ? atoi(e) : -1; }();
    static const int dbg = [] { const char* e = getenv("SFA_FWD_DBG"); return e ? atoi(e) : 0; }();
    static const int stages = [] { const char* e = getenv("SFA_FWD_STAGES"); return e ? atoi(e) : 0; }();
#define SFA_LAUNCH(D_, NQ_, NS_, POLY_, DBG_) \
    return launch_instance<D_, NQ_, NS_, BF16, POLY_, DBG_>(qm, km, vm, p, batch, stream)
    if (head_dim == 128) {
        if (nq == 1) SFA_LAUNCH(128, 1, 4, 0, 0);
        if constexpr (BF16) {
            if (dbg == 1) SFA_LAUNCH(128, 2, 4, 0, 1);
            if (dbg == 2) SFA_LAUNCH(128, 2, 4, 0, 2);
            if (dbg == 4) SFA_LAUNCH(128, 2, 4, 0, 4);
            if (dbg == 5) SFA_LAUNCH(128, 2, 4, 0, 5);
            if (dbg == 6) SFA_LAUNCH(128, 2, 4, 0, 6);
            if (poly == 2) SFA_LAUNCH(128, 2, 4, 2, 0);
            if (poly == 3) SFA_LAUNCH(128, 2, 4, 3, 0);
            if (stages == 5) SFA_LAUNCH(128, 2, 5, 0, 0);
        }
        SFA_LAUNCH(128, 2, 4, 0, 0);
    }
    if (nq == 1) SFA_LAUNCH(64, 1, 6, 0, 0);
    if constexpr (!BF16) {
        if (dbg == 5) SFA_LAUNCH(64, 2, 6, 0, 5);
        if (poly == 0) SFA_LAUNCH(64, 2, 6, 0, 0);
        if (poly == 4) SFA_LAUNCH(64, 2, 6, 4, 0);
    }
    // head_dim 64 is MUFU-bound (a tile's exp2 takes twice its MMA time): a quarter of the exponentials go to
    // the FMA-pipe polynomial (measured on C2: 0.404 -> 0.390 ms)
    SFA_LAUNCH(64, 2, 6, 2, 0);
#undef SFA_LAUNCH
}

}  // namespace

int fwd_error_flag_read_v1(int* value) {
    SFA_CUDA_CHECK(cudaMemcpyFromSymbol(value, g_fwd_error, sizeof(int)));
    return SFA_OK;
}
int fwd_probe_read_v1(unsigned long long* out, int n) {
    if (n > 32 + 2048) n = 32 + 2048;
    SFA_CUDA_CHECK(cudaMemcpyFromSymbol(out, g_fwd_probe, n * sizeof(unsigned long long)));
    return SFA_OK;
}
int fwd_error_flag_clear_v1() {
    const int zero = 0;
    SFA_CUDA_CHECK(cudaMemcpyToSymbol(g_fwd_error, &zero, sizeof(int)));
    return SFA_OK;
}

int launch_fwd_v1(const sfa_fwd_args& a, cudaStream_t stream) {
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
    const DeviceInfo* dev = current_device();
    if (!dev) return SFA_ERR_NO_DEVICE;

    alignas(64) CUtensorMap qm, km, vm;
    int rc;
    // seqlen_k == 0 still needs valid maps; a one-row extent keeps the encoder happy and is never read.
    const int rows_k = a.seqlen_k > 0 ? a.seqlen_k : 1;
    if ((rc = make_map(&qm, a.q, a.dtype, a.head_dim, a.heads_q, a.seqlen_q, a.batch, a.q_head_stride, a.q_row_stride, a.q_batch_stride))) return rc;
    if ((rc = make_map(&km, a.k, a.dtype, a.head_dim, a.heads_kv, rows_k, a.batch, a.k_head_stride, a.k_row_stride, a.k_batch_stride))) return rc;
    if ((rc = make_map(&vm, a.v, a.dtype, a.head_dim, a.heads_kv, rows_k, a.batch, a.v_head_stride, a.v_row_stride, a.v_batch_stride))) return rc;

    FwdParams p;
    p.o = a.o; p.lse = a.lse;
    p.o_bs = a.o_batch_stride; p.o_rs = a.o_row_stride; p.o_hs = a.o_head_stride;
    p.seqlen_q = a.seqlen_q; p.seqlen_k = a.seqlen_k; p.heads_q = a.heads_q; p.group = a.heads_q / a.heads_kv;
    p.causal = a.causal ? 1 : 0; p.is_bf16 = a.dtype == SFA_BF16;
    const float scale = a.softmax_scale > 0.f ? a.softmax_scale : 1.f / sqrtf((float)a.head_dim);
    p.scale_log2 = scale * kLog2e;
    {
        const char* e = getenv("SFA_FWD_STAGGER");
        p.stagger_cycles = e ? atoi(e) : (a.head_dim == 64 ? kStaggerCycles64 : 0);
        const char* lo = getenv("SFA_FWD_LOADORDER");
        p.load_order = lo ? atoi(lo) : 1;
    }
    void* flag_addr = nullptr;
    SFA_CUDA_CHECK(cudaGetSymbolAddress(&flag_addr, g_fwd_error));
    p.error_flag = reinterpret_cast<int*>(flag_addr);

    static int forced_nq = [] { const char* e = getenv("SFA_FWD_NQ"); return e ? atoi(e) : 0; }();
    const int nq = forced_nq == 1 || forced_nq == 2 ? forced_nq : (a.seqlen_q > kBlockM ? 2 : 1);
    return a.dtype == SFA_BF16 ? launch_dtype<true>(qm, km, vm, p, a.batch, a.head_dim, nq, stream)
                               : launch_dtype<false>(qm, km, vm, p, a.batch, a.head_dim, nq, stream);
}

}  // namespace sfa
