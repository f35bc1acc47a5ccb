// Host-buffer entry points: pinned host tensors in, pinned host tensors out (sfa_fwd_host, sfa_decode_host).
//
// The reference's host layer moves its inputs with blocking cudaMemcpy calls and a cudaDeviceSynchronize per call
// (src/flash_attn.cu:973-975,986) and its caller holds everything on the device; a caller whose tensors live in
// host memory pays the PCIe time on top.  Here the independent units of a call are cut into pieces and the
// host->device copy of piece i+1, the kernel of piece i and the device->host copy of piece i-1 run at the
// same time on three streams (PCIe is full duplex), so a call costs about max(upload, download) + one piece.
//
// Prefill pieces are (batch entry, group of kv-heads, block of query rows).  With every head in a piece the host rows
// are contiguous, so each piece moves as plain 1-D copies (a caller that owns a slice of heads - a rank of the
// multi-GPU driver - passes strided tensors and the same code issues pitched 2-D copies).  Under the causal mask a
// piece of rows [r0, r1) needs keys [0, r1 + shift) only, so K/V travel incrementally, piece by piece, and the first
// kernel starts after the first piece's upload; the kernel's bottom-right aligned mask makes "rows [r0, r1) against
// the key prefix" exactly the rows of the full problem.
//
// What a call costs beyond its upload is the last piece's kernel and download plus ~28 us of copy-engine idle time
// per piece (tools/e2e_pieces.py: C3 31.8 ms with 16 uniform pieces, 33.9 with 64, 38.2 with 256).  So the default
// plan is FEW LARGE pieces with geometrically smaller ones at the end of the call, and the staging buffers form rings
// (4 deep for Q / O / LSE, 3 for a unit's K/V) so that the upload never waits for the kernel two pieces back.
#include "internal.h"

#include <algorithm>
#include <atomic>
#include <math.h>
#include <mutex>
#include <vector>

namespace sfa {
namespace {

constexpr int kRing = 4;      // staging slots per kind (Q, O, LSE): the upload may run this many pieces ahead of the kernels
constexpr int kKvRing = 3;    // K/V staging slots (one per unit in flight)

struct Pipeline {
    std::mutex mu;                   // one call at a time enqueues on a device's pipeline
    bool ready = false;
    bool used = false;               // ev_last has been recorded: the next call orders itself behind it
    cudaStream_t s_in = nullptr, s_out = nullptr, s_comp[3] = {};   // s_comp: extra kernel streams (prefill pieces take turns)
    cudaEvent_t ev_start = nullptr, ev_last = nullptr, ev_in[kRing] = {}, ev_comp[kRing] = {}, ev_out[kRing] = {}, ev_kv_done[kKvRing][4] = {};
    // staging rings; grown on demand, outgrown buffers are retired (work may be queued on them)
    void* buf[16] = {};
    size_t cap[16] = {};
    std::vector<void*> retired;
    std::vector<cudaEvent_t> trace;   // host_trace: 5 timing events per piece of the last call (see fwd_host_trace)
};
enum { kQ0 = 0, kO0 = kRing, kL0 = 2 * kRing, kKV0 = 3 * kRing, kNumBufs = 3 * kRing + kKvRing };   // Q, O, LSE, K+V (decode: q+seqlens, o+lse, -, caches)
static_assert(kNumBufs <= 16, "staging table too small");

std::atomic<int> g_last_plan_pieces{0};   // pieces of the most recent sfa_fwd_host plan (tests, tools)
std::mutex g_pipes_mu;
std::vector<Pipeline*> g_pipes;

Pipeline* pipeline_for_current_device() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lk(g_pipes_mu);
    if ((int)g_pipes.size() <= dev) g_pipes.resize(dev + 1, nullptr);
    if (!g_pipes[dev]) g_pipes[dev] = new Pipeline();
    return g_pipes[dev];
}

int pipeline_init(Pipeline& p) {
    if (p.ready) return SFA_OK;
    SFA_CUDA_CHECK(cudaStreamCreateWithFlags(&p.s_in, cudaStreamNonBlocking));
    SFA_CUDA_CHECK(cudaStreamCreateWithFlags(&p.s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 3; ++i) SFA_CUDA_CHECK(cudaStreamCreateWithFlags(&p.s_comp[i], cudaStreamNonBlocking));
    std::vector<cudaEvent_t*> evs = {&p.ev_start, &p.ev_last};
    for (int i = 0; i < kRing; ++i) { evs.push_back(&p.ev_in[i]); evs.push_back(&p.ev_comp[i]); evs.push_back(&p.ev_out[i]); }
    for (int i = 0; i < kKvRing; ++i)
        for (int j = 0; j < 4; ++j) evs.push_back(&p.ev_kv_done[i][j]);
    for (cudaEvent_t* e : evs) SFA_CUDA_CHECK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    p.ready = true;
    return SFA_OK;
}

int ensure(Pipeline& p, int which, size_t bytes) {
    if (p.cap[which] >= bytes) return SFA_OK;
    if (p.buf[which]) p.retired.push_back(p.buf[which]);
    p.buf[which] = nullptr;
    p.cap[which] = 0;
    const size_t want = bytes + bytes / 8;
    SFA_CUDA_CHECK(cudaMalloc(&p.buf[which], want));
    p.cap[which] = want;
    return SFA_OK;
}

// rows x width_bytes between host and device; one plain copy when both sides are dense
int copy_rows(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width, size_t rows,
              cudaMemcpyKind kind, cudaStream_t s) {
    if (rows == 0 || width == 0) return SFA_OK;
    if (dst_pitch == width && src_pitch == width) SFA_CUDA_CHECK(cudaMemcpyAsync(dst, src, width * rows, kind, s));
    else SFA_CUDA_CHECK(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width, rows, kind, s));
    return SFA_OK;
}

int check_pinned(const void* p, const char* what) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess || at.type != cudaMemoryTypeHost) {
        cudaGetLastError();
        return fail(SFA_ERR_INVALID, std::string(what) + " must point to pinned host memory (cudaHostAlloc / cudaHostRegister)");
    }
    return SFA_OK;
}

}  // namespace

// ---- plan: pieces = (batch entry, group of kv-heads, block of query rows), in issue order
// With many batch entries a unit is one entry with ALL its heads (contiguous on the host).  With few entries (one
// long sequence) the heads are cut into groups first: under the causal mask the work of a block of rows grows
// with its position while its upload does not, so row blocks alone leave the heavy last blocks' kernels behind
// the end of the upload (C5, one sequence: 38.9 ms against a 30.1 ms copy floor); head groups are independent and
// uniform.  A group keeps at least 1 KB per row and tensor (pitched DMAs slow down below that).
std::vector<HostPiece> plan_fwd_host(int B, int Sq, int Sk, int Hq, int Hkv, int D, bool causal, int want_pieces) {
    std::vector<HostPiece> plan;
    if (B <= 0 || Sq <= 0 || Sk < 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv || D <= 0) return plan;
    const size_t es = 2;
    const double in_bytes = es * ((double)B * Sq * Hq * D + 2.0 * B * Sk * Hkv * D);
    const bool explicit_pieces = want_pieces > 0;
    const HostTuning& tune = host_tuning();
    if (want_pieces <= 0) {
        want_pieces = 8;                                          // + the tapered end below
        const int by_size = (int)(in_bytes / tune.piece_bytes);   // pieces of at least 8 MB keep the DMAs efficient
        if (want_pieces > by_size) want_pieces = by_size > B ? by_size : B;
    }
    const int hg_min = (int)(1024 / (D * es)) > 0 ? (int)(1024 / (D * es)) : 1;
    int n_hg = 1;
    if (B < 4) {
        n_hg = Hkv / hg_min;
        if (n_hg > want_pieces / B) n_hg = want_pieces / B;
        if (n_hg < 1) n_hg = 1;
    }
    const int hg = (Hkv + n_hg - 1) / n_hg;               // kv-heads per group
    n_hg = (Hkv + hg - 1) / hg;
    const int grp = Hq / Hkv;
    // Estimated kernel time against upload time.  A "heavy" call (C5: 29 ms of kernels, 29 ms of copies) is bound by
    // the SUM of its kernels, and small pieces of causal work do not fill the GPU (2 kv-heads x 8192 rows are 64 work
    // items for 148 SMs; a tapered C5 plan measured 34.4 - 34.6 ms like the uniform one, its 2-head units taking 3.1
    // and 3.6 ms of kernels against 1.95 for half of a 4-head unit): heavy calls keep uniform pieces.  A light call
    // (C3: 3.5 ms of kernels) only needs a short tail.
    const double flops = 4.0 * B * Hq * (double)Sq * Sk * D * (causal ? 0.5 : 1.0);
    bool heavy = flops / 1.1e15 > 0.4 * in_bytes / 50e9;
    if (tune.heavy >= 0) heavy = tune.heavy != 0;
    const bool taper = !explicit_pieces && tune.taper != 0 && !heavy;
    std::vector<int> groups;                               // kv-heads per head group of one batch entry
    {
        // light calls: half groups at the two ends (nothing computes during the first upload, nothing uploads during
        // the last kernel)
        const int edge = hg / 2;
        if (taper && n_hg >= 3 && edge >= 1 && Hkv - 2 * edge >= hg) {
            groups.push_back(edge);
            for (int left = Hkv - 2 * edge; left > 0; left -= hg) groups.push_back(left < hg ? left : hg);
            groups.push_back(edge);
        } else {
            for (int left = Hkv; left > 0; left -= hg) groups.push_back(left < hg ? left : hg);
        }
    }
    const int n_units = B * (int)groups.size();
    // Row blocks.  Light calls: about 8 pieces in all.  Heavy calls: ONE WAVE of work items per kernel - a piece of
    // hqn query heads x rows is hqn * rows / 256 items for the 148 persistent CTAs, and a launch of 172 items (4 heads x
    // 11008 rows) runs a second, nearly empty wave: C5 on one GPU with two kernel streams, pieces of 4 heads x
    // 32768 / 16384 / 11008 / 8192 / 5632 / 4096 rows: 33.5 / 34.1 / 37.3 / 32.8 / 36.0 / 38.0 ms; a rank's half of C5
    // 18.7 / 18.4 / 19.0 / 17.7 / 19.7 / 30.2 (profiles/r2_ks_sweep.txt).  At most ~96 pieces per call.
    const int target = (n_hg > 1 && !explicit_pieces) ? 8 : want_pieces;
    int n_r = (target + B * n_hg - 1) / (B * n_hg);
    if (heavy && !explicit_pieces) {
        int blocks = tune.wave_items / (hg * grp);
        if (blocks < 1) blocks = 1;
        n_r = (Sq + blocks * 256 - 1) / (blocks * 256);
        const int cap = 96 / (B * n_hg) > 1 ? 96 / (B * n_hg) : 1;
        if (n_r > cap) n_r = cap;
    }
    int rows_per = (Sq + n_r - 1) / n_r;
    rows_per = (rows_per + 255) / 256 * 256;              // whole 256-row blocks of the kernel
    n_r = (Sq + rows_per - 1) / rows_per;
    // row cuts of a unit: uniform blocks; the last block of the LAST unit of a tapered plan is cut again into 1/2,
    // 1/4, 1/8, 1/8 (the exposed tail is then the kernel and download of a small piece).  A refined piece keeps at
    // least `taper_bytes` of input.
    auto cuts_of = [&](bool last_unit, size_t bytes_per_row) {
        std::vector<int> c;
        for (int r = 0; r < n_r; ++r) c.push_back(r * rows_per);
        c.push_back(Sq);
        if (!taper || !last_unit) return c;
        int a0 = c[c.size() - 2];
        for (int level = 0; level < 3; ++level) {
            const int len = Sq - a0;
            const int half = (len / 2 + 255) / 256 * 256;
            if (half <= 0 || half >= len || (size_t)(len - half) * bytes_per_row < (size_t)tune.taper_bytes) break;
            a0 += half;                                   // the big part first
            c.push_back(a0);
        }
        std::sort(c.begin(), c.end());
        c.erase(std::unique(c.begin(), c.end()), c.end());
        return c;
    };
    int unit = 0;
    for (int b = 0; b < B; ++b) {
        int hk0 = 0;
        for (size_t g = 0; g < groups.size(); hk0 += groups[g], ++g, ++unit) {
            const int hkn = groups[g];
            const std::vector<int> cuts = cuts_of(unit == n_units - 1, ((size_t)hkn * grp + 2 * (size_t)hkn) * D * es);
            for (size_t r = 0; r + 1 < cuts.size(); ++r) plan.push_back({b, hk0, hkn, cuts[r], cuts[r + 1], unit});
        }
    }
    return plan;
}

int fwd_host_impl(const sfa_fwd_args& a, int want_pieces, cudaStream_t cur) {
    if (!a.q || !a.k || !a.v || !a.o) return fail(SFA_ERR_INVALID, "fwd_host: null tensor pointer");
    if (a.batch < 0 || a.seqlen_q < 0 || a.seqlen_k < 0 || a.heads_q <= 0 || a.heads_kv <= 0 || a.heads_q % a.heads_kv)
        return fail(SFA_ERR_INVALID, "fwd_host: bad dimension");
    if (a.head_dim != 64 && a.head_dim != 128) return fail(SFA_ERR_UNSUPPORTED, "fwd_host: head_dim must be 64 or 128");
    if (a.dtype != SFA_FP16 && a.dtype != SFA_BF16) return fail(SFA_ERR_INVALID, "fwd_host: dtype must be fp16 or bf16");
    const int B = a.batch, Sq = a.seqlen_q, Sk = a.seqlen_k, Hq = a.heads_q, Hkv = a.heads_kv, D = a.head_dim;
    // heads of a token must be dense (head stride == head_dim): a token's heads are one DMA row
    if (a.q_head_stride != D || a.k_head_stride != D || a.v_head_stride != D || a.o_head_stride != D)
        return fail(SFA_ERR_INVALID, "fwd_host: the heads of a token must be contiguous (head stride == head_dim)");
    if (B == 0 || Sq == 0) return SFA_OK;
    int rc;
    if ((rc = check_pinned(a.q, "fwd_host: q")) || (rc = check_pinned(a.k, "fwd_host: k")) ||
        (rc = check_pinned(a.v, "fwd_host: v")) || (rc = check_pinned(a.o, "fwd_host: o")))
        return rc;
    if (a.lse && (rc = check_pinned(a.lse, "fwd_host: lse"))) return rc;
    if (!current_device()) return SFA_ERR_NO_DEVICE;
    Pipeline* pp = pipeline_for_current_device();
    if (!pp) return fail(SFA_ERR_CUDA, "fwd_host: no current device");
    std::lock_guard<std::mutex> lk(pp->mu);
    Pipeline& P = *pp;
    if ((rc = pipeline_init(P))) return rc;

    const size_t es = 2;
    const int grp = Hq / Hkv;
    const std::vector<HostPiece> plan = plan_fwd_host(B, Sq, Sk, Hq, Hkv, D, a.causal != 0, want_pieces);
    int rows_max = 0, hg_max = 0, n_units = 0;
    for (const HostPiece& pc : plan) {
        rows_max = std::max(rows_max, pc.r1 - pc.r0);
        hg_max = std::max(hg_max, pc.hkn);
        n_units = std::max(n_units, pc.unit + 1);
    }
    const int shift = Sk - Sq;
    const size_t q_row_max = (size_t)hg_max * grp * D * es, kv_row_max = (size_t)hg_max * D * es;
    const size_t kv_half = ((size_t)Sk * kv_row_max + 255) / 256 * 256;
    for (int i = 0; i < kRing; ++i)
        if ((rc = ensure(P, kQ0 + i, rows_max * q_row_max)) || (rc = ensure(P, kO0 + i, rows_max * q_row_max)) ||
            (rc = ensure(P, kL0 + i, (size_t)hg_max * grp * rows_max * 4)))
            return rc;
    for (int i = 0; i < kKvRing && i < n_units; ++i)
        if ((rc = ensure(P, kKV0 + i, 2 * kv_half + 256))) return rc;
    const bool tracing = host_tuning().trace != 0;
    if (tracing) {
        for (cudaEvent_t e : P.trace) cudaEventDestroy(e);
        P.trace.clear();
    }
    auto mark = [&](cudaStream_t st) {   // host_trace only: a timing event on `st`
        if (!tracing) return;
        cudaEvent_t e = nullptr;
        if (cudaEventCreate(&e) == cudaSuccess) { cudaEventRecord(e, st); P.trace.push_back(e); }
    };

    // the previous call (maybe on another stream) may still be using the staging buffers
    if (P.used) SFA_CUDA_CHECK(cudaStreamWaitEvent(cur, P.ev_last, 0));
    SFA_CUDA_CHECK(cudaEventRecord(P.ev_start, cur));
    SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_in, P.ev_start, 0));
    SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_out, P.ev_start, 0));
    // The pieces' kernels alternate between the caller's stream and a second one: they are independent, and a
    // persistent kernel ends with SMs going idle one by one (its last, heaviest items) - the next piece's CTAs move in
    // as they free up instead of waiting for the whole grid to drain.
    int n_ks = host_tuning().kernel_streams;
    n_ks = n_ks < 1 ? 1 : (n_ks > 4 ? 4 : n_ks);
    for (int i = 1; i < n_ks; ++i) SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_comp[i - 1], P.ev_start, 0));
    const uint8_t* hq = reinterpret_cast<const uint8_t*>(a.q);
    const uint8_t* hk = reinterpret_cast<const uint8_t*>(a.k);
    const uint8_t* hv = reinterpret_cast<const uint8_t*>(a.v);
    uint8_t* ho = reinterpret_cast<uint8_t*>(a.o);
    bool comp_used[kRing] = {}, out_used[kRing] = {}, kv_used[kKvRing] = {};
    int keys_up = 0;   // keys of the current unit already on the device
    for (size_t piece = 0; piece < plan.size(); ++piece) {
        const HostPiece& pc = plan[piece];
        const int b = pc.b, hk0 = pc.hk0, hkn = pc.hkn;                    // kv-heads [hk0, hk0 + hkn)
        const int hq0 = hk0 * grp, hqn = hkn * grp;
        const size_t q_row = (size_t)hqn * D * es, kv_row = (size_t)hkn * D * es;
        const int kvb = pc.unit % kKvRing;
        uint8_t* dK = reinterpret_cast<uint8_t*>(P.buf[kKV0 + kvb]);
        uint8_t* dV = dK + kv_half;
        if (piece == 0 || plan[piece - 1].unit != pc.unit) {
            // this K/V buffer was last read by the kernels of the unit kKvRing units back
            if (kv_used[kvb]) {
                for (int i = 0; i < n_ks; ++i) SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_in, P.ev_kv_done[kvb][i], 0));
            }
            keys_up = 0;
        }
        const int r0 = pc.r0, r1 = pc.r1, rows = r1 - r0;
        int keys = a.causal ? r1 + shift : Sk;       // keys [0, keys) are visible to rows < r1
        keys = keys < 0 ? 0 : (keys > Sk ? Sk : keys);
        const int pb = (int)(piece % kRing);
        uint8_t* dQ = reinterpret_cast<uint8_t*>(P.buf[kQ0 + pb]);
        uint8_t* dO = reinterpret_cast<uint8_t*>(P.buf[kO0 + pb]);
        float* dL = reinterpret_cast<float*>(P.buf[kL0 + pb]);
        // ---- upload: the piece's Q rows and the K/V rows it is the first to need
        if (comp_used[pb]) SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_in, P.ev_comp[pb], 0));
        mark(P.s_in);
        if ((rc = copy_rows(dQ, q_row, hq + es * ((size_t)b * a.q_batch_stride + (size_t)r0 * a.q_row_stride + (size_t)hq0 * D),
                            es * a.q_row_stride, q_row, rows, cudaMemcpyHostToDevice, P.s_in)))
            return rc;
        if (keys > keys_up) {
            const size_t n = (size_t)(keys - keys_up);
            if ((rc = copy_rows(dK + keys_up * kv_row, kv_row,
                                hk + es * ((size_t)b * a.k_batch_stride + (size_t)keys_up * a.k_row_stride + (size_t)hk0 * D),
                                es * a.k_row_stride, kv_row, n, cudaMemcpyHostToDevice, P.s_in)) ||
                (rc = copy_rows(dV + keys_up * kv_row, kv_row,
                                hv + es * ((size_t)b * a.v_batch_stride + (size_t)keys_up * a.v_row_stride + (size_t)hk0 * D),
                                es * a.v_row_stride, kv_row, n, cudaMemcpyHostToDevice, P.s_in)))
                return rc;
            keys_up = keys;
        }
        mark(P.s_in);
        SFA_CUDA_CHECK(cudaEventRecord(P.ev_in[pb], P.s_in));
        // ---- kernel
        const int ki = (int)(piece % n_ks);
        cudaStream_t ks = ki == 0 ? cur : P.s_comp[ki - 1];
        SFA_CUDA_CHECK(cudaStreamWaitEvent(ks, P.ev_in[pb], 0));
        if (out_used[pb]) SFA_CUDA_CHECK(cudaStreamWaitEvent(ks, P.ev_out[pb], 0));
        sfa_fwd_args d = {};
        d.q = dQ; d.k = dK; d.v = dV; d.o = dO; d.lse = a.lse ? dL : nullptr;
        d.batch = 1; d.seqlen_q = rows; d.seqlen_k = keys; d.heads_q = hqn; d.heads_kv = hkn; d.head_dim = D;
        d.q_batch_stride = (int64_t)rows * hqn * D; d.q_row_stride = (int64_t)hqn * D; d.q_head_stride = D;
        d.k_batch_stride = (int64_t)Sk * hkn * D; d.k_row_stride = (int64_t)hkn * D; d.k_head_stride = D;
        d.v_batch_stride = d.k_batch_stride; d.v_row_stride = d.k_row_stride; d.v_head_stride = D;
        d.o_batch_stride = d.q_batch_stride; d.o_row_stride = d.q_row_stride; d.o_head_stride = D;
        d.softmax_scale = a.softmax_scale; d.causal = a.causal; d.dtype = a.dtype;
        mark(ks);
        if ((rc = launch_fwd(d, ks))) return rc;
        mark(ks);
        SFA_CUDA_CHECK(cudaEventRecord(P.ev_comp[pb], ks));
        comp_used[pb] = true;
        // ---- download
        SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_out, P.ev_comp[pb], 0));
        if ((rc = copy_rows(ho + es * ((size_t)b * a.o_batch_stride + (size_t)r0 * a.o_row_stride + (size_t)hq0 * D),
                            es * a.o_row_stride, dO, q_row, q_row, rows, cudaMemcpyDeviceToHost, P.s_out)))
            return rc;
        if (a.lse)   // host (batch, heads, Sq) fp32: one row of `rows` floats per head
            if ((rc = copy_rows(a.lse + ((size_t)b * Hq + hq0) * Sq + r0, (size_t)Sq * 4, dL, (size_t)rows * 4,
                                (size_t)rows * 4, hqn, cudaMemcpyDeviceToHost, P.s_out)))
                return rc;
        mark(P.s_out);
        SFA_CUDA_CHECK(cudaEventRecord(P.ev_out[pb], P.s_out));
        out_used[pb] = true;
        if (piece + 1 == plan.size() || plan[piece + 1].unit != pc.unit) {
            for (int i = 0; i < n_ks; ++i) SFA_CUDA_CHECK(cudaEventRecord(P.ev_kv_done[kvb][i], i == 0 ? cur : P.s_comp[i - 1]));
            kv_used[kvb] = true;
        }
    }
    for (int i = 0; i < kRing; ++i)
        if (out_used[i]) SFA_CUDA_CHECK(cudaStreamWaitEvent(cur, P.ev_out[i], 0));
    SFA_CUDA_CHECK(cudaEventRecord(P.ev_last, cur));
    P.used = true;
    g_last_plan_pieces = (int)plan.size();
    return SFA_OK;
}

// Split-KV decode on pinned host tensors: the batch is cut into pieces whose cache upload overlaps the previous
// piece's kernel and download.
int decode_host_impl(const sfa_decode_args& a, int want_pieces, cudaStream_t cur) {
    if (!a.q || !a.k_cache || !a.v_cache || !a.o) return fail(SFA_ERR_INVALID, "decode_host: null tensor pointer");
    if (a.batch < 0 || a.seqlen_k < 0 || a.heads_q <= 0 || a.heads_kv <= 0 || a.heads_q % a.heads_kv)
        return fail(SFA_ERR_INVALID, "decode_host: bad dimension");
    if (a.dtype != SFA_FP16 && a.dtype != SFA_BF16) return fail(SFA_ERR_INVALID, "decode_host: dtype must be fp16 or bf16");
    const int B = a.batch, M = a.seqlen_k, Hq = a.heads_q, Hkv = a.heads_kv, D = a.head_dim;
    if (a.q_head_stride != D || a.k_head_stride != D || a.v_head_stride != D || a.o_head_stride != D)
        return fail(SFA_ERR_INVALID, "decode_host: the heads of a token must be contiguous (head stride == head_dim)");
    if (B == 0) return SFA_OK;
    int rc;
    if ((rc = check_pinned(a.q, "decode_host: q")) || (rc = check_pinned(a.k_cache, "decode_host: k_cache")) ||
        (rc = check_pinned(a.v_cache, "decode_host: v_cache")) || (rc = check_pinned(a.o, "decode_host: o")))
        return rc;
    if (a.seqlens && (rc = check_pinned(a.seqlens, "decode_host: seqlens"))) return rc;
    if (a.lse && (rc = check_pinned(a.lse, "decode_host: lse"))) return rc;
    if (!current_device()) return SFA_ERR_NO_DEVICE;
    Pipeline* pp = pipeline_for_current_device();
    if (!pp) return fail(SFA_ERR_CUDA, "decode_host: no current device");
    std::lock_guard<std::mutex> lk(pp->mu);
    Pipeline& P = *pp;
    if ((rc = pipeline_init(P))) return rc;

    if (want_pieces <= 0) want_pieces = 8;
    const int n = want_pieces < B ? want_pieces : B;
    const int step = (B + n - 1) / n;
    const size_t es = 2, q_row = (size_t)Hq * D * es, kv_row = (size_t)Hkv * D * es;
    const size_t q_bytes = ((size_t)step * q_row + 255) / 256 * 256, cache_bytes = ((size_t)step * M * kv_row + 255) / 256 * 256;
    for (int i = 0; i < 2; ++i) {
        if ((rc = ensure(P, kQ0 + i, q_bytes + (size_t)step * 4 + 256)) ||                 // q, seqlens
            (rc = ensure(P, kO0 + i, q_bytes + (size_t)step * Hq * 4 + 256)) ||               // o, lse
            (rc = ensure(P, kKV0 + i, 2 * cache_bytes)))
            return rc;
    }
    // the previous call (maybe on another stream) may still be using the staging buffers
    if (P.used) SFA_CUDA_CHECK(cudaStreamWaitEvent(cur, P.ev_last, 0));
    SFA_CUDA_CHECK(cudaEventRecord(P.ev_start, cur));
    SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_in, P.ev_start, 0));
    SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_out, P.ev_start, 0));
    bool comp_used[2] = {false, false}, out_used[2] = {false, false};
    int piece = 0;
    for (int lo = 0; lo < B; lo += step, ++piece) {
        const int w = (B - lo < step) ? B - lo : step, pb = piece & 1;
        uint8_t* dQ = reinterpret_cast<uint8_t*>(P.buf[kQ0 + pb]);
        int32_t* dS = reinterpret_cast<int32_t*>(dQ + q_bytes);
        uint8_t* dO = reinterpret_cast<uint8_t*>(P.buf[kO0 + pb]);
        float* dL = reinterpret_cast<float*>(dO + q_bytes);
        uint8_t* dK = reinterpret_cast<uint8_t*>(P.buf[kKV0 + pb]);
        uint8_t* dV = dK + cache_bytes;
        if (comp_used[pb]) SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_in, P.ev_comp[pb], 0));
        if ((rc = copy_rows(dQ, q_row, reinterpret_cast<const uint8_t*>(a.q) + es * (size_t)lo * a.q_batch_stride,
                            es * a.q_batch_stride, q_row, w, cudaMemcpyHostToDevice, P.s_in)))
            return rc;
        for (int b = 0; b < w; ++b) {   // a sequence's cache: M rows of all kv heads
            if ((rc = copy_rows(dK + (size_t)b * M * kv_row, kv_row,
                                reinterpret_cast<const uint8_t*>(a.k_cache) + es * (size_t)(lo + b) * a.k_batch_stride,
                                es * a.k_row_stride, kv_row, M, cudaMemcpyHostToDevice, P.s_in)) ||
                (rc = copy_rows(dV + (size_t)b * M * kv_row, kv_row,
                                reinterpret_cast<const uint8_t*>(a.v_cache) + es * (size_t)(lo + b) * a.v_batch_stride,
                                es * a.v_row_stride, kv_row, M, cudaMemcpyHostToDevice, P.s_in)))
                return rc;
        }
        if (a.seqlens) SFA_CUDA_CHECK(cudaMemcpyAsync(dS, a.seqlens + lo, (size_t)w * 4, cudaMemcpyHostToDevice, P.s_in));
        SFA_CUDA_CHECK(cudaEventRecord(P.ev_in[pb], P.s_in));
        SFA_CUDA_CHECK(cudaStreamWaitEvent(cur, P.ev_in[pb], 0));
        if (out_used[pb]) SFA_CUDA_CHECK(cudaStreamWaitEvent(cur, P.ev_out[pb], 0));
        sfa_decode_args d = {};
        d.q = dQ; d.k_cache = dK; d.v_cache = dV; d.seqlens = a.seqlens ? dS : nullptr; d.o = dO; d.lse = a.lse ? dL : nullptr;
        d.batch = w; d.seqlen_k = M; d.heads_q = Hq; d.heads_kv = Hkv; d.head_dim = D;
        d.q_batch_stride = (int64_t)Hq * D; d.q_head_stride = D;
        d.k_batch_stride = (int64_t)M * Hkv * D; d.k_row_stride = (int64_t)Hkv * D; d.k_head_stride = D;
        d.v_batch_stride = d.k_batch_stride; d.v_row_stride = d.k_row_stride; d.v_head_stride = D;
        d.o_batch_stride = (int64_t)Hq * D; d.o_head_stride = D;
        d.softmax_scale = a.softmax_scale; d.num_splits = a.num_splits; d.dtype = a.dtype;
        d.workspace = nullptr; d.workspace_bytes = 0;   // split partials: the stream's own library scratch
        if ((rc = sfa_decode(&d, cur))) return rc;
        SFA_CUDA_CHECK(cudaEventRecord(P.ev_comp[pb], cur));
        comp_used[pb] = true;
        SFA_CUDA_CHECK(cudaStreamWaitEvent(P.s_out, P.ev_comp[pb], 0));
        if ((rc = copy_rows(reinterpret_cast<uint8_t*>(a.o) + es * (size_t)lo * a.o_batch_stride, es * a.o_batch_stride, dO, q_row,
                            q_row, w, cudaMemcpyDeviceToHost, P.s_out)))
            return rc;
        if (a.lse) SFA_CUDA_CHECK(cudaMemcpyAsync(a.lse + (size_t)lo * Hq, dL, (size_t)w * Hq * 4, cudaMemcpyDeviceToHost, P.s_out));
        SFA_CUDA_CHECK(cudaEventRecord(P.ev_out[pb], P.s_out));
        out_used[pb] = true;
    }
    for (int i = 0; i < 2; ++i)
        if (out_used[i]) SFA_CUDA_CHECK(cudaStreamWaitEvent(cur, P.ev_out[i], 0));
    SFA_CUDA_CHECK(cudaEventRecord(P.ev_last, cur));
    P.used = true;
    return SFA_OK;
}

HostTuning& host_tuning() {
    static HostTuning t;
    return t;
}

int host_last_plan_pieces() { return g_last_plan_pieces.load(); }

// host_trace: milliseconds from the first upload's start to (upload start, upload end, kernel start, kernel end,
// download end) of every piece of the last sfa_fwd_host call on the current device; returns the number of pieces.
int fwd_host_trace(float* out, int capacity) {
    Pipeline* pp = pipeline_for_current_device();
    if (!pp) return 0;
    std::lock_guard<std::mutex> lk(pp->mu);
    if (cudaDeviceSynchronize() != cudaSuccess) return fail(SFA_ERR_CUDA, "fwd_host_trace: device synchronize failed");
    const int n = (int)pp->trace.size() / 5;
    for (int i = 0; i < n * 5 && out && i < capacity * 5; ++i) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, pp->trace[0], pp->trace[i]);
        out[i] = ms;
    }
    return n;
}

void release_host_pipelines() {
    std::lock_guard<std::mutex> lk(g_pipes_mu);
    int cur = -1;
    cudaGetDevice(&cur);
    for (size_t i = 0; i < g_pipes.size(); ++i) {
        Pipeline* p = g_pipes[i];
        if (!p) continue;
        std::lock_guard<std::mutex> lk2(p->mu);
        cudaSetDevice((int)i);
        cudaDeviceSynchronize();
        for (int b = 0; b < 16; ++b) {
            if (p->buf[b]) cudaFree(p->buf[b]);
            p->buf[b] = nullptr;
            p->cap[b] = 0;
        }
        for (void* r : p->retired) cudaFree(r);
        p->retired.clear();
    }
    if (cur >= 0) cudaSetDevice(cur);
}

}  // namespace sfa
