// Cross-GPU split-KV decode: the KV length of one long sequence is sharded over the GPUs of a box; every rank
// runs the split-KV kernel over its shard and the per-rank results (O_r normalised, LSE_r) are exchanged and
// merged (the merge is the reference's flash_combine_kernel,
// src/flash_attn.cu:877-935, applied across devices instead of across splits).
//
// shard_exchange_kernel does the three steps in ONE kernel over peer memory: (1) merge this rank's splits,
// (2) store the result straight into every peer's workspace through NVLink (P2P stores into CUDA-IPC mapped
// buffers) and publish a per-row flag with release.sys semantics, (3) wait for the peers' flags (acquire.sys)
// and merge the world's partials.  A row's block depends only on the same row's block on the other ranks, so
// there is no intra-grid dependency.  Workspaces are double-buffered by step parity: a rank can be at most one
// step ahead of a peer (it needs the peer's flags of step s+1 to finish step s+1), so slot (s & 1) is quiescent
// when step s+2 writes it.  Flags hold the step number and are never reset.
#include "internal.h"
#include "merge.cuh"

#include <math.h>

namespace sfa {
namespace {

constexpr float kLn2 = 0.6931471805599453f;
constexpr int kMaxWorld = 8;

struct ShardParams {
    const float* o_part;      // (rows, splits, D) this rank's split partials
    const float2* ml_part;    // (rows, splits)   (m in log2 domain, l)
    int num_splits;
    void* o; float* lse;      // final output (B, Hq, D) in T and (B, Hq) fp32; written by every rank
    int64_t o_bs, o_hs;
    int heads_q, rows;
    int rank, world;
    uint32_t step;
    uint8_t* ws[kMaxWorld];   // ws[r] = rank r's workspace as mapped in this process (ws[rank] is local)
    float* payload;           // publish-only mode: (rows, D + 1) local buffer, no flags
    int* error_flag;
};

__host__ __device__ inline size_t slot_floats(int rows, int D) { return (size_t)rows * (D + 1); }
__host__ __device__ inline size_t slot_offset(int parity, int src, int world, int rows, int D) {
    return ((size_t)parity * world + src) * slot_floats(rows, D) * sizeof(float);
}
__host__ __device__ inline size_t flags_offset(int world, int rows, int D) {
    return 2 * (size_t)world * slot_floats(rows, D) * sizeof(float);
}
__host__ __device__ inline size_t flag_index(int parity, int src, int world, int rows, int row) {
    return ((size_t)parity * world + src) * rows + row;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ int g_shard_error;

// MODE 0: fused exchange + merge over peer memory.  MODE 1: publish this rank's (O, LSE) to p.payload only.
// MODE 2: merge `world` gathered payloads laid out (world, rows, D + 1) at p.payload (after an all-gather).
template <typename T, int D, int MODE>
__device__ __forceinline__ void shard_exchange_body(const ShardParams& p) {
    extern __shared__ float w_sm[];   // num_splits weights
    __shared__ MergeScratch<D> sc;
    __shared__ float lse_sm[kMaxWorld];
    __shared__ int ok_sm;
    const int hq = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int row = b * p.heads_q + hq;
    float o_n = 0.f, lse2 = -INFINITY;
    if constexpr (MODE != 2) {
        // ---- (1) merge this rank's splits (whole block); the rest of the kernel is the first D threads' work
        float acc, L, M;
        merge_splits<D>(p.o_part, p.ml_part, (int64_t)row * p.num_splits, p.num_splits, w_sm, sc, acc, L, M);
        o_n = L > 0.f ? acc / L : 0.f;
        lse2 = L > 0.f ? M + log2f(L) : -INFINITY;
    }
    const bool worker = tid < D;
    if constexpr (MODE == 1) {
        if (worker) p.payload[(size_t)row * D + tid] = o_n;
        if (tid == 0) p.payload[(size_t)p.rows * D + row] = lse2;
        return;
    }
    const int parity = p.step & 1;
    const float* src_base[kMaxWorld];
    if constexpr (MODE == 0) {
        // ---- (2) push to every peer (and to myself), then publish the row's flag
#pragma unroll 1
        for (int dst = 0; dst < p.world; ++dst) {
            float* slot = reinterpret_cast<float*>(p.ws[dst] + slot_offset(parity, p.rank, p.world, p.rows, D));
            if (worker) slot[(size_t)row * D + tid] = o_n;
            if (tid == 0) slot[(size_t)p.rows * D + row] = lse2;
        }
        __threadfence_system();
        __syncthreads();
        if (tid < p.world) {
            uint32_t* flags = reinterpret_cast<uint32_t*>(p.ws[tid] + flags_offset(p.world, p.rows, D));
            st_release_sys(flags + flag_index(parity, p.rank, p.world, p.rows, row), p.step);
        }
        // ---- (3) wait for every rank's row (bounded: a missing peer must not hang the GPU)
        if (tid == 0) ok_sm = 1;
        __syncthreads();
        if (tid < p.world) {
            const uint32_t* flags = reinterpret_cast<const uint32_t*>(p.ws[p.rank] + flags_offset(p.world, p.rows, D));
            const uint32_t* f = flags + flag_index(parity, tid, p.world, p.rows, row);
            const uint64_t t0 = timer_ns();
            uint32_t spins = 0;
            while (ld_acquire_sys(f) != p.step) {
                if ((++spins & 0x3FF) == 0 && timer_ns() - t0 > 2000000000ull) {
                    atomicCAS(p.error_flag, 0, 1 + tid);
                    ok_sm = 0;
                    break;
                }
            }
        }
        __syncthreads();
        if (!ok_sm) {
            // A peer's row never arrived: the caller must not read an unwritten output as if it were a result.  The
            // row becomes NaN (O and LSE) and the device flag names the missing rank (sfa_shard_watchdog).
            T* orow = reinterpret_cast<T*>(p.o) + (int64_t)b * p.o_bs + (int64_t)hq * p.o_hs;
            if (worker) orow[tid] = (T)__int_as_float(0x7fc00000);
            if (tid == 0 && p.lse) p.lse[row] = __int_as_float(0x7fc00000);
            return;
        }
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r)
            src_base[r] = reinterpret_cast<const float*>(p.ws[p.rank] + slot_offset(parity, r < p.world ? r : 0, p.world, p.rows, D));
    } else {
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r) src_base[r] = p.payload + (size_t)(r < p.world ? r : 0) * slot_floats(p.rows, D);
    }
    // ---- merge the world's partials for this row
    // L2 loads: the rows were written by peers over NVLink, a stale L1 line of an earlier step must not be hit
    if (tid < p.world) lse_sm[tid] = __ldcg(src_base[tid] + (size_t)p.rows * D + row);
    __syncthreads();
    float M = -INFINITY;
    for (int r = 0; r < p.world; ++r) M = fmaxf(M, lse_sm[r]);
    float W = 0.f, acc = 0.f;
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
        if (r < p.world) {
            const float w = lse_sm[r] == -INFINITY ? 0.f : exp2f(lse_sm[r] - M);
            W += w;
            if (worker) acc += w * __ldcg(src_base[r] + (size_t)row * D + tid);
        }
    }
    T* orow = reinterpret_cast<T*>(p.o) + (int64_t)b * p.o_bs + (int64_t)hq * p.o_hs;
    if (worker) orow[tid] = (T)(W > 0.f ? acc / W : 0.f);
    if (tid == 0 && p.lse) p.lse[row] = W > 0.f ? (M + log2f(W)) * kLn2 : -INFINITY;
}

template <typename T, int D, int MODE>
__global__ void __launch_bounds__(D * kMergeGroups) shard_exchange_kernel(const __grid_constant__ ShardParams p) {
    shard_exchange_body<T, D, MODE>(p);
}

// Single-GPU emulation for tests: the blocks of ALL ranks in one cooperative launch (blockIdx.z = rank), so that the
// blocks that spin on each other's flags are guaranteed to be co-resident.  (Separate launches that wait on one
// another on one GPU have no such guarantee and have raised Xid 109 on this driver.)
template <typename T, int D>
__global__ void __launch_bounds__(D * kMergeGroups) shard_exchange_emulated_kernel(const ShardParams* per_rank) {
    shard_exchange_body<T, D, 0>(per_rank[blockIdx.z]);
}

template <typename T, int D>
int launch_emulated(const ShardParams* dev_params, int world, int heads_q, int batch, int num_splits_max, cudaStream_t stream) {
    dim3 grid(heads_q, batch, world);
    dim3 block(D * kMergeGroups);
    size_t smem = (size_t)num_splits_max * sizeof(float);
    void* kargs[] = {(void*)&dev_params};
    SFA_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)shard_exchange_emulated_kernel<T, D>, grid, block, kargs, smem, stream));
    count_launch();
    return SFA_OK;
}

template <typename T, int D>
int launch_mode(const ShardParams& p, int mode, int batch, cudaStream_t stream) {
    dim3 grid(p.heads_q, batch);
    const size_t smem = (size_t)(p.num_splits > 0 ? p.num_splits : 1) * sizeof(float);
    if (mode == 0) {
        // A row's block spins until the same row's block of every peer has published: all of a launch's blocks must
        // be resident at once on every rank (CUDA does not promise in-order dispatch beyond one wave).
        static int per_sm = [] {
            int n = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shard_exchange_kernel<T, D, 0>, D * kMergeGroups, 1024) != cudaSuccess) n = 0;
            return n;
        }();
        const DeviceInfo* dev = current_device();
        if (!dev) return SFA_ERR_NO_DEVICE;
        if ((long long)p.heads_q * batch > (long long)per_sm * dev->sm_count)
            return fail(SFA_ERR_UNSUPPORTED, "sharded decode: batch * heads_q = " + std::to_string(p.heads_q * batch) +
                        " rows exceed one resident wave (" + std::to_string(per_sm * dev->sm_count) +
                        "); use the publish + all-gather + merge form (mode 1, exchange=\"nccl\")");
    }
    if (mode == 0) shard_exchange_kernel<T, D, 0><<<grid, D * kMergeGroups, smem, stream>>>(p);
    else if (mode == 1) shard_exchange_kernel<T, D, 1><<<grid, D * kMergeGroups, smem, stream>>>(p);
    else shard_exchange_kernel<T, D, 2><<<grid, D * kMergeGroups, smem, stream>>>(p);
    SFA_CUDA_CHECK(cudaGetLastError());
    count_launch();
    return SFA_OK;
}

}  // namespace

size_t shard_workspace_bytes(int world, int rows, int head_dim) {
    return flags_offset(world, rows, head_dim) + 2 * (size_t)world * rows * sizeof(uint32_t);
}

namespace {
int fill_params(const ShardLaunch& a, ShardParams& p) {
    if (a.world < 1 || a.world > kMaxWorld || a.rank < 0 || a.rank >= a.world)
        return fail(SFA_ERR_INVALID, "sharded decode: world must be 1..8 and 0 <= rank < world");
    if (a.head_dim != 64 && a.head_dim != 128) return fail(SFA_ERR_UNSUPPORTED, "head_dim must be 64 or 128");
    p = ShardParams{};
    p.o_part = a.o_part; p.ml_part = a.ml_part; p.num_splits = a.num_splits;
    p.o = a.o; p.lse = a.lse; p.o_bs = a.o_bs; p.o_hs = a.o_hs;
    p.heads_q = a.heads_q; p.rows = a.batch * a.heads_q;
    p.rank = a.rank; p.world = a.world; p.step = a.step;
    for (int r = 0; r < a.world; ++r) p.ws[r] = reinterpret_cast<uint8_t*>(a.peer_ws ? a.peer_ws[r] : nullptr);
    p.payload = a.payload;
    void* flag_addr = nullptr;
    SFA_CUDA_CHECK(cudaGetSymbolAddress(&flag_addr, g_shard_error));
    p.error_flag = reinterpret_cast<int*>(flag_addr);
    return SFA_OK;
}
}  // namespace

int launch_shard_exchange(const ShardLaunch& a, cudaStream_t stream) {
    ShardParams p;
    int rc = fill_params(a, p);
    if (rc) return rc;
    if (a.dtype == SFA_FP16)
        return a.head_dim == 128 ? launch_mode<__half, 128>(p, a.mode, a.batch, stream) : launch_mode<__half, 64>(p, a.mode, a.batch, stream);
    return a.head_dim == 128 ? launch_mode<__nv_bfloat16, 128>(p, a.mode, a.batch, stream)
                             : launch_mode<__nv_bfloat16, 64>(p, a.mode, a.batch, stream);
}

int launch_shard_exchange_emulated(const ShardLaunch* ranks, int world, cudaStream_t stream) {
    if (world < 1 || world > kMaxWorld) return fail(SFA_ERR_INVALID, "sharded decode: world must be 1..8");
    ShardParams host[kMaxWorld];
    int splits_max = 1;
    for (int r = 0; r < world; ++r) {
        int rc = fill_params(ranks[r], host[r]);
        if (rc) return rc;
        if (ranks[r].num_splits > splits_max) splits_max = ranks[r].num_splits;
    }
    ShardParams* dev = nullptr;
    SFA_CUDA_CHECK(cudaMalloc(&dev, sizeof(ShardParams) * world));
    cudaError_t e = cudaMemcpyAsync(dev, host, sizeof(ShardParams) * world, cudaMemcpyHostToDevice, stream);
    int rc = SFA_OK;
    if (e != cudaSuccess) rc = fail(SFA_ERR_CUDA, cudaGetErrorString(e));
    const ShardLaunch& a = ranks[0];
    if (!rc) {
        if (a.dtype == SFA_FP16)
            rc = a.head_dim == 128 ? launch_emulated<__half, 128>(dev, world, a.heads_q, a.batch, splits_max, stream)
                                   : launch_emulated<__half, 64>(dev, world, a.heads_q, a.batch, splits_max, stream);
        else
            rc = a.head_dim == 128 ? launch_emulated<__nv_bfloat16, 128>(dev, world, a.heads_q, a.batch, splits_max, stream)
                                   : launch_emulated<__nv_bfloat16, 64>(dev, world, a.heads_q, a.batch, splits_max, stream);
    }
    cudaStreamSynchronize(stream);     // test-only path: the parameter block is freed right away
    cudaFree(dev);
    return rc;
}

int shard_error_flag_read(int* value, int clear) {
    SFA_CUDA_CHECK(cudaMemcpyFromSymbol(value, g_shard_error, sizeof(int)));
    if (clear && *value) {
        const int zero = 0;
        SFA_CUDA_CHECK(cudaMemcpyToSymbol(g_shard_error, &zero, sizeof(int)));
    }
    return SFA_OK;
}

}  // namespace sfa
