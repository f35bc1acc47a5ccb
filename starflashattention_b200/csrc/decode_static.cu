// sfa_decode_static: split-KV decode that allocates nothing.  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
// harnesses that forbid allocation).
//
// The split-KV kernels keep their per-split partials (un-normalised O in fp32 plus (m, l) per row and split) in
// scratch memory that sfa_decode takes from the caller or from the library's grow-only per-stream buffer
// (cudaMalloc).  Here the scratch is ONE file-scope device array of kStaticScratchBytes that every GPU which loads
// the library holds until exit.  Consequences, all documented in include/star_flash_attn.h:
//  * a launch whose partials do not fit is rejected with SFA_ERR_INVALID and the number of bytes it needs;
//  * every call on a device shares the array: two calls that overlap on the GPU (different streams, or graphs
//    replayed concurrently) race on it;
//  * the array keeps the last call's contents.  Nothing in it has to start at zero: every CTA of the split kernel
//    writes its own (row, split) partial - empty chunks write (m, l) = (-inf, 0) - before the combine kernel reads it.
#include "internal.h"

#include <mutex>
#include <string>

namespace sfa {
namespace {

constexpr size_t kStaticScratchBytes = SFA_DECODE_STATIC_SCRATCH_BYTES;   // 16 MiB (include/star_flash_attn.h)

// In host code this name is a host-side stand-in for the symbol, NOT a device address: a kernel must be given the
// pointer cudaGetSymbolAddress returns for the current device.
static __device__ __align__(256) unsigned char g_decode_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices] = {};   // device address of the array, per device ordinal (looked up once per device)

// Device address of the scratch array on the CURRENT device.
int static_scratch(void** out) {
    int dev = -1;
    SFA_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(SFA_ERR_UNSUPPORTED, "sfa_decode_static: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* ptr = nullptr;
        SFA_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_decode_static_scratch));   // `dev` is current
        if (!ptr) return fail(SFA_ERR_CUDA, "sfa_decode_static: cudaGetSymbolAddress returned a null address");
        g_static_ptr[dev] = ptr;
    }
    *out = g_static_ptr[dev];
    return SFA_OK;
}

}  // namespace
}  // namespace sfa

extern "C" {

size_t sfa_decode_static_scratch_bytes(void) { return sfa::kStaticScratchBytes; }

int sfa_decode_static(const sfa_decode_args* args, void* stream) {
    if (!args) return sfa::fail(SFA_ERR_INVALID, "sfa_decode_static: null args");
    sfa_decode_args a = *args;
    a.workspace = nullptr;
    a.workspace_bytes = 0;
    if (a.batch <= 0 || a.heads_q <= 0 || a.heads_kv <= 0 || a.head_dim <= 0) return sfa_decode(&a, stream);   // nothing to size: its checks answer
    // split count and scratch size exactly as sfa_decode would choose them
    const int splits = sfa_decode_num_splits(&a);
    if (splits < 0) return splits;
    const size_t need = sfa_decode_workspace_bytes(a.batch, a.heads_q, a.head_dim, splits);
    if (need > sfa::kStaticScratchBytes)
        return sfa::fail(SFA_ERR_INVALID, "sfa_decode_static: batch * heads_q * num_splits partials need " + std::to_string(need) +
                                              " bytes of scratch, the static array holds " + std::to_string(sfa::kStaticScratchBytes) +
                                              " (lower num_splits or the batch, or call sfa_decode with a workspace)");
    void* scratch = nullptr;
    const int rc = sfa::static_scratch(&scratch);
    if (rc) return rc;
    a.num_splits = splits;
    a.workspace = scratch;                       // the caller-workspace path of sfa_decode: no allocation anywhere
    a.workspace_bytes = sfa::kStaticScratchBytes;
    return sfa_decode(&a, stream);
}

}  // extern "C"
