// Library-wide host state: last-error string, device properties, per-device workspace, launch counter.
#include "internal.h"

#include <atomic>
#include <mutex>
#include <vector>

namespace sfa {

static thread_local std::string g_last_error;
static std::atomic<uint64_t> g_launches{0};

void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

namespace {
// Scratch of one (device, stream): split-KV partials live here between the two launches of a decode call, so two
// streams must never share it.  Buffers that were outgrown are kept until sfa_release_workspaces(): a CUDA graph
// may have captured their address, and work queued earlier may still be using them.
struct StreamScratch {
    cudaStream_t stream = nullptr;
    void* ws = nullptr;
    size_t ws_bytes = 0;
};
constexpr int kSchedSlots = 4096;          // [0, 2048) rotate per launch, [2048, 4096) are given out once (captures)
struct PerDevice {
    DeviceInfo info;
    bool probed = false;
    std::vector<StreamScratch> scratch;
    std::vector<void*> retired;
    int* sched = nullptr;                  // kSchedSlots x 2 ints, zero between launches
    unsigned sched_next = 0, sched_captured = 0;
};
std::mutex g_mu;
std::vector<PerDevice> g_dev;
}  // namespace

const DeviceInfo* current_device() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) {
        cudaGetLastError();
        set_error("no CUDA device is available; this library has no CPU fallback");
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    if ((int)g_dev.size() <= dev) g_dev.resize(dev + 1);
    PerDevice& d = g_dev[dev];
    if (!d.probed) {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
            cudaGetLastError();
            set_error("cudaGetDeviceProperties failed");
            return nullptr;
        }
        d.info.device = dev;
        d.info.sm_count = prop.multiProcessorCount;
        d.info.cc_major = prop.major;
        d.info.cc_minor = prop.minor;
        d.probed = true;
    }
    if (d.info.cc_major != 10) {
        set_error("device compute capability is " + std::to_string(d.info.cc_major) + "." +
                  std::to_string(d.info.cc_minor) + "; the kernels are built for sm_100a only");
        return nullptr;
    }
    return &d.info;
}

void* get_workspace(size_t bytes, cudaStream_t stream) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lk(g_mu);
    if ((int)g_dev.size() <= dev) g_dev.resize(dev + 1);
    PerDevice& d = g_dev[dev];
    StreamScratch* sc = nullptr;
    for (StreamScratch& x : d.scratch)
        if (x.stream == stream) { sc = &x; break; }
    if (!sc) {
        d.scratch.push_back(StreamScratch{stream, nullptr, 0});
        sc = &d.scratch.back();
    }
    if (sc->ws_bytes >= bytes) return sc->ws;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cap);
    if (cap != cudaStreamCaptureStatusNone) {
        set_error("workspace must grow but the stream is being captured; warm the call up once on this stream "
                  "before capture or pass a caller-owned workspace");
        return nullptr;
    }
    // The outgrown buffer is retired, not freed: earlier work on this stream, or a captured graph, may still use it.
    if (sc->ws) d.retired.push_back(sc->ws);
    sc->ws = nullptr;
    sc->ws_bytes = 0;
    size_t want = bytes < (size_t(8) << 20) ? (size_t(8) << 20) : bytes + bytes / 4;
    if (cudaMalloc(&sc->ws, want) != cudaSuccess) {
        cudaGetLastError();
        set_error("cudaMalloc of the split-KV workspace failed");
        return nullptr;
    }
    sc->ws_bytes = want;
    return sc->ws;
}

int* fwd_sched_counters(cudaStream_t stream) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lk(g_mu);
    if ((int)g_dev.size() <= dev) g_dev.resize(dev + 1);
    PerDevice& d = g_dev[dev];
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cap);
    if (!d.sched) {
        if (cap != cudaStreamCaptureStatusNone) {
            set_error("the first prefill call on a device cannot be captured in a CUDA graph; warm it up once");
            return nullptr;
        }
        const size_t bytes = (size_t)kSchedSlots * 2 * sizeof(int);
        if (cudaMalloc(reinterpret_cast<void**>(&d.sched), bytes) != cudaSuccess ||
            cudaMemset(d.sched, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
            cudaGetLastError();
            d.sched = nullptr;
            set_error("allocation of the prefill schedule counters failed");
            return nullptr;
        }
    }
    unsigned slot;
    if (cap != cudaStreamCaptureStatusNone) {
        // a captured launch replays with the same counters for as long as its graph lives: a slot of its own
        if (d.sched_captured >= kSchedSlots / 2) {
            set_error("more than 2048 prefill launches captured in CUDA graphs on this device");
            return nullptr;
        }
        slot = kSchedSlots / 2 + d.sched_captured++;
    } else {
        slot = d.sched_next++ % (kSchedSlots / 2);
    }
    return d.sched + 2 * slot;
}

int fwd_sched_reset() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return SFA_ERR_NO_DEVICE;
    std::lock_guard<std::mutex> lk(g_mu);
    if ((int)g_dev.size() <= dev || !g_dev[dev].sched) return SFA_OK;
    SFA_CUDA_CHECK(cudaMemset(g_dev[dev].sched, 0, (size_t)kSchedSlots * 2 * sizeof(int)));
    return SFA_OK;
}

}  // namespace sfa

extern "C" {
const char* sfa_last_error(void) { return sfa::g_last_error.c_str(); }
int sfa_abi_version(void) { return SFA_ABI_VERSION; }
uint64_t sfa_launch_count(void) { return sfa::g_launches.load(); }
void sfa_release_workspaces(void) {
    sfa::release_host_pipelines();
    std::lock_guard<std::mutex> lk(sfa::g_mu);
    int cur = -1;
    cudaGetDevice(&cur);
    for (size_t i = 0; i < sfa::g_dev.size(); ++i) {
        sfa::PerDevice& d = sfa::g_dev[i];
        if (d.scratch.empty() && d.retired.empty()) continue;
        cudaSetDevice((int)i);
        cudaDeviceSynchronize();
        for (sfa::StreamScratch& x : d.scratch)
            if (x.ws) cudaFree(x.ws);
        for (void* r : d.retired) cudaFree(r);
        d.scratch.clear();
        d.retired.clear();
    }
    if (cur >= 0) cudaSetDevice(cur);
}
}
