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
struct PerDevice {
    DeviceInfo info;
    bool probed = false;
    void* ws = nullptr;
    size_t ws_bytes = 0;
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
    if (d.ws_bytes >= bytes) return d.ws;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cap);
    if (cap != cudaStreamCaptureStatusNone) {
        set_error("workspace must grow but the stream is being captured; warm the call up once before capture "
                  "or pass a caller-owned workspace");
        return nullptr;
    }
    // Work queued earlier may still read the old buffer.
    cudaDeviceSynchronize();
    if (d.ws) cudaFree(d.ws);
    d.ws = nullptr;
    d.ws_bytes = 0;
    size_t want = bytes < (size_t(8) << 20) ? (size_t(8) << 20) : bytes + bytes / 4;
    if (cudaMalloc(&d.ws, want) != cudaSuccess) {
        cudaGetLastError();
        set_error("cudaMalloc of the split-KV workspace failed");
        return nullptr;
    }
    d.ws_bytes = want;
    return d.ws;
}

}  // namespace sfa

extern "C" {
const char* sfa_last_error(void) { return sfa::g_last_error.c_str(); }
int sfa_abi_version(void) { return SFA_ABI_VERSION; }
uint64_t sfa_launch_count(void) { return sfa::g_launches.load(); }
void sfa_release_workspaces(void) {
    std::lock_guard<std::mutex> lk(sfa::g_mu);
    int cur = -1;
    cudaGetDevice(&cur);
    for (size_t i = 0; i < sfa::g_dev.size(); ++i) {
        if (!sfa::g_dev[i].ws) continue;
        cudaSetDevice((int)i);
        cudaDeviceSynchronize();
        cudaFree(sfa::g_dev[i].ws);
        sfa::g_dev[i].ws = nullptr;
        sfa::g_dev[i].ws_bytes = 0;
    }
    if (cur >= 0) cudaSetDevice(cur);
}
}
