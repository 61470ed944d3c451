// api.cu -- version / error plumbing of libyolohead.so.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace yh {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char *what, int kernels_launched) {
    g_launches.fetch_add(kernels_launched, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

// Ticket words (one zeroed device counter per launch of a per-image kernel, see common.cuh).
// The array is a __device__ symbol, so it has a different address on every device: the base
// is looked up per device.  Eager launches take slots round-robin from a ring of kRingSlots
// (a slot is cleared on the launch's own stream, and 256 launches later the kernel that used
// it has long finished).  A launch that is being CAPTURED into a CUDA graph keeps its slot for
// as long as the graph lives, so it gets one from a separate region that is never handed out
// again (kGraphSlots per device; capturing more launches than that is an error).
constexpr int kRingSlots = 256;
constexpr int kGraphSlots = 3840;
__device__ unsigned g_tickets[kRingSlots + kGraphSlots];
static std::atomic<unsigned> g_ticket_next[kMaxDevices];
static std::atomic<unsigned> g_graph_next[kMaxDevices];
static std::atomic<unsigned *> g_ticket_base[kMaxDevices];

__global__ void zero_ticket(unsigned *t) { *t = 0u; }

int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < kMaxDevices) ? dev : 0;
}

unsigned *fresh_ticket(cudaStream_t stream) {
    const int dev = current_device();
    unsigned *base = g_ticket_base[dev].load(std::memory_order_acquire);
    if (base == nullptr) {
        cudaGetSymbolAddress((void **)&base, g_tickets);   // this device's copy of the symbol
        g_ticket_base[dev].store(base, std::memory_order_release);
    }
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cap);
    unsigned *t;
    if (cap == cudaStreamCaptureStatusActive) {
        const unsigned g = g_graph_next[dev].fetch_add(1, std::memory_order_relaxed);
        if (g >= (unsigned)kGraphSlots) {
            set_error("more than %d per-image launches captured into CUDA graphs on device %d", kGraphSlots, dev);
            return nullptr;
        }
        t = base + kRingSlots + g;
    } else {
        t = base + (g_ticket_next[dev].fetch_add(1, std::memory_order_relaxed) % kRingSlots);
    }
#ifdef YH_TICKET_MEMSET
    cudaMemsetAsync(t, 0, sizeof(unsigned), stream);
#else
    zero_ticket<<<1, 1, 0, stream>>>(t);  // a kernel, not a memset node: see yh_ap_reduce
#endif
    return t;
}

// true the first time it is called for (flag array, current device): per-device one-time set-up
// such as cudaFuncSetAttribute, which is a per-device property of a kernel
bool first_use_on_device(std::atomic<bool> *seen) {
    const int dev = current_device();
    return !seen[dev].exchange(true, std::memory_order_acq_rel);
}

}  // namespace yh

extern "C" int yh_version(void) { return YH_VERSION; }
extern "C" const char *yh_last_error(void) { return yh::g_err; }
extern "C" long long yh_launch_count(void) { return yh::g_launches.load(std::memory_order_relaxed); }
