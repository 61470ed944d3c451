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

constexpr int kTicketSlots = 256;
__device__ unsigned g_tickets[kTicketSlots];
static std::atomic<unsigned> g_ticket_next{0};

__global__ void zero_ticket(unsigned *t) { *t = 0u; }

unsigned *fresh_ticket(cudaStream_t stream) {
    static unsigned *base = nullptr;
    if (base == nullptr) cudaGetSymbolAddress((void **)&base, g_tickets);
    unsigned *t = base + (g_ticket_next.fetch_add(1, std::memory_order_relaxed) % kTicketSlots);
#ifdef YH_TICKET_MEMSET
    cudaMemsetAsync(t, 0, sizeof(unsigned), stream);
#else
    zero_ticket<<<1, 1, 0, stream>>>(t);  // a kernel, not a memset node: see yh_ap_reduce
#endif
    return t;
}

}  // namespace yh

extern "C" int yh_version(void) { return YH_VERSION; }
extern "C" const char *yh_last_error(void) { return yh::g_err; }
extern "C" long long yh_launch_count(void) { return yh::g_launches.load(std::memory_order_relaxed); }
