// peer.cuh -- what the kernels that talk to other GPUs through peer-mapped memory share
// (partition.cu: yh_exchange_records, yh_peer_sum; exchange_images.cu: yh_exchange_images).
#pragma once

#include "common.cuh"

namespace yh {

constexpr int kMaxParts = 32;

struct OwnerTable {
    uint8_t of_class[256];
};

__device__ __forceinline__ int64_t part_count_of(int64_t cap, const int64_t *dev) {
    if (dev == nullptr) return cap;
    const int64_t m = *dev;
    return m < cap ? m : cap;
}

__device__ __forceinline__ int owner_of(const OwnerTable &tab, uint64_t flags) { return tab.of_class[flags >> 56]; }


constexpr int kGtWords = 4 * 32 + 2;  // 4 subsets x (C <= 32) counts + image count (2 words)

constexpr int kResWords = 4 * 32 * YH_MAX_THRESHOLDS + 4;  // AP table + totals + error flag of one rank

struct PeerCtl {
    unsigned ready[kMaxParts];
    unsigned done[kMaxParts];
    unsigned res_ready[kMaxParts];
    unsigned calls;              // exchanges this rank has started: the epoch when the caller passes 0
    unsigned wait_ms;            // how long a wait on a peer's flag may last (0 = the default); yh_peer_set_wait_ms
    unsigned pad[kMaxParts - 2];
    long long row[kMaxParts][kMaxParts];
    int gt[kMaxParts][kGtWords];
    double res[kMaxParts][kResWords];
};

struct PeerTable {
    PeerCtl *ctl[kMaxParts];
    uint64_t *key[kMaxParts];
    uint64_t *flags[kMaxParts];
};

struct ExchangeLocal {        // head of the workspace, zeroed per call
    unsigned finished_ctas;
    unsigned timed_out;
    unsigned overflow;
    unsigned pad;
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
constexpr unsigned kDefaultWaitMs = 4000;

// epoch 0 = "the library's own call counter" (kept in the rank's block, advanced by peer_begin):
// what a CUDA-graph replay needs, where kernel arguments are frozen
__device__ __forceinline__ unsigned epoch_of(unsigned epoch, const PeerCtl *mine) {
    return epoch ? epoch : *reinterpret_cast<const volatile unsigned *>(&mine->calls);
}

// true once *flag == epoch; false if the limit ran out
__device__ __forceinline__ bool wait_flag(const unsigned *flag, unsigned epoch, unsigned wait_ms = 0) {
    const unsigned long long kWaitLimitNs = (unsigned long long)(wait_ms ? wait_ms : kDefaultWaitMs) * 1000000ull;
    if (ld_acquire_sys(flag) == epoch) return true;
    const unsigned long long t0 = now_ns();
    while (ld_acquire_sys(flag) != epoch) {
        __nanosleep(200);
        if (now_ns() - t0 > kWaitLimitNs) return false;
    }
    return true;
}


}  // namespace yh
