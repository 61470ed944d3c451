// partition.cu -- yh_partition_records: stable partition of compact detection
// records by the rank that owns their class (sm_100a).
//
// Multi-GPU mAP ("scaling variant"): average precision is
// independent per class (metrics.py:126-141 loops classes, :382 sorts one class at
// a time), so instead of replicating every record on every GPU the ranks exchange
// records all-to-all by class owner and each sorts only its own classes.  This
// kernel lays a rank's records out as n_parts contiguous runs, one per owner, each
// run keeping image order -- the order Python's stable sorted() breaks confidence
// ties by -- so the exchange is one all_to_all_single with those run lengths.
//
// Three streaming launches over M records (16 B each): per-tile run counts, a scan
// over tiles, and the scatter (each record read twice, written once).
#include "common.cuh"

namespace yh {

constexpr int kPartThreads = 256;
constexpr int kPartWarps = kPartThreads / 32;
constexpr int kPartRounds = 4;
constexpr int kPartSpan = 32 * kPartRounds;         // consecutive records one warp owns
constexpr int kPartTile = kPartWarps * kPartSpan;   // 1024 records per CTA
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

// tile_cnt[tile][p] = records of the tile that go to part p
__global__ void __launch_bounds__(kPartThreads)
part_tile_counts(const uint64_t *__restrict__ flags, int64_t Mcap, const int64_t *__restrict__ Mdev,
                 const __grid_constant__ OwnerTable tab, unsigned *__restrict__ tile_cnt) {
    __shared__ unsigned s_cnt[kMaxParts];
    const int64_t M = part_count_of(Mcap, Mdev);
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    if (tid < kMaxParts) s_cnt[tid] = 0;
    __syncthreads();
    const int64_t first = (int64_t)blockIdx.x * kPartTile + wp * kPartSpan + lane;
#pragma unroll
    for (int r = 0; r < kPartRounds; ++r) {
        const int64_t idx = first + 32 * r;
        const bool ok = idx < M;
        const int p = ok ? owner_of(tab, __ldg(flags + idx)) : 0;
        const unsigned act = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const unsigned peers = __match_any_sync(act, p);
            if ((peers & ((1u << lane) - 1u)) == 0) atomicAdd(&s_cnt[p], __popc(peers));
        }
    }
    __syncthreads();
    if (tid < kMaxParts) tile_cnt[(size_t)blockIdx.x * kMaxParts + tid] = s_cnt[tid];
}

// One CTA per part: exclusive scan of column p of tile_cnt over the tiles, in place.  A
// thread sums a short run of consecutive tiles, the CTA scans the 1,024 sums, the thread
// writes its run back.  part_count[p] = size of the part.
__global__ void __launch_bounds__(1024)
part_scan(unsigned *__restrict__ tile_cnt, int ntiles, int64_t *__restrict__ part_count) {
    __shared__ unsigned s_w[32];
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const int per = (ntiles + 1023) / 1024;
    const int t0 = tid * per, t1 = min(ntiles, t0 + per);
    unsigned sum = 0;
    for (int t = t0; t < t1; ++t) sum += tile_cnt[(size_t)t * kMaxParts + p];
    unsigned inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) s_w[wp] = inc;
    __syncthreads();
    if (wp == 0) {
        const unsigned x = s_w[lane];
        unsigned xi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned u = __shfl_up_sync(0xffffffffu, xi, o);
            if (lane >= o) xi += u;
        }
        s_w[lane] = xi - x;
        if (lane == 31) part_count[p] = xi;
    }
    __syncthreads();
    unsigned run = s_w[wp] + inc - sum;
    for (int t = t0; t < t1; ++t) {
        const unsigned c = tile_cnt[(size_t)t * kMaxParts + p];
        tile_cnt[(size_t)t * kMaxParts + p] = run;
        run += c;
    }
}

__global__ void __launch_bounds__(kPartThreads)
part_scatter(const uint64_t *__restrict__ key, const uint64_t *__restrict__ flags, int64_t Mcap,
             const int64_t *__restrict__ Mdev, const __grid_constant__ OwnerTable tab, const unsigned *__restrict__ tile_pre,
             const int64_t *__restrict__ part_count, int n_parts, uint64_t *__restrict__ out_key,
             uint64_t *__restrict__ out_flags) {
    __shared__ unsigned s_w[kPartWarps][kMaxParts];  // per-warp counts, then running cursors
    __shared__ int64_t part_base[kMaxParts];         // records of the parts before p
    const int64_t M = part_count_of(Mcap, Mdev);
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    s_w[wp][lane] = 0;
    if (tid < kMaxParts) {
        int64_t b = 0;
        for (int q = 0; q < tid && q < n_parts; ++q) b += part_count[q];
        part_base[tid] = b;
    }
    __syncwarp();
    const int64_t first = (int64_t)blockIdx.x * kPartTile + wp * kPartSpan + lane;
    const unsigned lt = (1u << lane) - 1u;
    uint64_t k[kPartRounds], f[kPartRounds];
    int part[kPartRounds];
#pragma unroll
    for (int r = 0; r < kPartRounds; ++r) {
        const int64_t idx = first + 32 * r;
        const bool ok = idx < M;
        k[r] = ok ? __ldg(key + idx) : 0;
        f[r] = ok ? __ldg(flags + idx) : 0;
        part[r] = ok ? owner_of(tab, f[r]) : -1;
        const unsigned act = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const unsigned peers = __match_any_sync(act, part[r]);
            if ((peers & lt) == 0) s_w[wp][part[r]] += __popc(peers);
        }
        __syncwarp();
    }
    __syncthreads();
    if (tid < kMaxParts) {  // warp-exclusive prefix per part, plus the tile's own prefix
        unsigned run = tile_pre[(size_t)blockIdx.x * kMaxParts + tid];
#pragma unroll
        for (int q = 0; q < kPartWarps; ++q) {
            const unsigned c = s_w[q][tid];
            s_w[q][tid] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kPartRounds; ++r) {
        const bool ok = part[r] >= 0;
        const unsigned act = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const unsigned peers = __match_any_sync(act, part[r]);
            const unsigned cur = s_w[wp][part[r]];
            const int64_t pos = part_base[part[r]] + cur + __popc(peers & lt);
            out_key[pos] = k[r];
            out_flags[pos] = f[r];
            __syncwarp(act);
            if ((peers & lt) == 0) s_w[wp][part[r]] = cur + __popc(peers);
        }
        __syncwarp();
    }
}

}  // namespace yh

using namespace yh;

static inline size_t part_up256(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" size_t yh_partition_workspace_bytes(int64_t M) {
    if (M < 0) return 0;
    const size_t ntiles = (size_t)((M + kPartTile - 1) / kPartTile) + 1;
    return 256 + part_up256(sizeof(int64_t) * kMaxParts) + sizeof(unsigned) * ntiles * kMaxParts;
}

extern "C" int yh_partition_records(const uint64_t *key, const uint64_t *flags, int64_t M,
                                    const int64_t *M_dev, const uint8_t *owner_of_class_host, int C, int n_parts,
                                    uint64_t *out_key, uint64_t *out_flags, int64_t *part_count,
                                    void *workspace, size_t workspace_bytes, void *stream) {
    if (C < 1 || C > 256 || n_parts < 1 || n_parts > kMaxParts) {
        set_error("yh_partition_records: C=%d / n_parts=%d outside this build (C<=256, parts<=%d)", C, n_parts,
                  kMaxParts);
        return YH_ERR_UNSUPPORTED;
    }
    if (M < 0 || M >= (1ll << 31) || !owner_of_class_host || !part_count || !workspace ||
        (M > 0 && (!key || !flags || !out_key || !out_flags))) {
        set_error("yh_partition_records: bad argument (M=%lld)", (long long)M);
        return YH_ERR_BAD_ARG;
    }
    if (workspace_bytes < yh_partition_workspace_bytes(M)) {
        set_error("yh_partition_records: workspace too small");
        return YH_ERR_WORKSPACE;
    }
    OwnerTable tab;
    for (int c = 0; c < 256; ++c) {
        const int o = c < C ? owner_of_class_host[c] : 0;
        if (o >= n_parts) {
            set_error("yh_partition_records: class %d is owned by part %d >= n_parts %d", c, o, n_parts);
            return YH_ERR_BAD_ARG;
        }
        tab.of_class[c] = (uint8_t)o;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (M == 0) {
        cudaMemsetAsync(part_count, 0, sizeof(int64_t) * n_parts, st);
        return check_launch("yh_partition_records", 0);
    }
    char *ws = (char *)workspace;
    ws += (256 - ((uintptr_t)ws & 255)) & 255;
    int64_t *part_base = (int64_t *)ws;
    unsigned *tile_cnt = (unsigned *)(ws + part_up256(sizeof(int64_t) * kMaxParts));
    const int ntiles = (int)((M + kPartTile - 1) / kPartTile);
    part_tile_counts<<<ntiles, kPartThreads, 0, st>>>(flags, M, M_dev, tab, tile_cnt);
    part_scan<<<n_parts, 1024, 0, st>>>(tile_cnt, ntiles, part_count);
    part_scatter<<<ntiles, kPartThreads, 0, st>>>(key, flags, M, M_dev, tab, tile_cnt, part_count, n_parts, out_key, out_flags);
    return check_launch("yh_partition_records", 3);
}
