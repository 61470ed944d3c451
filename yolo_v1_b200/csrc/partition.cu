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
#include <string.h>

#include "peer.cuh"

namespace yh {

constexpr int kPartThreads = 256;
constexpr int kPartWarps = kPartThreads / 32;
constexpr int kPartRounds = 4;
constexpr int kPartSpan = 32 * kPartRounds;         // consecutive records one warp owns
constexpr int kPartTile = kPartWarps * kPartSpan;   // 1024 records per CTA
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


// ===================================================================== peer exchange
// yh_exchange_records: the same stable partition, but every run is written straight
// into its owner's receive buffer through peer-mapped memory (NVLink / NVSwitch) --
// the partition and the all-to-all are one kernel, and the run lengths never visit the
// host.  Per rank one peer-mapped block holds a PeerCtl and the receive arrays:
//
//   publish   row[me][d] = records I send to rank d, my ground-truth counts and image
//             count, stored into EVERY rank's PeerCtl; then ready[me] = epoch (release.sys)
//   scatter   every CTA waits for all ready[] flags (acquire.sys), derives from the full
//             R x R matrix where my run starts in each receive buffer
//             (sum of row[r][d] over r < me: rank order = image order, what the
//             reference's stable sort breaks confidence ties by), orders its 1,024-record
//             tile by owner in shared memory and writes each run with consecutive lanes on
//             consecutive remote addresses; the last CTA to finish sets done[me] = epoch
//             on every rank
//   finish    one CTA waits for all done[] flags, then publishes the received record
//             count, the summed ground-truth counts and the global image count for
//             yh_ap_reduce (M_dev) -- still no host read.
//
// If a receive buffer would overflow (known to all ranks from the same matrix before a
// single record moves) nothing is scattered, info[2] is set on every rank and the host
// falls back to the NCCL exchange and grows the buffers.  Waits give up after a few
// seconds (info[3]) instead of hanging the GPU.

// first kernel of an exchange: clears the per-call words and advances the call counter
__global__ void peer_begin(ExchangeLocal *__restrict__ loc, PeerCtl *__restrict__ mine) {
    if (threadIdx.x == 0) {
        loc->finished_ctas = 0;
        loc->timed_out = 0;
        loc->overflow = 0;
        mine->calls = mine->calls + 1;
    }
}

// A CTA of the exchange owns a contiguous range of tiles (the same range in the count and in
// the scatter kernel), so per-tile prefixes are never stored: the count kernel leaves one
// total per CTA and part, the scatter kernel re-derives the rest walking its tiles in order.
__device__ __forceinline__ void cta_tile_range(int64_t M, int &t0, int &t1) {
    const int ntiles = (int)((M + kPartTile - 1) / kPartTile);
    const int per = (ntiles + (int)gridDim.x - 1) / (int)gridDim.x;
    t0 = min(ntiles, (int)blockIdx.x * per);
    t1 = min(ntiles, t0 + per);
}

// cta_cnt[cta][p] = records of the CTA's tile range that go to part p
__global__ void __launch_bounds__(kPartThreads)
peer_count(const uint64_t *__restrict__ flags, int64_t Mcap, const int64_t *__restrict__ Mdev,
           const __grid_constant__ OwnerTable tab, unsigned *__restrict__ cta_cnt) {
    __shared__ unsigned s_cnt[kMaxParts];
    const int64_t M = part_count_of(Mcap, Mdev);
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    if (tid < kMaxParts) s_cnt[tid] = 0;
    __syncthreads();
    int t0, t1;
    cta_tile_range(M, t0, t1);
    for (int tile = t0; tile < t1; ++tile) {
        const int64_t first = (int64_t)tile * kPartTile + wp * kPartSpan + lane;
        uint64_t f[kPartRounds];
#pragma unroll
        for (int r = 0; r < kPartRounds; ++r) f[r] = first + 32 * r < M ? __ldg(flags + first + 32 * r) : 0;
#pragma unroll
        for (int r = 0; r < kPartRounds; ++r) {
            const bool ok = first + 32 * r < M;
            const int p = ok ? owner_of(tab, f[r]) : 0;
            const unsigned act = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const unsigned grp = __match_any_sync(act, p);
                if ((grp & ((1u << lane) - 1u)) == 0) atomicAdd(&s_cnt[p], __popc(grp));
            }
        }
    }
    __syncthreads();
    if (tid < kMaxParts) cta_cnt[(size_t)blockIdx.x * kMaxParts + tid] = s_cnt[tid];
}

// One CTA: per part an exclusive scan of the CTA totals (in place), then this rank's message --
// run lengths, ground-truth counts, image count -- goes to every rank and ready[me] is raised.
__global__ void __launch_bounds__(1024)
peer_scan_publish(unsigned *__restrict__ cta_cnt, int nctas, const __grid_constant__ PeerTable peers, int world,
                  int me, unsigned epoch, const int32_t *__restrict__ gt_local, int n_gt, long long n_images) {
    __shared__ unsigned s_warp[32];
    __shared__ long long s_total[kMaxParts];
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    for (int p = 0; p < world; ++p) {
        const unsigned c = tid < nctas ? cta_cnt[(size_t)tid * kMaxParts + p] : 0;
        unsigned inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) s_warp[wp] = inc;
        __syncthreads();
        if (wp == 0) {
            const unsigned x = s_warp[lane];
            unsigned xi = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned u = __shfl_up_sync(0xffffffffu, xi, o);
                if (lane >= o) xi += u;
            }
            s_warp[lane] = xi - x;
            if (lane == 31) s_total[p] = xi;
        }
        __syncthreads();
        if (tid < nctas) cta_cnt[(size_t)tid * kMaxParts + p] = s_warp[wp] + inc - c;
        __syncthreads();
    }
    const int per = world + n_gt + 2;  // words of one rank's message
    for (int q = tid; q < world * per; q += blockDim.x) {
        const int d = q / per, i = q - d * per;
        PeerCtl *c = peers.ctl[d];
        if (i < world)
            c->row[me][i] = s_total[i];
        else if (i < world + n_gt)
            c->gt[me][i - world] = gt_local[i - world];
        else
            c->gt[me][kGtWords - 2 + (i - world - n_gt)] = (int)((unsigned long long)n_images >> (32 * (i - world - n_gt)));
    }
    __threadfence_system();
    __syncthreads();
    if (tid < world) st_release_sys(&peers.ctl[tid]->ready[me], epoch_of(epoch, peers.ctl[me]));
}

__global__ void __launch_bounds__(kPartThreads)
peer_scatter(const uint64_t *__restrict__ key, const uint64_t *__restrict__ flags, int64_t Mcap,
             const int64_t *__restrict__ Mdev, const __grid_constant__ OwnerTable tab,
             const unsigned *__restrict__ cta_pre, const __grid_constant__ PeerTable peers, int world, int me,
             unsigned epoch, long long capacity, ExchangeLocal *__restrict__ loc) {
    __shared__ unsigned s_w[kPartWarps][kMaxParts];  // per-warp counts, then running cursors
    __shared__ long long dst_base[kMaxParts];        // where my run starts in rank d's receive buffer
    __shared__ unsigned tile_start[kMaxParts + 1];   // tile-local start of each part
    __shared__ uint64_t s_key[kPartTile], s_flg[kPartTile];
    __shared__ int s_state;
    const int64_t M = part_count_of(Mcap, Mdev);
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const PeerCtl *mine = peers.ctl[me];
    epoch = epoch_of(epoch, mine);
    if (tid == 0) s_state = 0;
    __syncthreads();
    // persistent CTAs: one wait for the peers' rows per CTA, one system fence at the end
    if (tid < world && !wait_flag(&mine->ready[tid], epoch, mine->wait_ms)) s_state = 1;
    __syncthreads();
    if (s_state) {
        if (tid == 0) loc->timed_out = 1;
        return;
    }
    if (tid < kMaxParts) {
        long long before = 0, total = 0;
        if (tid < world)
            for (int r = 0; r < world; ++r) {
                const long long c = __ldcv(&mine->row[r][tid]);  // written by a peer: never from L1
                if (r < me) before += c;
                total += c;
            }
        // start of my run in rank tid's buffer + what the CTAs before this one put there
        dst_base[tid] = before + cta_pre[(size_t)blockIdx.x * kMaxParts + tid];
        if (total > capacity) s_state = 2;
    }
    __syncthreads();
    const bool overflow = s_state == 2;
    if (overflow && tid == 0 && blockIdx.x == 0) loc->overflow = 1;
    int t0, t1;
    cta_tile_range(M, t0, t1);
    if (overflow) t1 = t0;
    const unsigned lt = (1u << lane) - 1u;
    for (int tile = t0; tile < t1; ++tile) {
        s_w[wp][lane] = 0;
        __syncwarp();
        const int64_t first = (int64_t)tile * kPartTile + wp * kPartSpan + lane;
        uint64_t k[kPartRounds], f[kPartRounds];
        int part[kPartRounds];
#pragma unroll
        for (int r = 0; r < kPartRounds; ++r) {
            const int64_t idx = first + 32 * r;
            const bool ok = idx < M;
            k[r] = ok ? __ldg(key + idx) : 0;
            f[r] = ok ? __ldg(flags + idx) : 0;
        }
#pragma unroll
        for (int r = 0; r < kPartRounds; ++r) {
            const bool ok = first + 32 * r < M;
            part[r] = ok ? owner_of(tab, f[r]) : -1;
            const unsigned act = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const unsigned peers_m = __match_any_sync(act, part[r]);
                if ((peers_m & lt) == 0) s_w[wp][part[r]] += __popc(peers_m);
            }
            __syncwarp();
        }
        __syncthreads();
        if (tid < kMaxParts) {  // per part: exclusive prefix over the warps, then over the parts (tile-local)
            unsigned run = 0;
#pragma unroll
            for (int q = 0; q < kPartWarps; ++q) {
                const unsigned c = s_w[q][tid];
                s_w[q][tid] = run;
                run += c;
            }
            unsigned inc = run;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned u = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += u;
            }
            tile_start[tid] = inc - run;
            if (tid == kMaxParts - 1) tile_start[kMaxParts] = inc;
            dst_base[tid] -= (long long)(inc - run);  // so that position = dst_base[p] + tile-local index
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < kPartRounds; ++r) {
            const bool ok = part[r] >= 0;
            const unsigned act = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const unsigned peers_m = __match_any_sync(act, part[r]);
                const unsigned cur = s_w[wp][part[r]];
                const unsigned pos = tile_start[part[r]] + cur + __popc(peers_m & lt);
                s_key[pos] = k[r];
                s_flg[pos] = f[r];
                __syncwarp(act);
                if ((peers_m & lt) == 0) s_w[wp][part[r]] = cur + __popc(peers_m);
            }
            __syncwarp();
        }
        __syncthreads();
        // consecutive threads -> consecutive addresses of one run in one receive buffer
        const unsigned n_tile = tile_start[kMaxParts];
        for (unsigned i = tid; i < n_tile; i += kPartThreads) {
            int p = 0;
#pragma unroll
            for (int h = kMaxParts / 2; h > 0; h >>= 1)
                if (tile_start[p + h] <= i) p += h;
            const long long pos = dst_base[p] + i;  // = run start + records before the tile + (i - tile_start[p])
            peers.key[p][pos] = s_key[i];
            peers.flags[p][pos] = s_flg[i];
        }
        __syncthreads();
        if (tid < kMaxParts) dst_base[tid] += (long long)tile_start[tid + 1];  // undo the shift, add the tile's count
        __syncthreads();
    }
    // completion: my stores are visible system-wide before the count moves; the last CTA raises done[me]
    __threadfence_system();
    __syncthreads();
    if (tid == 0) s_state = (atomicAdd(&loc->finished_ctas, 1u) == gridDim.x - 1) ? 3 : 0;
    __syncthreads();
    if (s_state == 3 && tid < world) {
        __threadfence_system();
        st_release_sys(&peers.ctl[tid]->done[me], epoch);
    }
}

__global__ void __launch_bounds__(256)
peer_finish(const __grid_constant__ PeerTable peers, int world, int me, unsigned epoch, int n_gt,
            const ExchangeLocal *__restrict__ loc, int32_t *__restrict__ gt_global, int64_t *__restrict__ info) {
    __shared__ int s_bad;
    const int tid = threadIdx.x;
    const PeerCtl *mine = peers.ctl[me];
    epoch = epoch_of(epoch, mine);
    if (tid == 0) s_bad = 0;
    __syncthreads();
    if (tid < world && !wait_flag(&mine->done[tid], epoch, mine->wait_ms)) s_bad = 1;
    __syncthreads();
    const bool bad = s_bad || loc->timed_out;
    const bool overflow = loc->overflow != 0;
    for (int i = tid; i < n_gt; i += blockDim.x) {
        int s = 0;
        for (int r = 0; r < world; ++r) s += __ldcv(&mine->gt[r][i]);
        gt_global[i] = s;
    }
    if (tid == 0) {
        long long total = 0, images = 0;
        for (int r = 0; r < world; ++r) {
            total += __ldcv(&mine->row[r][me]);
            images += (long long)(((unsigned long long)(unsigned)__ldcv(&mine->gt[r][kGtWords - 1]) << 32) |
                                  (unsigned)__ldcv(&mine->gt[r][kGtWords - 2]));
        }
        info[0] = (bad || overflow) ? 0 : total;
        info[1] = images;
        info[2] = overflow ? 1 : 0;
        info[3] = bad ? 1 : 0;
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

// Sum of the ranks' result vectors (AP table, totals, error flag) in rank order, through the
// same peer blocks: store mine into everybody's slot, raise res_ready, wait for the others,
// add.  Every class is non-zero on exactly one rank and the totals are integers, so the
// sum is exact in any order.  A rank that saw its peers' results has, by stream order,
// peers that are done with their receive buffers: this is also what orders two exchanges.
namespace yh {
__global__ void __launch_bounds__(1024)
peer_sum_results(const __grid_constant__ PeerTable peers, int world, int me, unsigned epoch, int n, int err_index,
                 double *__restrict__ vec) {
    __shared__ int s_bad;
    const int tid = threadIdx.x;
    const PeerCtl *mine = peers.ctl[me];
    epoch = epoch_of(epoch, mine);
    if (tid == 0) s_bad = 0;
    for (int q = tid; q < world * n; q += blockDim.x) {
        const int d = q / n, i = q - d * n;
        peers.ctl[d]->res[me][i] = vec[i];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < world) {
        st_release_sys(&peers.ctl[tid]->res_ready[me], epoch);
        if (!wait_flag(&mine->res_ready[tid], epoch, mine->wait_ms)) s_bad = 1;
    }
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {
        double s = 0.0;
        for (int r = 0; r < world; ++r) s += __ldcv(&mine->res[r][i]);
        vec[i] = (i == err_index && s_bad) ? 1.0 : s;
    }
}
}  // namespace yh

extern "C" int yh_peer_sum(double *vec, int n, int err_index, int world, int rank, void *const *peer_block_host,
                           unsigned epoch, void *stream) {
    if (!vec || n < 1 || n > kResWords || world < 1 || world > kMaxParts || rank < 0 || rank >= world ||
        !peer_block_host) {
        set_error("yh_peer_sum: bad argument (n=%d world=%d rank=%d)", n, world, rank);
        return YH_ERR_BAD_ARG;
    }
    PeerTable peers;
    for (int r = 0; r < kMaxParts; ++r) {
        peers.ctl[r] = (PeerCtl *)peer_block_host[r < world ? r : rank];
        peers.key[r] = nullptr;
        peers.flags[r] = nullptr;
        if (!peers.ctl[r]) {
            set_error("yh_peer_sum: peer block %d is null", r);
            return YH_ERR_BAD_ARG;
        }
    }
    peer_sum_results<<<1, 1024, 0, (cudaStream_t)stream>>>(peers, world, rank, epoch, n, err_index, vec);
    return check_launch("yh_peer_sum");
}

// ------------------------------------------------------------------ peer-mapped memory
// The one place this library allocates: a block other processes can map must come from
// cudaMalloc (not from the caller's caching allocator) to be exportable as an IPC handle.
extern "C" int yh_peer_alloc(size_t bytes, void **ptr, void *handle64) {
    if (!ptr || !handle64 || bytes == 0) {
        set_error("yh_peer_alloc: bad argument");
        return YH_ERR_BAD_ARG;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e == cudaSuccess) e = cudaMemset(p, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        set_error("yh_peer_alloc(%zu): %s", bytes, cudaGetErrorString(e));
        if (p) cudaFree(p);
        cudaGetLastError();
        return (int)e;
    }
    memcpy(handle64, &h, 64);
    *ptr = p;
    return 0;
}

// how long this rank's kernels wait for a peer's flag before they give up (info[3] / the error word of yh_peer_sum)
extern "C" int yh_peer_set_wait_ms(void *own_block, unsigned wait_ms) {
    if (!own_block) {
        set_error("yh_peer_set_wait_ms: bad argument");
        return YH_ERR_BAD_ARG;
    }
    const cudaError_t e = cudaMemcpy(&((yh::PeerCtl *)own_block)->wait_ms, &wait_ms, sizeof(unsigned), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        set_error("yh_peer_set_wait_ms: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return (int)e;
    }
    return 0;
}

extern "C" int yh_peer_open(const void *handle64, void **ptr) {
    if (!ptr || !handle64) {
        set_error("yh_peer_open: bad argument");
        return YH_ERR_BAD_ARG;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void *p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        set_error("yh_peer_open: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return (int)e;
    }
    *ptr = p;
    return 0;
}

extern "C" int yh_peer_close(void *ptr) {
    const cudaError_t e = ptr ? cudaIpcCloseMemHandle(ptr) : cudaSuccess;
    if (e != cudaSuccess) {
        set_error("yh_peer_close: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return (int)e;
    }
    return 0;
}

extern "C" int yh_peer_free(void *ptr) {
    const cudaError_t e = ptr ? cudaFree(ptr) : cudaSuccess;
    if (e != cudaSuccess) {
        set_error("yh_peer_free: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return (int)e;
    }
    return 0;
}

// Layout of a rank's peer block: PeerCtl | key[capacity] | flags[capacity]
extern "C" size_t yh_peer_block_bytes(int64_t capacity) {
    if (capacity < 0) return 0;
    return part_up256(sizeof(PeerCtl)) + 2 * part_up256(sizeof(uint64_t) * (size_t)capacity);
}
extern "C" size_t yh_peer_key_offset(int64_t capacity) {
    (void)capacity;
    return part_up256(sizeof(PeerCtl));
}
extern "C" size_t yh_peer_flags_offset(int64_t capacity) {
    return part_up256(sizeof(PeerCtl)) + part_up256(sizeof(uint64_t) * (size_t)(capacity < 0 ? 0 : capacity));
}

constexpr int kExchangeCtas = 148 * 6;   // <= 1024: peer_scan_publish scans one CTA total per thread

extern "C" size_t yh_exchange_workspace_bytes(int64_t M) {
    if (M < 0) return 0;
    return 256 + 256 + sizeof(unsigned) * (size_t)kExchangeCtas * kMaxParts;
}

extern "C" int yh_exchange_records(const uint64_t *key, const uint64_t *flags, int64_t M, const int64_t *M_dev,
                                   const uint8_t *owner_of_class_host, int C, int world, int rank,
                                   void *const *peer_block_host, int64_t capacity, unsigned epoch,
                                   const int32_t *gt_count_local, int64_t n_images_local,
                                   int32_t *gt_count_global, int64_t *info, void *workspace,
                                   size_t workspace_bytes, void *stream) {
    if (C < 1 || C > 32 || world < 1 || world > kMaxParts) {
        set_error("yh_exchange_records: C=%d / world=%d outside this build (C<=32, world<=%d)", C, world, kMaxParts);
        return YH_ERR_UNSUPPORTED;
    }
    if (M < 0 || M >= (1ll << 31) || rank < 0 || rank >= world || capacity < 0 ||
        !owner_of_class_host || !peer_block_host || !gt_count_local || !gt_count_global || !info || !workspace ||
        (M > 0 && (!key || !flags))) {
        set_error("yh_exchange_records: bad argument (M=%lld rank=%d world=%d)", (long long)M, rank, world);
        return YH_ERR_BAD_ARG;
    }
    if (workspace_bytes < yh_exchange_workspace_bytes(M)) {
        set_error("yh_exchange_records: workspace too small");
        return YH_ERR_WORKSPACE;
    }
    OwnerTable tab;
    for (int c = 0; c < 256; ++c) {
        const int o = c < C ? owner_of_class_host[c] : 0;
        if (o >= world) {
            set_error("yh_exchange_records: class %d is owned by rank %d >= world %d", c, o, world);
            return YH_ERR_BAD_ARG;
        }
        tab.of_class[c] = (uint8_t)o;
    }
    PeerTable peers;
    for (int r = 0; r < kMaxParts; ++r) {
        char *base = (char *)peer_block_host[r < world ? r : rank];
        if (!base) {
            set_error("yh_exchange_records: peer block %d is null", r);
            return YH_ERR_BAD_ARG;
        }
        peers.ctl[r] = (PeerCtl *)base;
        peers.key[r] = (uint64_t *)(base + yh_peer_key_offset(capacity));
        peers.flags[r] = (uint64_t *)(base + yh_peer_flags_offset(capacity));
    }
    cudaStream_t st = (cudaStream_t)stream;
    char *ws = (char *)workspace;
    ws += (256 - ((uintptr_t)ws & 255)) & 255;
    ExchangeLocal *loc = (ExchangeLocal *)ws;
    unsigned *cta_cnt = (unsigned *)(ws + 256);
    const int ntiles = (int)((M + kPartTile - 1) / kPartTile);
    int grid = ntiles < kExchangeCtas ? ntiles : kExchangeCtas;  // persistent: a CTA owns a contiguous tile range
    if (grid < 1) grid = 1;
    peer_begin<<<1, 32, 0, st>>>(loc, peers.ctl[rank]);
    peer_count<<<grid, kPartThreads, 0, st>>>(flags, M, M_dev, tab, cta_cnt);
    peer_scan_publish<<<1, 1024, 0, st>>>(cta_cnt, grid, peers, world, rank, epoch, gt_count_local, 4 * C,
                                          (long long)n_images_local);
    peer_scatter<<<grid, kPartThreads, 0, st>>>(key, flags, M, M_dev, tab, cta_cnt, peers, world, rank, epoch,
                                                (long long)capacity, loc);
    peer_finish<<<1, 256, 0, st>>>(peers, world, rank, epoch, 4 * C, loc, gt_count_global, info);
    return check_launch("yh_exchange_records", 5);
}
#include "exchange_images.cuh"
