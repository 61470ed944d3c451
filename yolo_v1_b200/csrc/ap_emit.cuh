// ap_emit.cuh -- included at the end of ap_reduce.cu: cumulative TP counts over
// the class-sorted records and the 11-point AP (metrics.py:415-436, 635-649).
#pragma once

namespace yh {

// the sorted records sit in buffer (passes run & 1): a = caller's arrays, b = workspace copies
__device__ __forceinline__ const uint64_t *sorted_flags(const SortCtl *ctl, const uint64_t *a, const uint64_t *b) {
    return final_buffer(ctl) ? b : a;
}

struct ApCtl {                        // zeroed per call, follows SortCtl in the workspace
    unsigned bit_total[kMaxBits];     // set bits per flag position over all records
    unsigned seg_bits[32][kMaxBits];  // set bits at positions before the start of class c
    unsigned long long binmax[1];     // [4*C*T*kLevels] best precision per recall bucket (flex)
};

// Per-tile population count of each flag bit.  Lane (b & 31) accumulates bit b.
__device__ __forceinline__ void count_bits_round(uint64_t flags, bool ok, int nbits, int lane,
                                                 unsigned acc[2]) {
    const unsigned lo = ok ? (unsigned)flags : 0u, hi = ok ? (unsigned)(flags >> 32) : 0u;
    unsigned ulo = __reduce_or_sync(0xffffffffu, lo), uhi = __reduce_or_sync(0xffffffffu, hi);
    if (nbits < 32) ulo &= (1u << nbits) - 1u;
    uhi &= nbits >= 64 ? 0xffffffffu : (nbits > 32 ? (1u << (nbits - 32)) - 1u : 0u);
    while (ulo) {
        const int b = __ffs(ulo) - 1;
        ulo &= ulo - 1;
        const unsigned m = __ballot_sync(0xffffffffu, (lo >> b) & 1u);
        if (lane == b) acc[0] += __popc(m);
    }
    while (uhi) {
        const int b = __ffs(uhi) - 1;
        uhi &= uhi - 1;
        const unsigned m = __ballot_sync(0xffffffffu, (hi >> b) & 1u);
        if (lane == b) acc[1] += __popc(m);
    }
}

// tile_bits[tile][b] = set bits b inside the tile (records in sorted order)
__global__ void __launch_bounds__(kSortThreads)
tile_bit_counts(const uint64_t *__restrict__ val_a, const uint64_t *__restrict__ val_b, int64_t Mcap,
                const int64_t *__restrict__ Mdev, const SortCtl *__restrict__ ctl, int nbits,
                unsigned *__restrict__ tile_bits) {
    __shared__ unsigned s_acc[kSortWarps][kMaxBits];
    const int64_t M = count_of(Mcap, Mdev);
    if ((int64_t)blockIdx.x * kSortTile >= M) return;  // rows of empty tiles are never read
    const uint64_t *val = sorted_flags(ctl, val_a, val_b);
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const int64_t first = (int64_t)blockIdx.x * kSortTile + wp * kWarpSpan + lane;
    unsigned acc[2] = {0, 0};
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int64_t idx = first + 32 * i;
        const bool ok = idx < M;
        const uint64_t f = ok ? __ldg(val + idx) : 0;
        count_bits_round(f, ok, nbits, lane, acc);
    }
    s_acc[wp][lane] = acc[0];
    s_acc[wp][lane + 32] = acc[1];
    __syncthreads();
    if (tid < kMaxBits) {
        unsigned t = 0;
#pragma unroll
        for (int q = 0; q < kSortWarps; ++q) t += s_acc[q][tid];
        tile_bits[(size_t)blockIdx.x * kMaxBits + tid] = t;
    }
}

// exclusive scan over tiles: one CTA per EIGHT flag bits (one 32-byte sector of a tile's row, so
// every access is a full sector and no two CTAs write the same one); a thread sums a short run of
// consecutive tiles, the CTA scans the 1,024 sums per bit, the thread writes its run back
constexpr int kScanCols = 8;
__global__ void __launch_bounds__(1024)
scan_tile_bits(unsigned *__restrict__ tile_bits, int64_t Mcap, const int64_t *__restrict__ Mdev,
               ApCtl *__restrict__ ap) {
    __shared__ unsigned s_w[kScanCols][32];
    // only the tiles that hold records (the count lives on the device; the capacity can be much larger)
    const int ntiles = (int)((count_of(Mcap, Mdev) + kSortTile - 1) / kSortTile);
    const int b0 = blockIdx.x * kScanCols, tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const int per = (ntiles + 1023) / 1024;
    const int t0 = tid * per, t1 = min(ntiles, t0 + per);
    unsigned sum[kScanCols];
#pragma unroll
    for (int j = 0; j < kScanCols; ++j) sum[j] = 0;
    for (int t = t0; t < t1; ++t) {
        const uint4 *row = reinterpret_cast<const uint4 *>(tile_bits + (size_t)t * kMaxBits + b0);
        const uint4 a = row[0], b = row[1];
        sum[0] += a.x, sum[1] += a.y, sum[2] += a.z, sum[3] += a.w;
        sum[4] += b.x, sum[5] += b.y, sum[6] += b.z, sum[7] += b.w;
    }
    unsigned inc[kScanCols];
#pragma unroll
    for (int j = 0; j < kScanCols; ++j) {
        inc[j] = sum[j];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned u = __shfl_up_sync(0xffffffffu, inc[j], o);
            if (lane >= o) inc[j] += u;
        }
        if (lane == 31) s_w[j][wp] = inc[j];
    }
    __syncthreads();
    if (wp < kScanCols) {  // warp j scans the 32 warp totals of bit b0 + j
        const unsigned x = s_w[wp][lane];
        unsigned xi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned u = __shfl_up_sync(0xffffffffu, xi, o);
            if (lane >= o) xi += u;
        }
        s_w[wp][lane] = xi - x;
        if (lane == 31) ap->bit_total[b0 + wp] = xi;
    }
    __syncthreads();
    unsigned run[kScanCols];
#pragma unroll
    for (int j = 0; j < kScanCols; ++j) run[j] = s_w[j][wp] + inc[j] - sum[j];
    for (int t = t0; t < t1; ++t) {
        uint4 *row = reinterpret_cast<uint4 *>(tile_bits + (size_t)t * kMaxBits + b0);
        const uint4 a = row[0], b = row[1];
        row[0] = make_uint4(run[0], run[1], run[2], run[3]);
        row[1] = make_uint4(run[4], run[5], run[6], run[7]);
        run[0] += a.x, run[1] += a.y, run[2] += a.z, run[3] += a.w;
        run[4] += b.x, run[5] += b.y, run[6] += b.z, run[7] += b.w;
    }
}

// seg_bits[c][b] = set bits b at sorted positions < start of class c (one CTA per class)
__global__ void __launch_bounds__(kSortThreads)
class_start_bits(const uint64_t *__restrict__ val_a, const uint64_t *__restrict__ val_b, int64_t Mcap,
                 const int64_t *__restrict__ Mdev, int nbits, const SortCtl *__restrict__ ctl,
                 const unsigned *__restrict__ tile_bits, ApCtl *__restrict__ ap) {
    __shared__ unsigned s_acc[kSortWarps][kMaxBits];
    const int64_t M = count_of(Mcap, Mdev);
    const uint64_t *val = sorted_flags(ctl, val_a, val_b);
    const int c = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const int64_t start = ctl->class_start[c];  // records whose class is < c
    const int64_t tile = start / kSortTile, tile0 = tile * kSortTile;
    unsigned acc[2] = {0, 0};
    // count inside [tile0, start); warp-strided rounds of 32 records
    for (int64_t base = tile0 + wp * 32; base < start; base += kSortThreads) {
        const int64_t idx = base + lane;
        const bool ok = idx < start && idx < M;
        const uint64_t f = ok ? __ldg(val + idx) : 0;
        count_bits_round(f, ok, nbits, lane, acc);
    }
    s_acc[wp][lane] = acc[0];
    s_acc[wp][lane + 32] = acc[1];
    __syncthreads();
    if (tid < kMaxBits) {
        unsigned t = 0;
#pragma unroll
        for (int q = 0; q < kSortWarps; ++q) t += s_acc[q][tid];
        const unsigned before = (start < M) ? tile_bits[(size_t)tile * kMaxBits + tid] : ap->bit_total[tid] - t;
        ap->seg_bits[c][tid] = before + t;
    }
}

constexpr int kEmitQueue = 128;  // queued true positives per warp between two drains

__device__ __forceinline__ void emit_queue(const uint2 *__restrict__ qa, const unsigned short *__restrict__ qb, int n,
                                           int lane, int C, int T, const int32_t *__restrict__ gt_count,
                                           ApCtl *__restrict__ ap) {
    for (int j = lane; j < n; j += 32) {
        const uint2 it = qa[j];
        const int c = qb[j] >> 8, b = qb[j] & 0xff;
        const int s = b / T, t = b - s * T;
        const unsigned ordinal = it.x;
        const int ngt = gt_count[s * C + c];
        const double prec = __ddiv_rn((double)ordinal, __dadd_rn((double)((long long)it.y + 1), 1e-6));
        const double rec = __ddiv_rn((double)ordinal, (double)ngt);
        int lvl = 0;
#pragma unroll
        for (int r = kLevels - 1; r >= 1; --r) {
            const double level = (r == kLevels - 1) ? 1.0 : __dmul_rn((double)r, 0.1);
            if (lvl == 0 && rec >= level) lvl = r;
        }
        atomicMax(&ap->binmax[(((size_t)s * C + c) * T + t) * kLevels + lvl],
                  (unsigned long long)__double_as_longlong(prec));
    }
}
// For every TP: ordinal within its class, rank within its class -> precision, recall
// -> atomicMax into the highest recall level it reaches.
__global__ void __launch_bounds__(kSortThreads)
emit_precisions(const uint64_t *__restrict__ val_a, const uint64_t *__restrict__ val_b, int64_t Mcap,
                const int64_t *__restrict__ Mdev, int C, int T, const SortCtl *__restrict__ ctl,
                const unsigned *__restrict__ tile_bits, const int32_t *__restrict__ gt_count,
                ApCtl *__restrict__ ap) {
    __shared__ unsigned s_acc[kSortWarps][kMaxBits];
    __shared__ uint2 s_qa[kSortWarps][kEmitQueue];
    __shared__ unsigned short s_qb[kSortWarps][kEmitQueue];
    const int64_t M = count_of(Mcap, Mdev);
    if ((int64_t)blockIdx.x * kSortTile >= M) return;
    const uint64_t *val = sorted_flags(ctl, val_a, val_b);
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const int nbits = 4 * T;  // the overall bit (4*T) only needs its total
    const int64_t first = (int64_t)blockIdx.x * kSortTile + wp * kWarpSpan + lane;
    uint64_t f[kSortItems];
    unsigned acc[2] = {0, 0};
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int64_t idx = first + 32 * i;
        const bool ok = idx < M;
        f[i] = ok ? __ldg(val + idx) : 0;
        count_bits_round(f[i], ok, nbits, lane, acc);
    }
    s_acc[wp][lane] = acc[0];
    s_acc[wp][lane + 32] = acc[1];
    __syncthreads();
    // running[b] for this warp = tile prefix + earlier warps of the tile; lane (b&31) holds bit b
    unsigned run[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int b = lane + 32 * h;
        unsigned r = tile_bits[(size_t)blockIdx.x * kMaxBits + b];
        for (int q = 0; q < wp; ++q) r += s_acc[q][b];
        run[h] = r;
    }
    const unsigned le = (lane == 31) ? 0xffffffffu : ((2u << lane) - 1u);
    // True positives are sparse and land on arbitrary lanes: each one is queued (ordinal, rank within its
    // class, class, bit) and the queue is worked off by all 32 lanes, so the two fp64 divisions and the
    // recall-level search of a TP never run with one active lane.
    int queued = 0;
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int64_t idx = first + 32 * i;
        const bool ok = idx < M;
        const unsigned lo = ok ? (unsigned)f[i] : 0u, hi = ok ? (unsigned)(f[i] >> 32) : 0u;
        const int c = (int)(f[i] >> 56);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const unsigned mine = h ? hi : lo;
            unsigned u = __reduce_or_sync(0xffffffffu, mine);
            const int lim = nbits - 32 * h;
            u &= lim >= 32 ? 0xffffffffu : (lim > 0 ? (1u << lim) - 1u : 0u);
            while (u) {
                const int bb = __ffs(u) - 1;
                u &= u - 1;
                const unsigned m = __ballot_sync(0xffffffffu, (mine >> bb) & 1u);
                const unsigned before = __shfl_sync(0xffffffffu, run[h], bb);
                if (lane == bb) run[h] += __popc(m);
                if (queued + 32 > kEmitQueue) {  // warp-uniform
                    __syncwarp();
                    emit_queue(s_qa[wp], s_qb[wp], queued, lane, C, T, gt_count, ap);
                    __syncwarp();
                    queued = 0;
                }
                if ((mine >> bb) & 1u) {
                    const int b = bb + 32 * h;
                    const unsigned ordinal = before + __popc(m & le) - ap->seg_bits[c][b];
                    const unsigned rank = (unsigned)(idx - (int64_t)ctl->class_start[c]);  // 0-based within class
                    const int at = queued + __popc(m & (le >> 1));
                    s_qa[wp][at] = make_uint2(ordinal, rank);
                    s_qb[wp][at] = (unsigned short)((c << 8) | b);
                }
                queued += __popc(m);
            }
        }
    }
    __syncwarp();
    emit_queue(s_qa[wp], s_qb[wp], queued, lane, C, T, gt_count, ap);
}

// ap[(s*C+c)*T+t]; totals[0] = TP of the overall pass, totals[1] = M
__global__ void finalize_ap(int C, int T, int64_t Mcap, const int64_t *__restrict__ Mdev,
                            const SortCtl *__restrict__ ctl, const int32_t *__restrict__ gt_count,
                            const ApCtl *__restrict__ apc, double *__restrict__ ap,
                            int64_t *__restrict__ totals, double *__restrict__ packed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t M = count_of(Mcap, Mdev);
    if (packed != nullptr) {  // everything compute() reads back, in one buffer of doubles
        if (i == 0) {
            packed[4 * C * T] = (double)apc->bit_total[4 * T];
            packed[4 * C * T + 1] = (double)M;
            packed[4 * C * T + 2] = (double)ctl->err;
        }
        if (i < 4 * C) packed[4 * C * T + 3 + i] = (double)gt_count[i];
    }
    if (i == 0) {
        totals[0] = apc->bit_total[4 * T];
        totals[1] = M;
    }
    if (i >= 4 * C * T) return;
    const int c = (i / T) % C, s = i / (T * C);
    const unsigned n_pred = ctl->class_start[c + 1] - ctl->class_start[c];
    const int ngt = gt_count[s * C + c];
    if (ngt == 0 || n_pred == 0) {  // metrics.py:375-379, 592-603
        ap[i] = 0.0;
        if (packed != nullptr) packed[i] = 0.0;
        return;
    }
    double p[kLevels];
    double best = 0.0;
#pragma unroll
    for (int r = kLevels - 1; r >= 1; --r) {
        const double v = __longlong_as_double((long long)apc->binmax[(size_t)i * kLevels + r]);
        best = v > best ? v : best;
        p[r] = best;
    }
    p[0] = 1.0;  // recall >= 0 always includes the (precision 1, recall 0) sentinel
    double a = 0.0;
#pragma unroll
    for (int r = 0; r < kLevels; ++r) a = __dadd_rn(a, __ddiv_rn(p[r], 11.0));
    ap[i] = a;
    if (packed != nullptr) packed[i] = a;
}

}  // namespace yh
