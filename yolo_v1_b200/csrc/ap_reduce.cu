// ap_reduce.cu -- record compaction, stable 64-bit radix sort, and the
// cumulative-TP / 11-point-AP reduction (sm_100a).
//
// Replaces the global half of mAPMetric.compute(): sorted(class_predictions,
// key=conf, reverse=True) (metrics.py:382, 606), np.cumsum of tp/fp, the
// precision/recall arrays and the 11-point interpolation (metrics.py:415-436,
// 635-649), for all classes x thresholds x ground-truth subsets at once.
//
// Pipeline (every step a bandwidth-bound streaming kernel over M records):
//   compact   fixed-stride per-image records -> dense arrays in image order
//   sort      LSD radix, 8-bit digits, one kernel per digit: tiles take dynamic
//             ids and publish per-digit counts; later tiles look back (chained
//             scan), so each pass reads and writes every record exactly once.
//             Stable, therefore equal confidences keep (image, per-image) order,
//             which is what Python's stable sorted() leaves.  Last pass = class.
//   bits      per-tile population count of every TP flag bit, scanned over tiles
//   emit      for every TP: ordinal m (cumulative TP in its class), rank i ->
//             precision m/(i+1+1e-6), recall m/n_gt; atomicMax into the recall
//             bucket it reaches
//   final     suffix-max over the 11 buckets, ap = sum(p_interp/11)
#include "common.cuh"

namespace yh {

constexpr int kSortThreads = 256;
#ifndef YH_SORT_ITEMS
#define YH_SORT_ITEMS 8
#endif
#ifndef YH_SORT_LOOK
#define YH_SORT_LOOK 4
#endif
#ifndef YH_SORT_MINBLOCKS
#define YH_SORT_MINBLOCKS 4
#endif
constexpr int kSortItems = YH_SORT_ITEMS;
constexpr int kSortTile = kSortThreads * kSortItems;  // 2048 records per tile
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kWarpSpan = 32 * kSortItems;            // records a warp owns inside a tile
constexpr int kPasses = 9;                             // 8 key digits + the class digit
constexpr int kMaxBits = 64;
constexpr int kLevels = 11;                            // np.linspace(0, 1, 11)
enum : unsigned { kFlagAgg = 1u << 30, kFlagIncl = 2u << 30, kValMask = (1u << 30) - 1u };

// ------------------------------------------------------------------ compaction
__global__ void __launch_bounds__(1024)
scan_counts_local(const int32_t *__restrict__ cnt, int N, int64_t *__restrict__ off,
                  int64_t *__restrict__ block_tot) {
    __shared__ int64_t s_w[32];
    const int i = blockIdx.x * 1024 + threadIdx.x, lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
    int64_t v = (i < N) ? cnt[i] : 0, inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int64_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_w[wp] = inc;
    __syncthreads();
    if (wp == 0) {
        int64_t x = s_w[lane], xi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t t = __shfl_up_sync(0xffffffffu, xi, o);
            if (lane >= o) xi += t;
        }
        s_w[lane] = xi - x;
        if (lane == 31) block_tot[blockIdx.x] = xi;
    }
    __syncthreads();
    if (i < N) off[i] = s_w[wp] + inc - v;
}

__global__ void __launch_bounds__(1024)
scan_block_totals(int64_t *__restrict__ block_tot, int nb, int64_t *__restrict__ total_out,
                  int64_t *__restrict__ running, int64_t *__restrict__ base_copy) {
    // single CTA; sequential chunks of 1024 with a running carry
    __shared__ int64_t s_w[32];
    __shared__ int64_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
    for (int base = 0; base < nb; base += 1024) {
        const int i = base + threadIdx.x;
        int64_t v = (i < nb) ? block_tot[i] : 0, inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int64_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_w[wp] = inc;
        __syncthreads();
        if (wp == 0) {
            int64_t x = s_w[lane], xi = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int64_t t = __shfl_up_sync(0xffffffffu, xi, o);
                if (lane >= o) xi += t;
            }
            s_w[lane] = xi - x;
        }
        __syncthreads();
        const int64_t carry = s_carry;
        if (i < nb) block_tot[i] = carry + s_w[wp] + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + s_w[wp] + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        // records already in the output arrays (chained calls append behind them)
        const int64_t base = running != nullptr ? *running : 0;
        *base_copy = base;
        *total_out = base + s_carry;
        if (running != nullptr) *running = base + s_carry;
    }
}

// one warp per image: add the block base to its offset and copy its records
__global__ void __launch_bounds__(256)
gather_records(const int32_t *__restrict__ cnt, const uint64_t *__restrict__ rec_key,
               const uint64_t *__restrict__ rec_flags, int N, int K, int64_t *__restrict__ off,
               const int64_t *__restrict__ block_base, const int64_t *__restrict__ base_copy,
               uint64_t *__restrict__ out_key, uint64_t *__restrict__ out_flags) {
    const int img = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (img >= N) return;
    const int64_t o = off[img] + block_base[img >> 10] + *base_copy;
    const int n = cnt[img];
    __syncwarp();
    if (lane == 0) off[img] = o;
    for (int q = lane; q < n; q += 32) {
        out_key[o + q] = __ldg(rec_key + (size_t)img * K + q);
        out_flags[o + q] = __ldg(rec_flags + (size_t)img * K + q);
    }
}

}  // namespace yh

using namespace yh;

extern "C" size_t yh_compact_workspace_bytes(int N) {
    if (N < 0) return 0;
    return 256 + sizeof(int64_t) * ((size_t)(N + 1023) / 1024 + 1);
}

extern "C" int yh_compact_records(const int32_t *n_keep, const uint64_t *rec_key,
                                  const uint64_t *rec_flags, int N, int K, int64_t *offsets,
                                  uint64_t *out_key, uint64_t *out_flags, int64_t *running_total,
                                  void *workspace, size_t workspace_bytes, void *stream) {
    if (N < 0 || K <= 0 || !n_keep || !rec_key || !rec_flags || !offsets || !out_key || !out_flags ||
        !workspace) {
        set_error("yh_compact_records: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (workspace_bytes < yh_compact_workspace_bytes(N)) {
        set_error("yh_compact_records: workspace too small");
        return YH_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int64_t *block_tot = (int64_t *)((char *)workspace + 256);
    const int nb = (N + 1023) / 1024;
    int64_t *base_copy = (int64_t *)workspace;
    if (N == 0) {
        scan_block_totals<<<1, 1024, 0, st>>>(block_tot, 0, offsets, running_total, base_copy);
        return check_launch("yh_compact_records");
    }
    scan_counts_local<<<nb, 1024, 0, st>>>(n_keep, N, offsets, block_tot);
    scan_block_totals<<<1, 1024, 0, st>>>(block_tot, nb, offsets + N, running_total, base_copy);
    gather_records<<<(N + 7) / 8, 256, 0, st>>>(n_keep, rec_key, rec_flags, N, K, offsets, block_tot, base_copy,
                                                out_key, out_flags);
    return check_launch("yh_compact_records", 3);
}

// ======================================================================= sort
#ifndef YH_HIST_CTAS_PER_SM
#define YH_HIST_CTAS_PER_SM 4
#endif
namespace yh {

constexpr int kShortPasses = 4;   // passes the host launches: 32 composite bits
constexpr int kShortBits = 8 * kShortPasses;
constexpr int kRunMax = 32;       // longest run of equal leading bits that fix_runs orders
constexpr int kFallbackCtasPerSm = 2;

struct SortCtl {                  // lives at the head of the workspace, zeroed per call
    unsigned tile_counter[16];    // dynamic tile ids, one counter per pass
    unsigned err;                 // set if a look-back spin gave up (never expected)
    unsigned overflow;            // the short plan could not finish the order: radix_fallback sorts by the full plan
    unsigned long long key_or;    // OR of all keys
    unsigned long long key_nand;  // OR of all ~keys: a bit varies over the input iff set in both
    unsigned work_count;          // runs queued by find_runs for fix_runs
    unsigned barrier;             // grid barrier of radix_fallback
    unsigned pad1[8];
    unsigned hist[kPasses][256];  // global digit histograms -> exclusive digit offsets
    unsigned class_start[40];     // records whose class is < c  (c = 0..32)
    unsigned tile_counter2[16];   // the same two, for the passes of radix_fallback
    unsigned hist2[kPasses][256];
};

// The sort key is the pair (class, confidence key).  Only the key bits that differ
// somewhere in the input can influence the order, and they form one range [lo, lo+nb)
// (sign / high exponent bits above it and the zero tail of a 24 x 24-bit product below
// it are common to all records).  The order is that of the composite
// (class << nb) | key[lo, lo+nb), nb + 5 bits.
//
// Short plan: the LSD passes run over the LEADING 32 bits of the composite only (four
// passes; `shift` = the bits left out).  Stable passes leave records that agree in those
// 32 bits in input order, so what is left to do is to order the few short runs of such
// records by their remaining bits (find_runs + fix_runs: 4 M products of two fp32 in 20
// classes leave a few thousand runs of two or three).  A run longer than kRunMax that is
// out of order, more runs than the work list holds, or more than 59 varying key bits
// raise `overflow`, and radix_fallback then sorts by all ceil((nb + 5) / 8) digits.
struct KeyPlan {
    int lo, nb, npasses, shift, full;
};
constexpr int kCompositeMaxBits = 59;  // (class << nb) | key bits must fit one word
__device__ __forceinline__ KeyPlan key_plan(const SortCtl *ctl) {
    const unsigned long long vary = ctl->key_or & ctl->key_nand;
    KeyPlan p;
    p.lo = vary ? __ffsll((long long)vary) - 1 : 0;
    p.nb = vary ? 64 - __clzll((long long)vary) - p.lo : 0;
    p.full = (p.nb + 5 + 7) / 8;
    p.shift = 0;
    p.npasses = p.full;
    if (p.nb > kCompositeMaxBits) {
        p.npasses = 0;  // everything is left to radix_fallback
    } else if (p.nb + 5 > kShortBits) {
        p.shift = p.nb + 5 - kShortBits;
        p.npasses = kShortPasses;
    }
    return p;
}
__device__ __forceinline__ bool plan_needs_fallback(const KeyPlan &pl, const SortCtl *ctl) {
    return pl.nb > kCompositeMaxBits || (pl.shift > 0 && ctl->overflow != 0);
}
// buffer (0 = caller's arrays, 1 = workspace copies) that holds the sorted records at the end
__device__ __forceinline__ int final_buffer(const SortCtl *ctl) {
    const KeyPlan pl = key_plan(ctl);
    return (plan_needs_fallback(pl, ctl) ? pl.full : pl.npasses) & 1;
}
// (class << nb) | key[lo, lo+nb) as one word.  Needs nb <= kCompositeMaxBits.
__device__ __forceinline__ uint64_t composite_of(uint64_t key, uint64_t val, const KeyPlan &pl) {
    const uint64_t mask = pl.nb ? (~0ull >> (64 - pl.nb)) : 0ull;
    return ((key >> pl.lo) & mask) | (((val >> 56) & 31ull) << pl.nb);
}
// digit `pass` of the composite without its `shift` low bits
__device__ __forceinline__ unsigned digit_of(uint64_t key, uint64_t val, int pass, const KeyPlan &pl) {
    if (pl.nb <= kCompositeMaxBits) return (unsigned)(composite_of(key, val, pl) >> (pl.shift + 8 * pass)) & 0xffu;
    // more than 59 varying bits (radix_fallback only): the class digit is assembled by hand
    const unsigned cls = (unsigned)(val >> 56) & 31u;
    const int sh = 8 * pass;
    unsigned d;
    if (sh >= pl.nb) {
        d = cls >> (sh - pl.nb);
    } else {
        d = (unsigned)((key >> pl.lo) >> sh);
        const int left = pl.nb - sh;  // key bits available to this digit
        if (left < 8) d = (d & ((1u << left) - 1u)) | (cls << left);
    }
    return d & 0xffu;
}
// The same digit from per-pass constants (composite plans only): the key's share is
// (key >> a) & m8, the class's share (cls << l) >> r.
struct DigitSel {
    int a, l, r;
    unsigned m8;
};
__device__ __forceinline__ DigitSel digit_sel(int pass, const KeyPlan &pl) {
    const int w = pl.shift + 8 * pass;  // first composite bit of the digit
    DigitSel s;
    s.a = pl.lo + (w < pl.nb ? w : 0);
    s.m8 = w < pl.nb ? (unsigned)(((~0ull >> (64 - pl.nb)) >> w) & 0xffull) : 0u;
    s.l = w < pl.nb ? pl.nb - w : 0;
    s.r = w < pl.nb ? 0 : w - pl.nb;
    return s;
}
__device__ __forceinline__ unsigned digit_by(uint64_t key, uint64_t val, const DigitSel &s) {
    const unsigned cls = (unsigned)(val >> 56) & 31u;
    return (((unsigned)(key >> s.a) & s.m8) | ((cls << s.l) >> s.r)) & 0xffu;
}
__device__ __forceinline__ int64_t count_of(int64_t cap, const int64_t *dev) {
    if (dev == nullptr) return cap;
    const int64_t m = *dev;
    return m < cap ? m : cap;
}

__global__ void __launch_bounds__(256) zero_fill(uint4 *__restrict__ p, size_t n16) {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) p[i] = z;
}

// which key bits vary
__global__ void __launch_bounds__(256)
key_range(const uint64_t *__restrict__ key, int64_t Mcap, const int64_t *__restrict__ Mdev,
          SortCtl *__restrict__ ctl) {
    const int64_t M = count_of(Mcap, Mdev);
    uint64_t o = 0, n = 0;
    // four independent loads per thread and round: the kernel is latency-bound otherwise
    const int64_t stride = (int64_t)gridDim.x * 256;
    int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    for (; i + 3 * stride < M; i += 4 * stride) {
        const uint64_t k0 = __ldg(key + i), k1 = __ldg(key + i + stride), k2 = __ldg(key + i + 2 * stride),
                       k3 = __ldg(key + i + 3 * stride);
        o |= (k0 | k1) | (k2 | k3);
        n |= ~((k0 & k1) & (k2 & k3));
    }
    for (; i < M; i += stride) {
        const uint64_t k = __ldg(key + i);
        o |= k;
        n |= ~k;
    }
    // one pair of global atomics per CTA: they all hit the same two words and serialise in L2
    __shared__ unsigned s_or[4];
    if (threadIdx.x < 4) s_or[threadIdx.x] = 0;
    __syncthreads();
    const unsigned olo = __reduce_or_sync(0xffffffffu, (unsigned)o), ohi = __reduce_or_sync(0xffffffffu, (unsigned)(o >> 32));
    const unsigned nlo = __reduce_or_sync(0xffffffffu, (unsigned)n), nhi = __reduce_or_sync(0xffffffffu, (unsigned)(n >> 32));
    if ((threadIdx.x & 31) == 0) {
        atomicOr(&s_or[0], olo);
        atomicOr(&s_or[1], ohi);
        atomicOr(&s_or[2], nlo);
        atomicOr(&s_or[3], nhi);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicOr(&ctl->key_or, ((unsigned long long)s_or[1] << 32) | s_or[0]);
        atomicOr(&ctl->key_nand, ((unsigned long long)s_or[3] << 32) | s_or[2]);
    }
}

// Digit histograms of `npasses` digits (and the class histogram) in one read of the records:
// a CTA counts in shared memory and adds what it found to hist[][] / class_hist[].
__device__ __forceinline__ void histogram_records(const uint64_t *key, const uint64_t *val, int64_t M,
                                                  const KeyPlan &pl, unsigned (*s_h)[256], unsigned *s_c,
                                                  unsigned *hist, unsigned *class_hist) {
    for (int i = threadIdx.x; i < kPasses * 256; i += blockDim.x) (&s_h[0][0])[i] = 0;
    if (class_hist != nullptr && threadIdx.x < 32) s_c[threadIdx.x] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < M; i0 += 4 * stride) {
        uint64_t k[4], v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {  // all eight loads in flight before the first atomic
            const int64_t i = i0 + u * stride;
            k[u] = i < M ? __ldcg(key + i) : 0;
            v[u] = i < M ? __ldcg(val + i) : 0;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (i0 + u * stride >= M) break;
            if (pl.nb <= kCompositeMaxBits) {
                const uint64_t comp = composite_of(k[u], v[u], pl) >> pl.shift;
#pragma unroll
                for (int p = 0; p < kPasses; ++p)
                    if (p < pl.npasses) atomicAdd(&s_h[p][(unsigned)(comp >> (8 * p)) & 0xffu], 1u);
            } else {
                for (int p = 0; p < pl.npasses; ++p) atomicAdd(&s_h[p][digit_of(k[u], v[u], p, pl)], 1u);
            }
            if (class_hist != nullptr) atomicAdd(&s_c[(unsigned)(v[u] >> 56) & 31u], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < pl.npasses * 256; i += blockDim.x) {
        const unsigned c = (&s_h[0][0])[i];
        if (c) atomicAdd(hist + i, c);
    }
    if (class_hist != nullptr && threadIdx.x < 32 && s_c[threadIdx.x]) atomicAdd(&class_hist[threadIdx.x], s_c[threadIdx.x]);
}

__global__ void __launch_bounds__(256)
digit_histograms(const uint64_t *__restrict__ key, const uint64_t *__restrict__ val, int64_t Mcap,
                 const int64_t *__restrict__ Mdev, SortCtl *__restrict__ ctl) {
    __shared__ unsigned s_h[kPasses][256];
    __shared__ unsigned s_c[32];
    const KeyPlan pl = key_plan(ctl);
    histogram_records(key, val, count_of(Mcap, Mdev), pl, s_h, s_c, &ctl->hist[0][0], ctl->class_start);
}

// hist[p][d] -> number of records whose digit is < d, by one CTA of 256 threads
__device__ __forceinline__ void scan_digit_row(unsigned *row, unsigned *s_w /* [8] */) {
    const int d = threadIdx.x, lane = d & 31, wp = d >> 5;
    const unsigned v = __ldcg(row + d);
    unsigned inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_w[wp] = inc;
    __syncthreads();
    unsigned base = 0;
    for (int q = 0; q < wp; ++q) base += s_w[q];
    row[d] = base + inc - v;
    __syncthreads();
}
// one CTA per pass; the last CTA turns the class histogram into class_start[0..32]
__global__ void __launch_bounds__(256) digit_offsets(SortCtl *__restrict__ ctl) {
    __shared__ unsigned s_w[8];
    const int p = blockIdx.x, d = threadIdx.x, lane = d & 31;
    if (p < kPasses) {
        scan_digit_row(ctl->hist[p], s_w);
    } else if (d < 32) {
        const unsigned v = ctl->class_start[d];
        unsigned inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        ctl->class_start[d] = inc - v;
        if (d == 31) ctl->class_start[32] = inc;
    }
}

struct SortSmem {
    uint64_t x_key[kSortTile], x_val[kSortTile];  // 32 KB exchange
    unsigned hist[kSortWarps][256];
    unsigned gbase[256];   // global position of the first record of digit d of this tile
    unsigned lstart[256];  // tile-local position of the first record of digit d
    unsigned wsum[kSortWarps];
    unsigned tile;
};

// One pass of the stable LSD sort.  A tile (2,048 records) is ranked in registers,
// exchanged through shared memory so that records of one digit leave the CTA as one
// contiguous run (coalesced stores), and chained to its predecessors with a
// decoupled look-back.
// Persistent CTAs: the grid is sized for the SMs, not for the record capacity (the count is only
// known on the device, and a CTA per possible tile means a tail of empty CTAs in every pass).
// Tiles are drawn in order from the pass's counter, so a tile's predecessors are always held by
// CTAs that are running or done -- what the look-back needs.
// kCoherent: the input was written earlier in this same launch (radix_fallback): read it past L1.
template <bool kCoherent, bool kComposite>
__device__ __forceinline__ void radix_pass_tiles(const uint64_t *kin, const uint64_t *vin, uint64_t *kout,
                                                 uint64_t *vout, int64_t M, int pass, const KeyPlan &pl,
                                                 unsigned *status, SortCtl *ctl, SortSmem &sm) {
    // radix_fallback (the coherent reader) keeps its own counters and histograms
    unsigned *tile_counter = kCoherent ? &ctl->tile_counter2[pass] : &ctl->tile_counter[pass];
    const unsigned *hist_row = kCoherent ? ctl->hist2[pass] : ctl->hist[pass];
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const DigitSel sel = digit_sel(pass, pl);
  for (;;) {
    __syncthreads();  // everybody is done with the previous tile's shared memory
    if (tid == 0) sm.tile = atomicAdd(tile_counter, 1u);
    for (int i = tid; i < kSortWarps * 256; i += kSortThreads) (&sm.hist[0][0])[i] = 0;
    __syncthreads();
    const unsigned tile = sm.tile;
    const int64_t tile0 = (int64_t)tile * kSortTile;
    if (tile0 >= M) return;
    const int64_t first = tile0 + wp * kWarpSpan + lane;

    uint64_t k[kSortItems], v[kSortItems];
    unsigned dg[kSortItems], rk[kSortItems];
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int64_t idx = first + 32 * i;
        const bool ok = idx < M;
        if (kCoherent) {
            k[i] = ok ? __ldcg(kin + idx) : 0;
            v[i] = ok ? __ldcg(vin + idx) : 0;
        } else {
            k[i] = ok ? kin[idx] : 0;
            v[i] = ok ? vin[idx] : 0;
        }
    }
    // stable rank of every record among the records of its warp with the same digit
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const bool ok = (first + 32 * i) < M;
        dg[i] = kComposite ? digit_by(k[i], v[i], sel) : digit_of(k[i], v[i], pass, pl);
        const unsigned act = __ballot_sync(0xffffffffu, ok);
        unsigned peers = 0, pre = 0;
#ifndef YH_SORT_MATCH
        // peers from eight ballots (one per digit bit): measured 191 us for the four passes against
        // 205 us with one MATCH.ANY per record (-DYH_SORT_MATCH), whose pipe is the slower one
        peers = act;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool bit = (dg[i] >> b) & 1u;
            const unsigned mb = __ballot_sync(0xffffffffu, bit);
            peers &= bit ? mb : ~mb;
        }
        if (ok) pre = sm.hist[wp][dg[i]];
#else
        if (ok) {
            peers = __match_any_sync(act, dg[i]);
            pre = sm.hist[wp][dg[i]];
        }
#endif
        __syncwarp();
        if (ok) {
            rk[i] = pre + __popc(peers & lt);
            if ((peers & lt) == 0) sm.hist[wp][dg[i]] = pre + __popc(peers);
        }
        __syncwarp();
    }
    __syncthreads();
    volatile unsigned *st = status;
    unsigned tot = 0;
    {   // thread d owns digit d: warp offsets inside the tile, the tile total, the local layout
        const int d = tid;
#pragma unroll
        for (int q = 0; q < kSortWarps; ++q) {
            const unsigned c = sm.hist[q][d];
            sm.hist[q][d] = tot;
            tot += c;
        }
        // successors can already add this tile's counts while it is still looking back itself
        st[(size_t)tile * 256 + d] = (tile == 0 ? kFlagIncl : kFlagAgg) | tot;
        // exclusive scan of tot over the 256 digits -> tile-local run starts
        unsigned inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) sm.wsum[wp] = inc;
        __syncthreads();
        unsigned wbase = 0;
#pragma unroll
        for (int q = 0; q < kSortWarps; ++q) wbase += (q < wp) ? sm.wsum[q] : 0u;
        sm.lstart[d] = wbase + inc - tot;
    }
    __syncthreads();
    // exchange: records of one digit become contiguous inside the tile
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        if ((first + 32 * i) < M) {
            const unsigned q = sm.lstart[dg[i]] + sm.hist[wp][dg[i]] + rk[i];
            sm.x_key[q] = k[i];
            sm.x_val[q] = v[i];
        }
    }
    // decoupled look-back, thread d for digit d: walk back over the tiles that have only
    // published their own counts until one with an inclusive prefix is met.  It runs after
    // the exchange was issued, so most predecessors have published by now.
    {
        const int d = tid;
        unsigned excl = 0;
        if (tile > 0) {
            // kLook status words are fetched per round trip, then consumed nearest first
            constexpr int kLook = YH_SORT_LOOK;
            int64_t t = (int64_t)tile - 1;
            unsigned spins = 0;
            bool done = false;
            while (!done) {
                unsigned sv[kLook];
#pragma unroll
                for (int j = 0; j < kLook; ++j) sv[j] = (t - j >= 0) ? st[(size_t)(t - j) * 256 + d] : (unsigned)kFlagIncl;
                int used = 0;
#pragma unroll
                for (int j = 0; j < kLook; ++j) {
                    if (done || used < j) continue;      // stopped at an earlier word of this batch
                    const unsigned flag = sv[j] >> 30;
                    if (flag == 0) continue;              // not published yet: reread from here
                    excl += sv[j] & kValMask;
                    used = j + 1;
                    done = flag == 2;
                }
                t -= used;
                if (!done && used < kLook) {
                    __nanosleep(20);
                    if (++spins > (1u << 22)) {
                        atomicExch(&ctl->err, 1u);
                        break;
                    }
                }
            }
            st[(size_t)tile * 256 + d] = kFlagIncl | (excl + tot);
        }
        sm.gbase[d] = __ldcg(hist_row + d) + excl;
    }
    __syncthreads();
    const int count = (int)((M - tile0) < kSortTile ? (M - tile0) : kSortTile);
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int q = tid + kSortThreads * i;
        if (q < count) {
            const uint64_t kk = sm.x_key[q], vv = sm.x_val[q];
            const unsigned d = kComposite ? digit_by(kk, vv, sel) : digit_of(kk, vv, pass, pl);
            const unsigned pos = sm.gbase[d] + (unsigned)q - sm.lstart[d];
            kout[pos] = kk;
            vout[pos] = vv;
        }
    }
  }
}

// Pass p of the short plan reads buffer (p & 1) and writes the other one.
__global__ void __launch_bounds__(kSortThreads, YH_SORT_MINBLOCKS)
radix_pass(uint64_t *__restrict__ key_a, uint64_t *__restrict__ val_a, uint64_t *__restrict__ key_b,
           uint64_t *__restrict__ val_b, int64_t Mcap, const int64_t *__restrict__ Mdev, int pass,
           SortCtl *__restrict__ ctl, unsigned *__restrict__ status) {
    __shared__ SortSmem sm;
    const KeyPlan pl = key_plan(ctl);
    if (pass >= pl.npasses) return;
    const int64_t M = count_of(Mcap, Mdev);
    const uint64_t *kin = (pass & 1) ? key_b : key_a, *vin = (pass & 1) ? val_b : val_a;
    uint64_t *kout = (pass & 1) ? key_a : key_b, *vout = (pass & 1) ? val_a : val_b;
    radix_pass_tiles<false, true>(kin, vin, kout, vout, M, pass, pl, status, ctl, sm);
}

// ---- what the short plan leaves: runs of records that agree in class and leading key bits ----
// All keys agree outside [lo, lo+nb), so comparing whole keys orders the remaining bits, and two
// records agree in the composite's leading 32 bits iff they agree in class and in key >> (lo + shift).
// find_runs: a record smaller than its predecessor inside such a run marks the run as out of order;
// the thread at the FIRST such record of a run (no earlier one on the way back to the run's start)
// queues the run's start.  Read-only, so no thread sees a half-ordered run.
__global__ void __launch_bounds__(256)
find_runs(const uint64_t *__restrict__ key, const uint64_t *__restrict__ val, int64_t Mcap,
          const int64_t *__restrict__ Mdev, SortCtl *__restrict__ ctl, unsigned *__restrict__ work,
          unsigned work_cap) {
    const KeyPlan pl = key_plan(ctl);
    if (pl.shift == 0 || pl.npasses == 0) return;
    const int64_t M = count_of(Mcap, Mdev);
    const int top = pl.lo + pl.shift;
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x + 1; i < M; i += stride) {
        const uint64_t k = __ldg(key + i), kp = __ldg(key + i - 1);
        if ((k >> top) != (kp >> top) || !(kp > k)) continue;
        const unsigned cls = (unsigned)(__ldg(val + i) >> 56);
        if ((unsigned)(__ldg(val + i - 1) >> 56) != cls) continue;
        // walk back to the start of the run
        int64_t j = i - 1;
        uint64_t kj = kp;
        bool mine = true;
        for (;;) {
            if (j == 0) break;
            if (i - j >= kRunMax) {  // too long for fix_runs
                atomicExch(&ctl->overflow, 1u);
                mine = false;
                break;
            }
            const uint64_t kq = __ldg(key + j - 1);
            if ((kq >> top) != (k >> top) || (unsigned)(__ldg(val + j - 1) >> 56) != cls) break;
            if (kq > kj) {  // an earlier record of the run is out of order: its thread queues the run
                mine = false;
                break;
            }
            kj = kq;
            --j;
        }
        if (!mine) continue;
        const unsigned at = atomicAdd(&ctl->work_count, 1u);
        if (at < work_cap) work[at] = (unsigned)j;
        else atomicExch(&ctl->overflow, 1u);
    }
}

// fix_runs: one thread per queued run -- stable insertion sort of at most kRunMax records by key.
__global__ void __launch_bounds__(128)
fix_runs(uint64_t *__restrict__ key, uint64_t *__restrict__ val, int64_t Mcap, const int64_t *__restrict__ Mdev,
         SortCtl *__restrict__ ctl, const unsigned *__restrict__ work, unsigned work_cap) {
    const KeyPlan pl = key_plan(ctl);
    if (pl.shift == 0 || pl.npasses == 0) return;
    const int64_t M = count_of(Mcap, Mdev);
    const int top = pl.lo + pl.shift;
    unsigned n = ctl->work_count;
    if (n > work_cap) n = work_cap;
    for (unsigned w = blockIdx.x * 128 + threadIdx.x; w < n; w += gridDim.x * 128) {
        const int64_t s = work[w];
        uint64_t rk[kRunMax], rv[kRunMax];
        rk[0] = key[s];
        rv[0] = val[s];
        const unsigned cls = (unsigned)(rv[0] >> 56);
        int len = 1;
        bool fits = true;
        while (s + len < M) {
            const uint64_t k = key[s + len], v = val[s + len];
            if ((k >> top) != (rk[0] >> top) || (unsigned)(v >> 56) != cls) break;
            if (len == kRunMax) {
                fits = false;
                break;
            }
            // insert behind every record that is not greater (stable)
            int q = len;
            while (q > 0 && rk[q - 1] > k) {
                rk[q] = rk[q - 1];
                rv[q] = rv[q - 1];
                --q;
            }
            rk[q] = k;
            rv[q] = v;
            ++len;
        }
        if (!fits) {
            atomicExch(&ctl->overflow, 1u);
            continue;
        }
        for (int q = 0; q < len; ++q) {
            key[s + q] = rk[q];
            val[s + q] = rv[q];
        }
    }
}

// ---- fallback: the full plan, all passes in one launch ----
// Co-resident persistent CTAs separated by a grid barrier (a monotonic counter).  Returns at once
// unless the short plan raised `overflow` or does not apply (more than 59 varying key bits).
__device__ __forceinline__ void grid_barrier(unsigned *bar, unsigned &target, unsigned *err) {
    __syncthreads();
    if (threadIdx.x == 0) {
        target += gridDim.x;
        __threadfence();
        atomicAdd(bar, 1u);
        unsigned seen, spins = 0;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(bar) : "memory");
            if (seen < target) __nanosleep(100);
        } while (seen < target && ++spins < (1u << 26));
        if (seen < target) atomicExch(err, 1u);  // gave up after ~7 s: the result is flagged invalid
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kSortThreads, kFallbackCtasPerSm)
radix_fallback(uint64_t *key_a, uint64_t *val_a, uint64_t *key_b, uint64_t *val_b, int64_t Mcap,
               const int64_t *__restrict__ Mdev, SortCtl *ctl, unsigned *status, int ntiles_cap) {
    __shared__ SortSmem sm;
    __shared__ unsigned s_w[8];
    KeyPlan pl = key_plan(ctl);
    if (!plan_needs_fallback(pl, ctl)) return;
    pl.shift = 0;
    pl.npasses = pl.full;
    const int64_t M = count_of(Mcap, Mdev);
    unsigned target = 0;
    // the status words the short passes used; tile_counter2 / hist2 are still zero from the call's start
    {
        const size_t n = (size_t)pl.full * ntiles_cap * 256;
        for (size_t i = (size_t)blockIdx.x * kSortThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kSortThreads)
            status[i] = 0;
    }
    // digit histograms of the full plan (counted in the exchange buffer, which is idle here)
    histogram_records(key_a, val_a, M, pl, reinterpret_cast<unsigned (*)[256]>(sm.x_key), sm.wsum,
                      &ctl->hist2[0][0], nullptr);
    grid_barrier(&ctl->barrier, target, &ctl->err);
    for (int p = blockIdx.x; p < pl.full; p += gridDim.x) scan_digit_row(ctl->hist2[p], s_w);
    grid_barrier(&ctl->barrier, target, &ctl->err);
    for (int p = 0; p < pl.full; ++p) {
        const uint64_t *kin = (p & 1) ? key_b : key_a, *vin = (p & 1) ? val_b : val_a;
        uint64_t *kout = (p & 1) ? key_a : key_b, *vout = (p & 1) ? val_a : val_b;
        radix_pass_tiles<true, false>(kin, vin, kout, vout, M, p, pl, status + (size_t)p * ntiles_cap * 256, ctl, sm);
        grid_barrier(&ctl->barrier, target, &ctl->err);
    }
}

}  // namespace yh

#include "ap_emit.cuh"

namespace yh {

struct ApLayout {
    size_t ctl, apc, status, zero_end, tile_bits, work, key_alt, val_alt, total;
    int ntiles;
    unsigned work_cap;
};

// CTAs of radix_fallback that are resident at once on the current device (cached per device)
static int fallback_resident_ctas() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 1;
    if (cached[dev] == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radix_fallback, kSortThreads, 0);
        if (per_sm > kFallbackCtasPerSm) per_sm = kFallbackCtasPerSm;
        cached[dev] = (sms > 0 && per_sm > 0) ? sms * per_sm : 1;
    }
    return cached[dev];
}

static inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

static ApLayout ap_layout(int64_t M, int C, int T) {
    ApLayout L;
    L.ntiles = (int)((M + kSortTile - 1) / kSortTile);
    if (L.ntiles < 1) L.ntiles = 1;
    size_t o = 0;
    L.ctl = o; o = up256(o + sizeof(SortCtl));
    L.apc = o; o = up256(o + sizeof(ApCtl) + sizeof(unsigned long long) * (size_t)4 * C * T * kLevels);
    // look-back status words, one row per pass; only the short plan's rows are zeroed up front
    // (radix_fallback clears the rows it uses itself)
    L.status = o;
    L.zero_end = up256(o + sizeof(unsigned) * (size_t)kShortPasses * L.ntiles * 256);
    o = up256(o + sizeof(unsigned) * (size_t)kPasses * L.ntiles * 256);
    L.tile_bits = o; o = up256(o + sizeof(unsigned) * (size_t)L.ntiles * kMaxBits);
    L.work_cap = (unsigned)(M / 8 > 4096 ? M / 8 : 4096);  // out-of-order runs the short plan may leave
    L.work = o; o = up256(o + sizeof(unsigned) * (size_t)L.work_cap);
    L.key_alt = o; o = up256(o + sizeof(uint64_t) * (size_t)(M > 0 ? M : 1));
    L.val_alt = o; o = up256(o + sizeof(uint64_t) * (size_t)(M > 0 ? M : 1));
    L.total = o;
    return L;
}

}  // namespace yh

extern "C" size_t yh_ap_workspace_bytes(int64_t M, int C, int T) {
    if (M < 0 || C < 1 || C > 32 || T < 1 || T > YH_MAX_THRESHOLDS) return 0;
    return ap_layout(M, C, T).total;
}

extern "C" int yh_ap_reduce(uint64_t *key, uint64_t *flags, int64_t M, const int64_t *M_dev,
                            const int32_t *gt_count, int C, int T, double *ap, int64_t *totals,
                            double *packed, void *workspace, size_t workspace_bytes, void *stream) {
    if (C < 1 || C > 32 || T < 1 || T > YH_MAX_THRESHOLDS) {
        set_error("yh_ap_reduce: C=%d / T=%d outside what this build carries (C<=32, T<=%d)", C, T,
                  YH_MAX_THRESHOLDS);
        return YH_ERR_UNSUPPORTED;
    }
    if (M < 0 || M >= (1ll << 30) || (M > 0 && (!key || !flags)) || !gt_count || !ap || !totals || !workspace) {
        set_error("yh_ap_reduce: bad argument (M=%lld)", (long long)M);
        return YH_ERR_BAD_ARG;
    }
    const ApLayout L = ap_layout(M, C, T);
    if (workspace_bytes < L.total) {
        set_error("yh_ap_reduce: workspace too small (%zu < %zu)", workspace_bytes, L.total);
        return YH_ERR_WORKSPACE;
    }
    if ((uintptr_t)workspace % 256) {
        set_error("yh_ap_reduce: workspace must be 256-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char *ws = (char *)workspace;
    SortCtl *ctl = (SortCtl *)(ws + L.ctl);
    ApCtl *apc = (ApCtl *)(ws + L.apc);
    unsigned *status = (unsigned *)(ws + L.status);
    unsigned *tile_bits = (unsigned *)(ws + L.tile_bits);
    uint64_t *kb = (uint64_t *)(ws + L.key_alt), *vb = (uint64_t *)(ws + L.val_alt);
    // zeroed by a kernel of ours rather than a memset node: inside a captured graph a memset does not
    // inherit the stream's priority and would queue behind a concurrent lower-priority kernel
    {
        const size_t n16 = L.zero_end / 16;  // the zeroed prefix is a multiple of 256 bytes
        int zb = (int)((n16 + 256 * 8 - 1) / (256 * 8));
        if (zb > 148 * 8) zb = 148 * 8;
        zero_fill<<<zb < 1 ? 1 : zb, 256, 0, st>>>(reinterpret_cast<uint4 *>(ws), n16);
    }
    int launched = 2;
    if (M > 0) {
        int hb = (int)((M + 256 * 16 - 1) / (256 * 16));
        if (hb > 148 * YH_HIST_CTAS_PER_SM) hb = 148 * YH_HIST_CTAS_PER_SM;  // every CTA ends with global atomics
        key_range<<<hb, 256, 0, st>>>(key, M, M_dev, ctl);
        digit_histograms<<<hb, 256, 0, st>>>(key, flags, M, M_dev, ctl);
        digit_offsets<<<kPasses + 1, 256, 0, st>>>(ctl);
        // short plan: four passes over the composite's leading 32 bits (passes beyond the plan's count
        // return at once: the count is only known on the device), then the runs they leave;
        // radix_fallback returns at once unless the short plan gave up
        const int sort_grid = L.ntiles < 148 * YH_SORT_MINBLOCKS ? L.ntiles : 148 * YH_SORT_MINBLOCKS;
        for (int p = 0; p < kShortPasses; ++p)
            radix_pass<<<sort_grid, kSortThreads, 0, st>>>(key, flags, kb, vb, M, M_dev, p, ctl,
                                                           status + (size_t)p * L.ntiles * 256);
        unsigned *work = (unsigned *)(ws + L.work);
        int fb = (int)((M + 256 * 4 - 1) / (256 * 4));
        if (fb > 148 * 8) fb = 148 * 8;
        find_runs<<<fb, 256, 0, st>>>(key, flags, M, M_dev, ctl, work, L.work_cap);
        fix_runs<<<148, 128, 0, st>>>(key, flags, M, M_dev, ctl, work, L.work_cap);
        // the grid barrier needs every CTA resident: size the grid from what this device really holds
        const int fall_cap = fallback_resident_ctas();
        const int fall_grid = L.ntiles < fall_cap ? L.ntiles : fall_cap;
        radix_fallback<<<fall_grid, kSortThreads, 0, st>>>(key, flags, kb, vb, M, M_dev, ctl, status, L.ntiles);
        const int nbits = 4 * T + 1;
        tile_bit_counts<<<L.ntiles, kSortThreads, 0, st>>>(flags, vb, M, M_dev, ctl, nbits, tile_bits);
        scan_tile_bits<<<kMaxBits / kScanCols, 1024, 0, st>>>(tile_bits, M, M_dev, apc);
        class_start_bits<<<C, kSortThreads, 0, st>>>(flags, vb, M, M_dev, nbits, ctl, tile_bits, apc);
        emit_precisions<<<L.ntiles, kSortThreads, 0, st>>>(flags, vb, M, M_dev, C, T, ctl, tile_bits, gt_count, apc);
        launched += 3 + kShortPasses + 3 + 4;
    }
    finalize_ap<<<(4 * C * T + 127) / 128, 128, 0, st>>>(C, T, M, M_dev, ctl, gt_count, apc, ap, totals, packed);
    return check_launch("yh_ap_reduce", launched);
}
