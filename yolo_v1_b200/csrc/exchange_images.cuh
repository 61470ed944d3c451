// exchange_images.cuh -- yh_exchange_images: compaction, partition by class owner and the all-to-all of one
// rank's detection records in ONE pass over the fixed-stride output of yh_map_update (part of partition.cu's
// translation unit: it shares the peer-block protocol -- peer_begin, ready / done flags, peer_finish).
//
// yh_compact_records + yh_exchange_records read and write every record three times on the way to its owner
// (dense copy; count by owner; ordered tile -> remote store).  But the records of an image leave yh_map_update
// sorted by class, and the owner of a class is fixed, so an image's records already ARE one contiguous run per
// owner.  All that is needed is, per image, how many records each owner gets (8 small counters), their prefix
// over the images, and one copy: source = the image's run, destination = the owner's receive buffer.
//
//   xi_count         one warp per image: owner of every record, eight 8-bit counters per image packed into one
//                    word; per-CTA (64 images) totals
//   xi_scan_publish  one CTA per batch: exclusive prefix of the CTA totals behind the batches before it; after
//                    the last batch the rank's run lengths, ground-truth counts and image count go to every rank
//                    and ready[me] is raised (as peer_scan_publish)
//   xi_scatter       a CTA waits for all ready[] flags and derives where this rank's runs start; a warp turns the
//                    counters of its 32 images into offsets and then copies, owner by owner, the records of those
//                    images to one contiguous destination range -- consecutive lanes on consecutive (remote)
//                    addresses, the source found by a 5-step search over the 32 offsets.  The last CTA of the last
//                    batch raises done[me] everywhere.
//   peer_finish      unchanged.
//
// Order in a receive buffer: source rank, then batch (update() call), then image, then the per-image order --
// the same sequence yh_exchange_records produces from the dense arrays, hence the same stable-sort tie order.
#pragma once

namespace yh {

constexpr int kXiOwners = 8;          // counters are 8 x 8 bits per image, 4 x 16 bits x 2 per CTA
constexpr int kXiThreads = 256;
constexpr int kXiIpw = 8;             // images per warp: few, so that many warps are in flight (the copy is latency-bound)
constexpr int kXiImages = kXiIpw * kXiThreads / 32;   // images per CTA

struct XiLocal {                      // head of the workspace (after ExchangeLocal), zeroed by xi_begin
    unsigned long long run_total[kXiOwners];   // records of the batches scanned so far, per owner
};

__global__ void xi_begin(ExchangeLocal *__restrict__ loc, XiLocal *__restrict__ xl, PeerCtl *__restrict__ mine) {
    if (threadIdx.x == 0) {
        loc->finished_ctas = 0;
        loc->timed_out = 0;
        loc->overflow = 0;
        mine->calls = mine->calls + 1;
    }
    if (threadIdx.x < kXiOwners) xl->run_total[threadIdx.x] = 0;
}

// bytes 0..3 of x -> four 16-bit fields
__device__ __forceinline__ unsigned long long spread_bytes(unsigned x) {
    unsigned long long v = x;
    v = (v | (v << 16)) & 0x0000ffff0000ffffull;
    v = (v | (v << 8)) & 0x00ff00ff00ff00ffull;
    return v;
}

__global__ void __launch_bounds__(kXiThreads)
xi_count(const int32_t *__restrict__ n_keep, const uint64_t *__restrict__ rec_flags, int N, int K,
         const __grid_constant__ OwnerTable tab, unsigned long long *__restrict__ img_cnt,
         unsigned *__restrict__ cta_tot) {
    __shared__ unsigned long long s_lo[kXiThreads / 32], s_hi[kXiThreads / 32];
    const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
    const int img0 = blockIdx.x * kXiImages + wp * kXiIpw;
    unsigned long long mine = 0;   // lane i < kXiIpw keeps the counters of image img0 + i
#pragma unroll 2
    for (int i = 0; i < kXiIpw; ++i) {
        const int img = img0 + i;
        if (img >= N) break;       // warp-uniform
        const int n = n_keep[img];
        const uint64_t *f = rec_flags + (size_t)img * K;
        unsigned lo = 0, hi = 0;   // four 8-bit counters each: at most 4 records per lane, 98 per image
        for (int q = lane; q < n; q += 32) {
            const int o = owner_of(tab, __ldg(f + q));
            if (o < 4) lo += 1u << (8 * o); else hi += 1u << (8 * (o - 4));
        }
        lo = __reduce_add_sync(0xffffffffu, lo);
        hi = __reduce_add_sync(0xffffffffu, hi);
        if (lane == i) mine = ((unsigned long long)hi << 32) | lo;
    }
    if (lane < kXiIpw && img0 + lane < N) img_cnt[img0 + lane] = mine;
    // CTA totals: 16-bit fields (64 images x 255 records fit)
    unsigned long long tl = spread_bytes((unsigned)mine), th = spread_bytes((unsigned)(mine >> 32));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        tl += __shfl_xor_sync(0xffffffffu, tl, o);
        th += __shfl_xor_sync(0xffffffffu, th, o);
    }
    if (lane == 0) {
        s_lo[wp] = tl;
        s_hi[wp] = th;
    }
    __syncthreads();
    if (threadIdx.x < kXiOwners) {
        unsigned t = 0;
        for (int q = 0; q < kXiThreads / 32; ++q) {
            const unsigned long long v = threadIdx.x < 4 ? s_lo[q] : s_hi[q];
            t += (unsigned)(v >> (16 * (threadIdx.x & 3))) & 0xffffu;
        }
        cta_tot[(size_t)blockIdx.x * kXiOwners + threadIdx.x] = t;
    }
}

// One CTA.  cta_tot[nblk][8] -> exclusive prefix over the batch's CTAs, behind the rank's earlier batches
// (run_total).  publish: this was the last batch -- send the rank's message and raise ready[me].
__global__ void __launch_bounds__(1024)
xi_scan_publish(unsigned *__restrict__ cta_tot, int nblk, XiLocal *__restrict__ xl, int publish,
                const __grid_constant__ PeerTable peers, int world, int me, unsigned epoch,
                const int32_t *__restrict__ gt_local, int n_gt, long long n_images) {
    __shared__ unsigned s_warp[32];
    __shared__ unsigned long long s_carry[kXiOwners];
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    if (tid < kXiOwners) s_carry[tid] = xl->run_total[tid];
    __syncthreads();
    for (int p = 0; p < world; ++p) {
        for (int base = 0; base < nblk; base += 1024) {
            const int i = base + tid;
            const unsigned c = i < nblk ? cta_tot[(size_t)i * kXiOwners + p] : 0;
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
            }
            __syncthreads();
            const unsigned long long carry = s_carry[p];
            // a rank's run in one receive buffer stays below 2^32 records (capacity is checked on the host)
            if (i < nblk) cta_tot[(size_t)i * kXiOwners + p] = (unsigned)(carry + s_warp[wp] + inc - c);
            __syncthreads();
            if (tid == 1023) s_carry[p] = carry + s_warp[wp] + inc;
            __syncthreads();
        }
    }
    if (tid < kXiOwners) xl->run_total[tid] = s_carry[tid];
    if (!publish) return;
    const int per = world + n_gt + 2;  // words of one rank's message
    for (int q = tid; q < world * per; q += blockDim.x) {
        const int d = q / per, i = q - d * per;
        PeerCtl *c = peers.ctl[d];
        if (i < world)
            c->row[me][i] = (long long)s_carry[i];
        else if (i < world + n_gt)
            c->gt[me][i - world] = gt_local[i - world];
        else
            c->gt[me][kGtWords - 2 + (i - world - n_gt)] = (int)((unsigned long long)n_images >> (32 * (i - world - n_gt)));
    }
    __threadfence_system();
    __syncthreads();
    if (tid < world) st_release_sys(&peers.ctl[tid]->ready[me], epoch_of(epoch, peers.ctl[me]));
}

__global__ void __launch_bounds__(kXiThreads)
xi_scatter(const int32_t *__restrict__ n_keep, const uint64_t *__restrict__ rec_key,
           const uint64_t *__restrict__ rec_flags, int N, int K, const unsigned long long *__restrict__ img_cnt,
           const unsigned *__restrict__ cta_base, const __grid_constant__ PeerTable peers, int world, int me,
           unsigned epoch, long long capacity, ExchangeLocal *__restrict__ loc, unsigned total_ctas) {
    __shared__ long long dst_base[kXiOwners];   // where this CTA's first record for owner r goes in r's buffer
    __shared__ unsigned long long s_lo[kXiThreads / 32], s_hi[kXiThreads / 32];
    __shared__ unsigned short s_off[kXiThreads / 32][32];
    __shared__ uint8_t s_first[kXiThreads / 32][32];
    __shared__ int s_state;
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const PeerCtl *mine = peers.ctl[me];
    epoch = epoch_of(epoch, mine);
    if (tid == 0) s_state = 0;
    __syncthreads();
    if (tid < world && !wait_flag(&mine->ready[tid], epoch, mine->wait_ms)) s_state = 1;
    __syncthreads();
    if (s_state) {
        if (tid == 0) loc->timed_out = 1;
        return;
    }
    if (tid < kXiOwners) {
        long long before = 0, total = 0;
        if (tid < world)
            for (int r = 0; r < world; ++r) {
                const long long c = __ldcv(&mine->row[r][tid]);  // written by a peer: never from L1
                if (r < me) before += c;
                total += c;
            }
        dst_base[tid] = before + cta_base[(size_t)blockIdx.x * kXiOwners + tid];
        if (total > capacity) s_state = 2;
    }
    __syncthreads();
    const bool overflow = s_state == 2;
    if (overflow && tid == 0 && blockIdx.x == 0) loc->overflow = 1;
    if (!overflow) {
        const int img = blockIdx.x * kXiImages + wp * kXiIpw + lane;   // lane < kXiIpw <-> one of the warp's images
        const unsigned long long c8 = (lane < kXiIpw && img < N) ? img_cnt[img] : 0ull;
        // offsets of my image's runs inside the CTA: inclusive scan of 16-bit fields over the lanes, then the warps
        unsigned long long il = spread_bytes((unsigned)c8), ih = spread_bytes((unsigned)(c8 >> 32));
        const unsigned long long ml = il, mh = ih;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long ul = __shfl_up_sync(0xffffffffu, il, o), uh = __shfl_up_sync(0xffffffffu, ih, o);
            if (lane >= o) {
                il += ul;
                ih += uh;
            }
        }
        if (lane == 31) {
            s_lo[wp] = il;
            s_hi[wp] = ih;
        }
        __syncthreads();
        unsigned long long wl = 0, wh = 0;   // records of the warps before mine
        for (int q = 0; q < wp; ++q) {
            wl += s_lo[q];
            wh += s_hi[q];
        }
        const unsigned long long exl = il - ml, exh = ih - mh;          // exclusive inside the warp
        // first record of every owner's run inside my image: byte-wise exclusive prefix of the counters
        const unsigned long long first8 = c8 * 0x0101010101010101ull - c8;
        const size_t src0 = (size_t)(blockIdx.x * kXiImages + wp * kXiIpw) * K;
#pragma unroll 1
        for (int r = 0; r < world; ++r) {
            const int sh = 16 * (r & 3);
            const unsigned ex = (unsigned)((r < 4 ? exl : exh) >> sh) & 0xffffu;
            const unsigned cnt = (unsigned)(c8 >> (8 * r)) & 0xffu;
            const unsigned wbase = (unsigned)((r < 4 ? wl : wh) >> sh) & 0xffffu;
            s_off[wp][lane] = (unsigned short)ex;
            s_first[wp][lane] = (uint8_t)(first8 >> (8 * r));
            const unsigned total = __shfl_sync(0xffffffffu, ex + cnt, 31);
            __syncwarp();
            uint64_t *dk = peers.key[r] + dst_base[r] + wbase;
            uint64_t *df = peers.flags[r] + dst_base[r] + wbase;
            // four records per lane and round: the loads of one round are all in flight before the first store
            for (unsigned j0 = lane; j0 < total; j0 += 128) {
                uint64_t kv[4], fv[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const unsigned j = j0 + 32 * u;
                    if (j >= total) break;
                    int i = 0;   // the last image whose offset is <= j (empty images share an offset with their successor)
#pragma unroll
                    for (int h = 16; h > 0; h >>= 1)
                        if (s_off[wp][i + h] <= j) i += h;
                    const size_t s = src0 + (size_t)i * K + s_first[wp][i] + (j - s_off[wp][i]);
                    kv[u] = __ldg(rec_key + s);
                    fv[u] = __ldg(rec_flags + s);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const unsigned j = j0 + 32 * u;
                    if (j >= total) break;
                    dk[j] = kv[u];
                    df[j] = fv[u];
                }
            }
            __syncwarp();
        }
    }
    // completion: my stores are visible system-wide before the count moves; the last CTA of the last batch
    // raises done[me]
    __threadfence_system();
    __syncthreads();
    if (tid == 0) s_state = (atomicAdd(&loc->finished_ctas, 1u) == total_ctas - 1) ? 3 : 0;
    __syncthreads();
    if (s_state == 3 && tid < world) {
        __threadfence_system();
        st_release_sys(&peers.ctl[tid]->done[me], epoch);
    }
}

}  // namespace yh

static inline size_t xi_up256(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" size_t yh_exchange_images_workspace_bytes(int64_t n_images_total, int n_chunks) {
    if (n_images_total < 0 || n_chunks < 0) return 0;
    const size_t blocks = (size_t)n_images_total / yh::kXiImages + (size_t)n_chunks + 1;
    return 256 + 256 + 256 + xi_up256(sizeof(unsigned long long) * (size_t)n_images_total) + 256 * (size_t)n_chunks +
           sizeof(unsigned) * blocks * yh::kXiOwners;
}

extern "C" int yh_exchange_images(int n_chunks, const int32_t *const *n_keep, const uint64_t *const *rec_key,
                                  const uint64_t *const *rec_flags, const int *n_images, int K,
                                  const uint8_t *owner_of_class_host, int C, int world, int rank,
                                  void *const *peer_block_host, int64_t capacity, unsigned epoch,
                                  const int32_t *gt_count_local, int32_t *gt_count_global, int64_t *info,
                                  void *workspace, size_t workspace_bytes, void *stream) {
    using namespace yh;
    if (C < 1 || C > 32 || world < 1 || world > kXiOwners) {
        set_error("yh_exchange_images: C=%d / world=%d outside this build (C<=32, world<=%d)", C, world, kXiOwners);
        return YH_ERR_UNSUPPORTED;
    }
    if (n_chunks < 0 || K < 1 || K > 255 || rank < 0 || rank >= world || capacity < 0 || capacity >= (1ll << 32) ||
        !owner_of_class_host || !peer_block_host || !gt_count_local || !gt_count_global || !info || !workspace ||
        (n_chunks > 0 && (!n_keep || !rec_key || !rec_flags || !n_images))) {
        set_error("yh_exchange_images: bad argument (chunks=%d K=%d rank=%d world=%d)", n_chunks, K, rank, world);
        return YH_ERR_BAD_ARG;
    }
    int64_t total_images = 0;
    unsigned total_ctas = 0;
    for (int c = 0; c < n_chunks; ++c) {
        if (n_images[c] < 0 || (n_images[c] > 0 && (!n_keep[c] || !rec_key[c] || !rec_flags[c]))) {
            set_error("yh_exchange_images: batch %d is malformed", c);
            return YH_ERR_BAD_ARG;
        }
        total_images += n_images[c];
        total_ctas += (unsigned)((n_images[c] + kXiImages - 1) / kXiImages);
    }
    if (workspace_bytes < yh_exchange_images_workspace_bytes(total_images, n_chunks)) {
        set_error("yh_exchange_images: workspace too small");
        return YH_ERR_WORKSPACE;
    }
    OwnerTable tab;
    for (int c = 0; c < 256; ++c) {
        const int o = c < C ? owner_of_class_host[c] : 0;
        if (o >= world) {
            set_error("yh_exchange_images: class %d is owned by rank %d >= world %d", c, o, world);
            return YH_ERR_BAD_ARG;
        }
        tab.of_class[c] = (uint8_t)o;
    }
    PeerTable peers;
    for (int r = 0; r < kMaxParts; ++r) {
        char *base = (char *)peer_block_host[r < world ? r : rank];
        if (!base) {
            set_error("yh_exchange_images: peer block %d is null", r);
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
    XiLocal *xl = (XiLocal *)(ws + 256);
    unsigned long long *img_cnt = (unsigned long long *)(ws + 512);
    unsigned *cta_tot = (unsigned *)(ws + 512 + xi_up256(sizeof(unsigned long long) * (size_t)total_images) +
                                     256 * (size_t)n_chunks);
    int64_t images_total_ll = total_images;
    int launches = 1;
    xi_begin<<<1, 32, 0, st>>>(loc, xl, peers.ctl[rank]);
    // pass 1: counters and prefixes, batch by batch; the last scan publishes
    size_t img_off = 0, blk_off = 0;
    int last = -1;
    for (int c = 0; c < n_chunks; ++c)
        if (n_images[c] > 0) last = c;
    if (last < 0) {   // nothing to send: publish zeros
        xi_scan_publish<<<1, 1024, 0, st>>>(cta_tot, 0, xl, 1, peers, world, rank, epoch, gt_count_local, 4 * C,
                                            (long long)images_total_ll);
        ++launches;
    }
    for (int c = 0; c < n_chunks; ++c) {
        if (n_images[c] == 0) continue;
        const int nblk = (n_images[c] + kXiImages - 1) / kXiImages;
        xi_count<<<nblk, kXiThreads, 0, st>>>(n_keep[c], rec_flags[c], n_images[c], K, tab, img_cnt + img_off,
                                              cta_tot + blk_off * kXiOwners);
        xi_scan_publish<<<1, 1024, 0, st>>>(cta_tot + blk_off * kXiOwners, nblk, xl, c == last ? 1 : 0, peers, world, rank,
                                            epoch, gt_count_local, 4 * C, (long long)images_total_ll);
        launches += 2;
        img_off += (size_t)n_images[c];
        blk_off += (size_t)nblk;
    }
    // pass 2: the copy
    img_off = 0;
    blk_off = 0;
    for (int c = 0; c < n_chunks; ++c) {
        if (n_images[c] == 0) continue;
        const int nblk = (n_images[c] + kXiImages - 1) / kXiImages;
        xi_scatter<<<nblk, kXiThreads, 0, st>>>(n_keep[c], rec_key[c], rec_flags[c], n_images[c], K, img_cnt + img_off,
                                                cta_tot + blk_off * kXiOwners, peers, world, rank, epoch,
                                                (long long)capacity, loc, total_ctas);
        ++launches;
        img_off += (size_t)n_images[c];
        blk_off += (size_t)nblk;
    }
    if (total_ctas == 0) {   // no scatter CTA exists to raise done[me]: a one-thread stand-in does
        xi_scatter<<<1, kXiThreads, 0, st>>>(nullptr, nullptr, nullptr, 0, K, img_cnt, cta_tot, peers, world, rank, epoch,
                                             (long long)capacity, loc, 1u);
        ++launches;
    }
    peer_finish<<<1, 256, 0, st>>>(peers, world, rank, epoch, 4 * C, loc, gt_count_global, info);
    return check_launch("yh_exchange_images", launches + 1);
}
