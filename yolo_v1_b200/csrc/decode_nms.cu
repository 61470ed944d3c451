// decode_nms.cu -- yh_decode_nms: batched decode + per-class NMS, one warp per image.
//
// Replaces mAPMetric._parse_predictions + _apply_nms (metrics.py:173-218, 258-296)
// and YOLOInference.parse_predictions + non_max_suppression (inference.py:141-210,
// 251-317).  Building blocks in head_image.cuh / head_nms.cuh.
#include "head_nms.cuh"

namespace yh {

#ifndef YH_WARPS_PER_CTA
// 8 warps x 4 CTAs = 32 images in flight per SM (64 registers per thread; 4 x (8 x 7,088 B + static + 1 KB) of
// the SM's 228 KB).  4 x 7 = 28 warps measured 1.5-2 % slower (r02: 0.319 vs 0.314 ms for map_update).
#define YH_WARPS_PER_CTA 8
#define YH_CTAS_PER_SM 4
#endif
constexpr int kWarpsPerCta = YH_WARPS_PER_CTA;   // warps (= images in flight) per CTA
constexpr int kCtasPerSm = YH_CTAS_PER_SM;     // resident CTAs per SM the kernel is built for (registers, shared memory)

// Overlay of `stage` after NMS, for the global (confidence desc, candidate asc) output order: the kept
// detections are bucketed by the leading bits of their key and ranked inside their bucket only.
constexpr int kRankBuckets = 64;
struct GlobalOrderOverlay {
    uint64_t key[kCand];            // by bucket-grouped slot
    uint8_t k[kPad], bucket[kPad];  // candidate and bucket of that slot
    int hist[kRankBuckets];         // bucket sizes, then running cursor
    uint8_t start[kRankBuckets + 8];  // first slot of each bucket; [kRankBuckets] = number of kept
};
static_assert(sizeof(GlobalOrderOverlay) <= sizeof(float2) * kHalfPairs, "overlay too large");

// (x, y, w, h) of candidate k as the reference stores them (metrics.py:200-203, inference.py:184-193): the centre
// in the reference's doubles, rebuilt from the raw values (an optional output, off the throughput path)
static __device__ __noinline__ void emit_box(const float *__restrict__ img, int k, double *__restrict__ o) {
    const int cell = k >> 1;
    const float *b = img + cell * kD + (k & 1) * 5;
    o[0] = abs_coord(cell % kS, __ldg(b));
    o[1] = abs_coord(cell / kS, __ldg(b + 1));
    o[2] = (double)__ldg(b + 2);
    o[3] = (double)__ldg(b + 3);
}

// one kept detection -> output slot `pos` of image `img`
__device__ __forceinline__ void emit_kept(const WarpWindow &w, const float *__restrict__ pimg, int img, int pos, int k,
                                          uint8_t *__restrict__ keep_k, uint8_t *__restrict__ keep_cls,
                                          double *__restrict__ keep_conf, double *__restrict__ keep_box) {
    const size_t o = (size_t)img * kCand + pos;
    keep_k[o] = (uint8_t)k;
    if (keep_cls != nullptr) keep_cls[o] = w.cand.cls[k];
    if (keep_conf != nullptr) keep_conf[o] = from_orderable(~w.cand.key[k]);
    if (keep_box != nullptr) emit_box(pimg, k, keep_box + 4 * o);
}

// Global order of the kept detections (inference.py:302): position = number of kept detections with a smaller
// (key, candidate).  The keys' leading 32 bits, taken relative to the image's smallest and shifted so that the
// largest lands in bucket 63, are a monotone bucket number: a counting sort over the buckets places every
// detection up to the order inside its bucket, and that is settled by comparing it with its bucket mates only
// (about one per bucket instead of all ~60 kept).  Any key distribution is handled -- equal keys just share a bucket.
__device__ __forceinline__ void emit_global_order(WarpWindow &w, const float *__restrict__ pimg, int img,
                                                  const Mask128 &keep, int lane, uint8_t *__restrict__ keep_k,
                                                  uint8_t *__restrict__ keep_cls, double *__restrict__ keep_conf,
                                                  double *__restrict__ keep_box) {
    GlobalOrderOverlay &go = *reinterpret_cast<GlobalOrderOverlay *>(w.stage);
    __syncwarp();  // the NMS overlay is dead
    go.hist[lane] = 0;
    go.hist[lane + 32] = 0;
    unsigned lo = 0xffffffffu, hi = 0u;
    uint64_t key[kSlots];
    int kk[kSlots];
#pragma unroll
    for (int s = 0; s < kSlots; ++s) {
        kk[s] = -1;
        if ((keep.word(s) >> lane) & 1u) {
            kk[s] = w.order[lane + 32 * s];
            key[s] = w.cand.key[kk[s]];
            const unsigned h32 = (unsigned)(key[s] >> 32);
            lo = min(lo, h32);
            hi = max(hi, h32);
        }
    }
    lo = __reduce_min_sync(kFull, lo);
    hi = __reduce_max_sync(kFull, hi);
    const int sh = max(0, 26 - __clz((hi - lo) | 1u));  // (hi - lo) >> sh < 64
    __syncwarp();
    int arrival[kSlots], bkt[kSlots];
#pragma unroll
    for (int s = 0; s < kSlots; ++s)
        if (kk[s] >= 0) {
            bkt[s] = (int)(((unsigned)(key[s] >> 32) - lo) >> sh);
            arrival[s] = atomicAdd(&go.hist[bkt[s]], 1);
        }
    __syncwarp();
    // exclusive scan over the 64 buckets, two per lane
    const int c0 = go.hist[2 * lane], c1 = go.hist[2 * lane + 1];
    int inc = c0 + c1;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(kFull, inc, off);
        if (lane >= off) inc += v;
    }
    const int n_kept = __shfl_sync(kFull, inc, 31);
    go.start[2 * lane] = (uint8_t)(inc - c0 - c1);
    go.start[2 * lane + 1] = (uint8_t)(inc - c1);
    if (lane == 0) go.start[kRankBuckets] = (uint8_t)n_kept;
    __syncwarp();
#pragma unroll
    for (int s = 0; s < kSlots; ++s)
        if (kk[s] >= 0) {
            const int slot = go.start[bkt[s]] + arrival[s];
            go.key[slot] = key[s];
            go.k[slot] = (uint8_t)kk[s];
            go.bucket[slot] = (uint8_t)bkt[s];
        }
    __syncwarp();
#pragma unroll 1
    for (int q = lane; q < n_kept; q += 32) {
        const uint64_t mine = go.key[q];
        const int k = go.k[q], b = go.bucket[q];
        const int first = go.start[b], last = go.start[b + 1];
        int pos = first;
        for (int r = first; r < last; ++r) {
            const uint64_t other = go.key[r];
            pos += (other < mine) || (other == mine && go.k[r] < k);
        }
        emit_kept(w, pimg, img, pos, k, keep_k, keep_cls, keep_conf, keep_box);
    }
}

template <int MODE>
__global__ void __launch_bounds__(kWarpsPerCta * 32, kCtasPerSm)
decode_nms_kernel(const float *__restrict__ pred, int N, double conf_thr, double nms_thr, int order,
                  int32_t *__restrict__ n_keep, uint8_t *__restrict__ keep_k,
                  uint8_t *__restrict__ keep_cls, double *__restrict__ keep_conf,
                  double *__restrict__ keep_box, int32_t *__restrict__ n_cand,
                  uint8_t *__restrict__ bad_range, unsigned *__restrict__ ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    WarpWindow *windows = reinterpret_cast<WarpWindow *>(smem_raw);
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    WarpWindow &w = windows[wib];
    const int warps_total = gridDim.x * kWarpsPerCta;
    const FilterParams fp = filter_params(nms_thr);

    // first image by position, further ones from the ticket counter (see map_update.cu)
    for (int img = blockIdx.x * kWarpsPerCta + wib; img < N;) {
        unsigned next_ticket = 0;
        if (lane == 0) next_ticket = atomicAdd(ticket, 1u);
        const float *pimg = pred + (size_t)img * kImgFloats;
#ifdef YH_PREFETCH_AHEAD
        if (img + YH_PREFETCH_AHEAD * warps_total < N)   // see map_update.cu
            prefetch_image(pimg + (size_t)YH_PREFETCH_AHEAD * warps_total * kImgFloats, lane);
#endif
        bool bad = false;
        ImageFilter filt;
        const Mask128 valid = decode_image(w, pimg, lane, conf_thr, bad_range != nullptr ? &bad : nullptr, filt);
        filt.ok = filt.ok && fp.thr_ok;
        if (bad_range != nullptr) {
            const unsigned anybad = __ballot_sync(kFull, bad);
            if (lane == 0) bad_range[img] = anybad ? 1 : 0;
        }
        int n;
        const Mask128 keep = order_and_nms<MODE>(w, pimg, valid, nms_thr, fp, filt, lane, n);
        if (lane == 0) {
            n_keep[img] = keep.count();
            if (n_cand != nullptr) n_cand[img] = n;
        }
        if (order == YH_ORDER_GLOBAL_CONF) {
            emit_global_order(w, pimg, img, keep, lane, keep_k, keep_cls, keep_conf, keep_box);
        } else {
            class_group_offsets(w, keep, lane);
#pragma unroll 1
            for (int s = 0; s < kSlots; ++s) {
                const int p = lane + 32 * s;
                if (!((keep.word(s) >> lane) & 1u)) continue;
                const int k = w.order[p];
                const int pos = w.cls_off[w.cand.cls[k]] + count_below(keep, p) - w.cls_kbase[w.cand.cls[k]];
                emit_kept(w, pimg, img, pos, k, keep_k, keep_cls, keep_conf, keep_box);
            }
        }
        __syncwarp();
        asm volatile("" : "+r"(next_ticket)::"memory");  // keep the (slow) ticket read down here
        const unsigned drawn = __shfl_sync(kFull, next_ticket, 0);
        img = drawn < (unsigned)N ? warps_total + (int)drawn : N;
    }
}

template <int MODE>
static int launch(const float *pred, int N, double conf_thr, double nms_thr, int order, int32_t *n_keep,
                  uint8_t *keep_k, uint8_t *keep_cls, double *keep_conf, double *keep_box,
                  int32_t *n_cand, uint8_t *bad_range, cudaStream_t st) {
    const size_t smem = sizeof(WarpWindow) * kWarpsPerCta;
    static std::atomic<bool> attr_done[kMaxDevices];
    if (first_use_on_device(attr_done)) {
        cudaFuncSetAttribute(decode_nms_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    int ctas = (N + kWarpsPerCta - 1) / kWarpsPerCta;
    const int cap = 148 * kCtasPerSm;
    if (ctas > cap) ctas = cap;
    unsigned *ticket = fresh_ticket(st);
    if (ticket == nullptr) return YH_ERR_UNSUPPORTED;
    decode_nms_kernel<MODE><<<ctas, kWarpsPerCta * 32, smem, st>>>(
        pred, N, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls, keep_conf, keep_box, n_cand, bad_range,
        ticket);
    return check_launch("yh_decode_nms");
}

}  // namespace yh

using namespace yh;

extern "C" int yh_decode_nms(const float *pred, int N, int S, int B, int C, double conf_thr,
                             double nms_thr, int iou_mode, int order, int32_t *n_keep,
                             uint8_t *keep_k, uint8_t *keep_cls, double *keep_conf,
                             double *keep_box, int32_t *n_cand, uint8_t *bad_range, void *stream) {
    const bool fast = shape_ok(S, B, C);
    if (!fast && !generic_shape_ok(S, B, C, true)) {
        set_error("yh_decode_nms: shape S=%d,B=%d,C=%d outside this build (S*S*B <= 255, B <= 8, C <= 32)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !pred || !n_keep || !keep_k) {
        set_error("yh_decode_nms: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (!fast) {   // any other head shape: one thread per image (generic.cu)
        if (order != YH_ORDER_CLASS_GROUPED && order != YH_ORDER_GLOBAL_CONF) {
            set_error("yh_decode_nms: unknown order %d", order);
            return YH_ERR_BAD_ARG;
        }
        if (iou_mode != YH_IOU_METRICS && iou_mode != YH_IOU_INFER) {
            set_error("yh_decode_nms: unknown iou_mode %d", iou_mode);
            return YH_ERR_BAD_ARG;
        }
        if (N == 0) return 0;
        return generic_decode_nms(pred, N, S, B, C, conf_thr, nms_thr, iou_mode, order, n_keep, keep_k, keep_cls, keep_conf,
                                  keep_box, n_cand, bad_range, (cudaStream_t)stream);
    }
    if ((uintptr_t)pred % 8) {
        set_error("yh_decode_nms: pred must be 8-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    if (order != YH_ORDER_CLASS_GROUPED && order != YH_ORDER_GLOBAL_CONF) {
        set_error("yh_decode_nms: unknown order %d", order);
        return YH_ERR_BAD_ARG;
    }
    if (N == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (iou_mode == YH_IOU_METRICS)
        return launch<YH_IOU_METRICS>(pred, N, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls, keep_conf,
                                      keep_box, n_cand, bad_range, st);
    if (iou_mode == YH_IOU_INFER)
        return launch<YH_IOU_INFER>(pred, N, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls, keep_conf,
                                    keep_box, n_cand, bad_range, st);
    set_error("yh_decode_nms: unknown iou_mode %d", iou_mode);
    return YH_ERR_BAD_ARG;
}
