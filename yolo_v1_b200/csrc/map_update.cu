// map_update.cu -- yh_map_update: decode + NMS + ground-truth parse + TP/FP matching
// fused into one pass over (pred, target), one warp per image; plus yh_parse_gt.
//
// Replaces mAPMetric.update (metrics.py:78-100) and the per-detection half of
// compute(): best-IoU ground truth and greedy TP/FP marking for every IoU
// threshold and every ground-truth subset (metrics.py:391-413, 465-486, 613-633).
// The reference does that matching after a global sort, but a detection only ever
// meets ground truths and rival detections of its own image and class, and their
// relative order in the global stable sort equals their per-image order, so the
// flags can be finalised here (verified against the
// reference's tp arrays in tests/golden/map_small.npz).
#include "head_nms.cuh"

namespace yh {

constexpr int kWarpsPerCtaM = 4;
constexpr int kCtasPerSmM = 7;

struct Thresholds {
    double t[YH_MAX_THRESHOLDS];
    int n;
};

// Overlay of `stage` once the target image has been parsed into registers.
struct MatchOverlay {
    double gx1[kCells], gy1[kCells], gx2[kCells], gy2[kCells], garea[kCells];  // 1,960 B, compact, cell order
    uint8_t gcls[kCells + 7], gbucket[kCells + 7];                              // 112 B
    unsigned long long flags[kCand];  // TP bits by sorted position                 784 B
    uint8_t bgt[kCand][4];            // by item: best ground truth per subset (255 = none)
    uint8_t ik[kPad], icls[kPad], ipos[kPad];  // by item: candidate, class, sorted position
};
static_assert(sizeof(MatchOverlay) <= sizeof(float2) * kHalfPairs, "overlay too large");

// Best IoU of a matched detection against the ground truths of one size bucket
// (1 small / 2 medium / 3 large), indexed by candidate k.  The three tables reuse
// the candidate's geometry slots (xa, ya, wh): a lane overwrites them only after it
// has finished with its own candidate's corners, and nobody else reads them.
__device__ __forceinline__ double *bucket_iou(WarpWindow &w, int bucket) {
    return bucket == 1 ? w.cand.xa : (bucket == 2 ? w.cand.ya : reinterpret_cast<double *>(w.cand.wh));
}

// Ground truths of one target image -> MatchOverlay (compact, in cell order).
// The target tensor is almost empty (an object in about 1 cell of 18), so it is not
// staged: tconf0/tconf1 are the slot-0 confidences of cells `lane` and `lane + 32`
// (loaded by the caller long before, 49 sectors per image), and only object cells
// fetch their box and class scores.  Returns the number of ground truths; cls_mask
// gets bit c for every class present.  gt_hist (shared, 4*kC ints) accumulates the
// per-(subset, class) counts of this CTA.
__device__ __forceinline__ int parse_targets(WarpWindow &w, const float *__restrict__ img, int lane,
                                             float tconf0, float tconf1, int *gt_hist, unsigned &cls_mask,
                                             uint8_t *__restrict__ out_cell, uint8_t *__restrict__ out_cls,
                                             float4 *__restrict__ out_raw) {
    const bool gt0 = tconf0 > 0.f, gt1 = (lane + 32 < kCells) && tconf1 > 0.f;  // metrics.py:239
    const unsigned m0 = __ballot_sync(kFull, gt0), m1 = __ballot_sync(kFull, gt1);
    const unsigned lt = (1u << lane) - 1u;
    MatchOverlay &mo = *reinterpret_cast<MatchOverlay *>(w.stage);
    const double small_a = __dmul_rn(__ddiv_rn(32.0, 448.0), __ddiv_rn(32.0, 448.0));   // metrics.py:511
    const double medium_a = __dmul_rn(__ddiv_rn(96.0, 448.0), __ddiv_rn(96.0, 448.0));  // metrics.py:512
    unsigned mine = 0;
    __syncwarp();  // every lane is done with what the overlay region held before
#pragma unroll 1
    for (int cs = 0; cs < 2; ++cs) {
        if (!(cs ? gt1 : gt0)) continue;
        const int cell = lane + 32 * cs;
        const float *c = img + cell * kD;
        const float2 xy = __ldg(reinterpret_cast<const float2 *>(c));
        const float2 wh = __ldg(reinterpret_cast<const float2 *>(c) + 1);
        const int cls = cell_class_global(c);
        const int g = cs ? __popc(m0) + __popc(m1 & lt) : __popc(m0 & lt);
        const double xa = abs_coord(cell % kS, xy.x), ya = abs_coord(cell / kS, xy.y);
        const Corners cr = corners_of(xa, ya, (double)wh.x, (double)wh.y);
        mo.gx1[g] = cr.x1; mo.gy1[g] = cr.y1; mo.gx2[g] = cr.x2; mo.gy2[g] = cr.y2;
        mo.garea[g] = cr.area;
        mo.gcls[g] = (uint8_t)cls;
        const int bucket = (cr.area < small_a) ? 1 : ((cr.area < medium_a) ? 2 : 3);
        mo.gbucket[g] = (uint8_t)bucket;
        if (out_raw != nullptr) {  // enough to rebuild the reference's tuple later (yh_list_form)
            out_cell[g] = (uint8_t)cell;
            out_cls[g] = (uint8_t)cls;
            out_raw[g] = make_float4(xy.x, xy.y, wh.x, wh.y);
        }
        atomicAdd(&gt_hist[cls], 1);
        atomicAdd(&gt_hist[bucket * kC + cls], 1);
        mine |= 1u << cls;
    }
    cls_mask = __reduce_or_sync(kFull, mine);
    __syncwarp();
    return __popc(m0) + __popc(m1);
}

__global__ void __launch_bounds__(kWarpsPerCtaM * 32, kCtasPerSmM)
map_update_kernel(const float *__restrict__ pred, const float *__restrict__ tgt, int N,
                  double conf_thr, double nms_thr, const __grid_constant__ Thresholds thr,
                  int32_t *__restrict__ n_keep, uint64_t *__restrict__ rec_key,
                  uint64_t *__restrict__ rec_flags, uint8_t *__restrict__ rec_k,
                  uint64_t *__restrict__ rec_box, int32_t *__restrict__ n_gt, uint8_t *__restrict__ gt_cell,
                  uint8_t *__restrict__ gt_cls, float4 *__restrict__ gt_raw,
                  int32_t *__restrict__ gt_count, unsigned *__restrict__ ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    WarpWindow *windows = reinterpret_cast<WarpWindow *>(smem_raw);
    __shared__ int gt_hist[4 * kC];
    __shared__ double s_thr[YH_MAX_THRESHOLDS];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    WarpWindow &w = windows[wib];
    const int warps_total = gridDim.x * kWarpsPerCtaM;
    const int T = thr.n;
    const unsigned lt = (1u << lane) - 1u;
    for (int i = threadIdx.x; i < 4 * kC; i += blockDim.x) gt_hist[i] = 0;
    if (threadIdx.x < YH_MAX_THRESHOLDS) s_thr[threadIdx.x] = thr.t[threadIdx.x];
    __syncthreads();

    // A warp's first image is fixed by its position; every further one is taken from a
    // global ticket counter, so warps that drew cheap images take more of them and the
    // grid drains evenly.  The ticket for the next round is drawn a whole image ahead.
    for (int img = blockIdx.x * kWarpsPerCtaM + wib; img < N;) {
        unsigned next_ticket = 0;
        if (lane == 0) next_ticket = atomicAdd(ticket, 1u);
        const float *pimg = pred + (size_t)img * kImgFloats, *timg = tgt + (size_t)img * kImgFloats;
        // slot-0 confidence of my two target cells: in flight while the predictions are decoded
        const float tconf0 = __ldg(timg + lane * kD + 4);
        const float tconf1 = __ldg(timg + (lane + 32 < kCells ? lane + 32 : lane) * kD + 4);
        // ---- predictions: decode, order, NMS
        const Mask128 valid = decode_image(w, pimg, lane, conf_thr, nullptr, timg, tconf0, tconf1);
        int n;
        const Mask128 keep = order_and_nms<YH_IOU_METRICS, false>(w, valid, nms_thr, lane, n);
        publish_keep(w, keep, lane);
        const bool lists = gt_raw != nullptr;
        if (rec_box != nullptr) {
            // the kept boxes as the reference stores them: (x_abs, y_abs) fp64, (w, h) fp32.
            // Written now: the matching below reuses the geometry slots of the table.
#pragma unroll 1
            for (int s = 0; s < kSlots; ++s) {
                if (!((w.keepw[s] >> lane) & 1u)) continue;
                const int k = w.order[lane + 32 * s];
                uint64_t *o = rec_box + 3 * ((size_t)img * kCand + w.keepbase[s] + __popc(w.keepw[s] & lt));
                o[0] = (uint64_t)__double_as_longlong(w.cand.xa[k]);
                o[1] = (uint64_t)__double_as_longlong(w.cand.ya[k]);
                o[2] = *reinterpret_cast<const uint64_t *>(&w.cand.wh[k]);
            }
        }

        // ---- targets: parse into the overlay
        unsigned cls_mask;
        const int ngt = parse_targets(w, timg, lane, tconf0, tconf1, gt_hist, cls_mask,
                                      lists ? gt_cell + (size_t)img * kCells : nullptr,
                                      lists ? gt_cls + (size_t)img * kCells : nullptr,
                                      lists ? gt_raw + (size_t)img * kCells : nullptr);
        MatchOverlay &mo = *reinterpret_cast<MatchOverlay *>(w.stage);
        if (lane == 0) {
            n_keep[img] = keep.count();
            n_gt[img] = ngt;
        }

        // ---- items: kept detections whose class has a ground truth, in sorted order
        int n_items = 0;
#pragma unroll 1
        for (int s = 0; s < kSlots; ++s) {
            const int p = lane + 32 * s;
            const bool kept = (w.keepw[s] >> lane) & 1u;
            int k = 0, c = 0;
            if (kept) {
                k = w.order[p];
                c = w.cand.cls[k];
                mo.flags[p] = 0;
            }
            const bool is_item = kept && ((cls_mask >> c) & 1u);
            const unsigned im = __ballot_sync(kFull, is_item);
            if (is_item) {
                const int it = n_items + __popc(im & lt);
                mo.ik[it] = (uint8_t)k;
                mo.icls[it] = (uint8_t)c;
                mo.ipos[it] = (uint8_t)p;
            }
            n_items += __popc(im);
        }
        __syncwarp();

        // ---- best ground truth of every item per size bucket; "all sizes" follows from the three
#pragma unroll 1
        for (int it = lane; it < n_items; it += 32) {
            const int k = mo.ik[it], c = mo.icls[it];
            const Corners dc = cand_corners(w, k);
            double b1 = 0.0, b2 = 0.0, b3 = 0.0;
            int g1 = 255, g2 = 255, g3 = 255;
            for (int g = 0; g < ngt; ++g) {
                if (mo.gcls[g] != c) continue;
                Corners gc;
                gc.x1 = mo.gx1[g]; gc.y1 = mo.gy1[g]; gc.x2 = mo.gx2[g]; gc.y2 = mo.gy2[g];
                gc.area = mo.garea[g];
                const double iou = iou_exact<YH_IOU_METRICS>(dc, gc);
                const int b = mo.gbucket[g];
                // strict: the first ground truth wins ties, IoU 0 never matches (metrics.py:393-403)
                if (b == 1 && iou > b1) { b1 = iou; g1 = g; }
                if (b == 2 && iou > b2) { b2 = iou; g2 = g; }
                if (b == 3 && iou > b3) { b3 = iou; g3 = g; }
            }
            // over all sizes: the largest of the three, the earliest ground truth on a tie
            double b0 = b1;
            int g0 = g1;
            if (b2 > b0 || (b2 == b0 && g2 < g0)) { b0 = b2; g0 = g2; }
            if (b3 > b0 || (b3 == b0 && g3 < g0)) { b0 = b3; g0 = g3; }
            w.cand.xa[k] = b1;  // bucket_iou(w, 1..3)[k]: this candidate's geometry is no longer needed
            w.cand.ya[k] = b2;
            reinterpret_cast<double *>(w.cand.wh)[k] = b3;
            *reinterpret_cast<unsigned *>(mo.bgt[it]) =
                (unsigned)g0 | ((unsigned)g1 << 8) | ((unsigned)g2 << 16) | ((unsigned)g3 << 24);
        }
        __syncwarp();

        // ---- greedy TP/FP: an item is a TP at threshold t iff its best IoU >= t and no
        // earlier detection of its class claimed the same ground truth with IoU >= t
        // (metrics.py:406-413).  "Earlier" = lower sorted position = lower item index.
        // One lane per (item, ground-truth subset): an image has few items (kept boxes whose
        // class has a ground truth), so a lane per item leaves most of the warp idle.
        const int n_work = 4 * n_items;
#pragma unroll 1
        for (int base = 0; base < n_work; base += 32) {
            const int wq = base + lane, q = lane & 3;
            const bool act = wq < n_work;
            const int it = act ? wq >> 2 : 0;
            uint64_t flags = 0;
            const unsigned g = act ? mo.bgt[it][q] : 255u;  // my best ground truth in subset q
            if (g != 255u) {
                const int k = mo.ik[it], c = mo.icls[it];
                // subset 0: the best over all sizes is g, so the IoU sits in g's bucket table
                const double *tab = bucket_iou(w, q ? q : mo.gbucket[g]);
                const double best = tab[k];
                double rival = 0.0;  // best IoU an earlier item brought to the same ground truth
                for (int e = it - 1; e >= 0 && mo.icls[e] == c; --e)
                    if (mo.bgt[e][q] == g) rival = fmax(rival, tab[mo.ik[e]]);
                unsigned bits = 0;
#pragma unroll 1
                for (int t = 0; t < T; ++t) {
                    const double th = s_thr[t];
                    if (best >= th && !(rival >= th)) bits |= 1u << t;
                }
                flags = (uint64_t)bits << (q * T);
                if (q == 0 && best >= 0.5 && !(rival >= 0.5)) flags |= 1ull << (4 * T);
            }
            // the four subsets of an item sit on four neighbouring lanes
            flags |= __shfl_xor_sync(kFull, flags, 1);
            flags |= __shfl_xor_sync(kFull, flags, 2);
            if (act && q == 0) mo.flags[mo.ipos[it]] = flags;
        }
        __syncwarp();

        // ---- records, in sorted order among the kept
#pragma unroll 1
        for (int s = 0; s < kSlots; ++s) {
            if (!((w.keepw[s] >> lane) & 1u)) continue;
            const int p = lane + 32 * s;
            const int k = w.order[p], c = w.cand.cls[k];
            const size_t o = (size_t)img * kCand + w.keepbase[s] + __popc(w.keepw[s] & lt);
            const double cf = w.cand.conf[k];
            rec_key[o] = ~orderable(cf == 0.0 ? 0.0 : cf);
            rec_flags[o] = mo.flags[p] | ((uint64_t)c << 56);
            rec_k[o] = (uint8_t)k;
        }
        asm volatile("" : "+r"(next_ticket)::"memory");  // keep the (slow) ticket read down here
        const unsigned drawn = __shfl_sync(kFull, next_ticket, 0);
        img = drawn < (unsigned)N ? warps_total + (int)drawn : N;
    }

    __syncthreads();
    for (int i = threadIdx.x; i < 4 * kC; i += blockDim.x)
        if (gt_hist[i]) atomicAdd(&gt_count[i], gt_hist[i]);
}

// Ground truths in list form (metrics.py:220-256), one warp per image.
__global__ void __launch_bounds__(128)
parse_gt_kernel(const float *__restrict__ tgt, int N, int32_t *__restrict__ n_gt,
                uint8_t *__restrict__ gt_cls, double *__restrict__ gt_box) {
    const int lane = threadIdx.x & 31;
    const int img = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (img >= N) return;
    const float *base = tgt + (size_t)img * kImgFloats;
    bool is_gt[2];
#pragma unroll
    for (int cs = 0; cs < 2; ++cs) {
        const int cell = lane + 32 * cs;
        is_gt[cs] = cell < kCells && __ldg(base + cell * kD + 4) > 0.f;
    }
    const unsigned m0 = __ballot_sync(kFull, is_gt[0]), m1 = __ballot_sync(kFull, is_gt[1]);
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll 1
    for (int cs = 0; cs < 2; ++cs)
        if (is_gt[cs]) {
            const int cell = lane + 32 * cs;
            const float *c = base + cell * kD;
            float cls_p[kC];
#pragma unroll
            for (int i = 0; i < kC; ++i) cls_p[i] = __ldg(c + kB * 5 + i);
            float pv;
            const int cls = argmax_first(cls_p, kC, pv);
            const int g = cs ? __popc(m0) + __popc(m1 & lt) : __popc(m0 & lt);
            const size_t o = (size_t)img * kCells + g;
            gt_cls[o] = (uint8_t)cls;
            gt_box[4 * o + 0] = abs_coord(cell % kS, __ldg(c + 0));
            gt_box[4 * o + 1] = abs_coord(cell / kS, __ldg(c + 1));
            gt_box[4 * o + 2] = (double)__ldg(c + 2);
            gt_box[4 * o + 3] = (double)__ldg(c + 3);
        }
    if (lane == 0) n_gt[img] = __popc(m0) + __popc(m1);
}

}  // namespace yh

using namespace yh;

extern "C" int yh_map_update(const float *pred, const float *tgt, int N, int S, int B, int C,
                             double conf_thr, double nms_thr, const double *thresholds_host, int T,
                             int32_t *n_keep, uint64_t *rec_key, uint64_t *rec_flags, uint8_t *rec_k,
                             void *rec_box, int32_t *n_gt, uint8_t *gt_cell, uint8_t *gt_cls,
                             float *gt_raw, int32_t *gt_count, void *stream) {
    if (!shape_ok(S, B, C)) {
        set_error("yh_map_update: only S=7,B=2,C=20 is built (got %d,%d,%d)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (T < 1 || T > YH_MAX_THRESHOLDS) {
        set_error("yh_map_update: T=%d thresholds; this build carries 1..%d", T, YH_MAX_THRESHOLDS);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !pred || !tgt || !thresholds_host || !n_keep || !rec_key || !rec_flags || !rec_k ||
        !n_gt || !gt_count || ((gt_raw != nullptr) != (gt_cell != nullptr)) ||
        ((gt_raw != nullptr) != (gt_cls != nullptr))) {
        set_error("yh_map_update: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (((uintptr_t)pred % 8) || ((uintptr_t)tgt % 8) || ((uintptr_t)rec_box % 8) || ((uintptr_t)gt_raw % 16)) {
        set_error("yh_map_update: pred/tgt/rec_box must be 8-byte aligned, gt_raw 16-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    if (N == 0) return 0;
    Thresholds th;
    th.n = T;
    for (int i = 0; i < YH_MAX_THRESHOLDS; ++i) th.t[i] = i < T ? thresholds_host[i] : 2.0;
    const size_t smem = sizeof(WarpWindow) * kWarpsPerCtaM;
    static std::atomic<bool> attr_done[kMaxDevices];
    if (first_use_on_device(attr_done)) {
        cudaFuncSetAttribute(map_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    int ctas = (N + kWarpsPerCtaM - 1) / kWarpsPerCtaM;
    const int cap = 148 * kCtasPerSmM;
    if (ctas > cap) ctas = cap;
    unsigned *ticket = fresh_ticket((cudaStream_t)stream);
    if (ticket == nullptr) return YH_ERR_UNSUPPORTED;
    map_update_kernel<<<ctas, kWarpsPerCtaM * 32, smem, (cudaStream_t)stream>>>(
        pred, tgt, N, conf_thr, nms_thr, th, n_keep, rec_key, rec_flags, rec_k, (uint64_t *)rec_box, n_gt, gt_cell,
        gt_cls, (float4 *)gt_raw, gt_count, ticket);
    return check_launch("yh_map_update");
}

extern "C" int yh_parse_gt(const float *tgt, int N, int S, int B, int C, int32_t *n_gt,
                           uint8_t *gt_cls, double *gt_box, void *stream) {
    if (!shape_ok(S, B, C)) {
        set_error("yh_parse_gt: only S=7,B=2,C=20 is built (got %d,%d,%d)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !tgt || !n_gt || !gt_cls || !gt_box) {
        set_error("yh_parse_gt: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (N == 0) return 0;
    parse_gt_kernel<<<(N + 3) / 4, 128, 0, (cudaStream_t)stream>>>(tgt, N, n_gt, gt_cls, gt_box);
    return check_launch("yh_parse_gt");
}
