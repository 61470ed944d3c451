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

struct Thresholds {
    double t[YH_MAX_THRESHOLDS];
    int n;
};

// Ground truths of one target image -> MatchOverlay (compact, in cell order).
// Returns the number of ground truths.  gt_hist (shared, 4*kC ints) accumulates
// the per-(subset, class) counts of this CTA.
__device__ __forceinline__ int parse_targets(WarpWindow &w, int lane, int *gt_hist,
                                             uint8_t *__restrict__ out_cell, uint8_t *__restrict__ out_cls,
                                             float4 *__restrict__ out_raw) {
    const float *st = reinterpret_cast<const float *>(w.stage);
    // pass 1: pull the (few) ground-truth cells into registers before the overlay is written
    bool is_gt[2];
    int cls[2] = {0, 0};
    float x[2], y[2], bw[2], bh[2];
#pragma unroll
    for (int cs = 0; cs < 2; ++cs) {
        const int cell = lane + 32 * cs;
        const bool live = cell < kCells;
        const float *c = st + (live ? cell : 0) * kD;
        is_gt[cs] = live && (c[4] > 0.f);  // slot-0 confidence only (metrics.py:239)
        x[cs] = c[0]; y[cs] = c[1]; bw[cs] = c[2]; bh[cs] = c[3];
    }
#pragma unroll 1
    for (int cs = 0; cs < 2; ++cs) {
        if (!(cs == 0 ? is_gt[0] : is_gt[1])) continue;  // object cells are rare (about 1 in 18)
        const float *c = st + (lane + 32 * cs) * kD;
        float cls_p[kC];
#pragma unroll
        for (int i = 0; i < kC / 2; ++i) {
            float2 v = *reinterpret_cast<const float2 *>(c + kB * 5 + 2 * i);
            cls_p[2 * i] = v.x;
            cls_p[2 * i + 1] = v.y;
        }
        float pv;
        const int a = argmax_first(cls_p, kC, pv);
        if (cs == 0) cls[0] = a; else cls[1] = a;
    }
    const unsigned m0 = __ballot_sync(kFull, is_gt[0]), m1 = __ballot_sync(kFull, is_gt[1]);
    const unsigned lt = (1u << lane) - 1u;
    __syncwarp();  // all lanes hold their cells; `stage` may be overwritten now
    MatchOverlay &mo = *reinterpret_cast<MatchOverlay *>(w.stage);
    const double small_a = __dmul_rn(__ddiv_rn(32.0, 448.0), __ddiv_rn(32.0, 448.0));   // metrics.py:511
    const double medium_a = __dmul_rn(__ddiv_rn(96.0, 448.0), __ddiv_rn(96.0, 448.0));  // metrics.py:512
#pragma unroll
    for (int cs = 0; cs < 2; ++cs)
        if (is_gt[cs]) {
            const int g = cs ? __popc(m0) + __popc(m1 & lt) : __popc(m0 & lt);
            const int cell = lane + 32 * cs;
            const double xa = abs_coord(cell % kS, x[cs]), ya = abs_coord(cell / kS, y[cs]);
            const Corners cr = corners_of(xa, ya, (double)bw[cs], (double)bh[cs]);
            mo.gx1[g] = cr.x1; mo.gy1[g] = cr.y1; mo.gx2[g] = cr.x2; mo.gy2[g] = cr.y2;
            mo.garea[g] = cr.area;
            mo.gcls[g] = (uint8_t)cls[cs];
            const int bucket = (cr.area < small_a) ? 1 : ((cr.area < medium_a) ? 2 : 3);
            mo.gbucket[g] = (uint8_t)bucket;
            if (out_raw != nullptr) {  // enough to rebuild the reference's tuple later (yh_list_form)
                out_cell[g] = (uint8_t)cell;
                out_cls[g] = (uint8_t)cls[cs];
                out_raw[g] = make_float4(x[cs], y[cs], bw[cs], bh[cs]);
            }
            atomicAdd(&gt_hist[cls[cs]], 1);
            atomicAdd(&gt_hist[bucket * kC + cls[cs]], 1);
        }
    __syncwarp();
    return __popc(m0) + __popc(m1);
}

__global__ void __launch_bounds__(kWarpsPerCtaM * 32)
map_update_kernel(const float *__restrict__ pred, const float *__restrict__ tgt, int N,
                  double conf_thr, double nms_thr, Thresholds thr, int32_t *__restrict__ n_keep,
                  uint64_t *__restrict__ rec_key, uint64_t *__restrict__ rec_flags,
                  uint8_t *__restrict__ rec_k, float4 *__restrict__ rec_raw,
                  int32_t *__restrict__ n_gt, uint8_t *__restrict__ gt_cell,
                  uint8_t *__restrict__ gt_cls, float4 *__restrict__ gt_raw,
                  int32_t *__restrict__ gt_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    WarpWindow *windows = reinterpret_cast<WarpWindow *>(smem_raw);
    __shared__ int gt_hist[4 * kC];
    __shared__ double s_thr[YH_MAX_THRESHOLDS];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    WarpWindow &w = windows[wib];
    const int warps_total = gridDim.x * kWarpsPerCtaM;
    const int T = thr.n;
    for (int i = threadIdx.x; i < 4 * kC; i += blockDim.x) gt_hist[i] = 0;
    if (threadIdx.x < YH_MAX_THRESHOLDS) s_thr[threadIdx.x] = thr.t[threadIdx.x];
    __syncthreads();

    for (int img = blockIdx.x * kWarpsPerCtaM + wib; img < N; img += warps_total) {
        // ---- predictions: decode, order, NMS
        stage_image(w, pred + (size_t)img * kImgFloats, lane);
        prefetch_image(tgt + (size_t)img * kImgFloats, lane);
        if (img + warps_total < N) prefetch_image(pred + (size_t)(img + warps_total) * kImgFloats, lane);
        const Mask128 valid = decode_image(w, lane, conf_thr, nullptr);
        int n;
        const Mask128 keep = order_and_nms<YH_IOU_METRICS>(w, valid, nms_thr, lane, n);
        publish_keep(w, keep, lane);
        const unsigned lt = (1u << lane) - 1u;

        // ---- targets: parse into the overlay
        stage_image(w, tgt + (size_t)img * kImgFloats, lane);
        const bool lists = gt_raw != nullptr;
        const int ngt = parse_targets(w, lane, gt_hist, lists ? gt_cell + (size_t)img * kCells : nullptr,
                                      lists ? gt_cls + (size_t)img * kCells : nullptr,
                                      lists ? gt_raw + (size_t)img * kCells : nullptr);
        MatchOverlay &mo = *reinterpret_cast<MatchOverlay *>(w.stage);
        if (lane == 0) {
            n_keep[img] = keep.count();
            n_gt[img] = ngt;
        }

        // ---- best ground truth of every kept detection, per subset
        // subset 0 = all ground truths, 1/2/3 = small/medium/large only
        unsigned matched = 0;  // bit s: my slot-s detection met a ground truth of its class
#pragma unroll 1
        for (int s = 0; s < kSlots; ++s) {
            if (!((w.keepw[s] >> lane) & 1u)) continue;
            const int p = lane + 32 * s;
            const int k = w.order[p], c = w.cand.cls[k];
            double b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
            int g0 = 255, g1 = 255, g2 = 255, g3 = 255;
            bool met = false;
            for (int g = 0; g < ngt; ++g) {
                if (mo.gcls[g] != c) continue;
                met = true;
                Corners gc;
                gc.x1 = mo.gx1[g]; gc.y1 = mo.gy1[g]; gc.x2 = mo.gx2[g]; gc.y2 = mo.gy2[g];
                gc.area = mo.garea[g];
                const double iou = iou_exact<YH_IOU_METRICS>(cand_corners(w, k), gc);
                const int b = mo.gbucket[g];
                if (iou > b0) { b0 = iou; g0 = g; }  // strict: first GT wins ties, IoU 0 never matches
                if (b == 1 && iou > b1) { b1 = iou; g1 = g; }
                if (b == 2 && iou > b2) { b2 = iou; g2 = g; }
                if (b == 3 && iou > b3) { b3 = iou; g3 = g; }
            }
            mo.biou[0][p] = b0; mo.biou[1][p] = b1; mo.biou[2][p] = b2; mo.biou[3][p] = b3;
            mo.bgt[0][p] = (uint8_t)g0; mo.bgt[1][p] = (uint8_t)g1;
            mo.bgt[2][p] = (uint8_t)g2; mo.bgt[3][p] = (uint8_t)g3;
            if (met) matched |= 1u << s;
        }
        __syncwarp();

        // ---- greedy TP/FP: a detection is a TP at threshold t iff its best IoU >= t
        // and no earlier detection of its class claimed the same ground truth with
        // IoU >= t (metrics.py:406-413).  "Earlier" = lower sorted position.
#pragma unroll 1
        for (int s = 0; s < kSlots; ++s) {
            if (!((w.keepw[s] >> lane) & 1u)) continue;
            const int p = lane + 32 * s;
            const int k = w.order[p], c = w.cand.cls[k];
            const int lo = w.cls_start[c];
            uint64_t flags = 0;
            if ((matched >> s) & 1u) {
#pragma unroll 1
                for (int q = 0; q < 4; ++q) {
                    const int g = mo.bgt[q][p];
                    if (g == 255) continue;
                    const double mine = mo.biou[q][p];
                    double rival = 0.0;  // best IoU an earlier detection brought to the same GT
#pragma unroll 1
                    for (int e = lo; e < p; ++e)
                        if (kept_at(w, e) && mo.bgt[q][e] == g) rival = fmax(rival, mo.biou[q][e]);
#pragma unroll 1
                    for (int t = 0; t < T; ++t) {
                        const double th = s_thr[t];
                        if (mine >= th && !(rival >= th)) flags |= 1ull << (q * T + t);
                    }
                    if (q == 0 && mine >= 0.5 && !(rival >= 0.5)) flags |= 1ull << (4 * T);
                }
            }
            // emission slot: sorted order among the kept
            const int j = w.keepbase[s] + __popc(w.keepw[s] & lt);
            const size_t o = (size_t)img * kCand + j;
            const double cf = w.cand.conf[k];
            rec_key[o] = ~orderable(cf == 0.0 ? 0.0 : cf);
            rec_flags[o] = flags | ((uint64_t)c << 56);
            rec_k[o] = (uint8_t)k;
            if (rec_raw != nullptr) {  // raw (x, y, w, h) of the box; yh_list_form rebuilds the tuple
                const float *raw = pred + (size_t)img * kImgFloats + (k / kB) * kD + (k % kB) * 5;
                rec_raw[o] = make_float4(__ldg(raw + 0), __ldg(raw + 1), __ldg(raw + 2), __ldg(raw + 3));
            }
        }
        __syncwarp();
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
                             float *rec_raw, int32_t *n_gt, uint8_t *gt_cell, uint8_t *gt_cls,
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
    if (((uintptr_t)pred % 8) || ((uintptr_t)tgt % 8) || ((uintptr_t)rec_raw % 16) || ((uintptr_t)gt_raw % 16)) {
        set_error("yh_map_update: pred/tgt must be 8-byte aligned, rec_raw/gt_raw 16-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    if (N == 0) return 0;
    Thresholds th;
    th.n = T;
    for (int i = 0; i < YH_MAX_THRESHOLDS; ++i) th.t[i] = i < T ? thresholds_host[i] : 2.0;
    const size_t smem = sizeof(WarpWindow) * kWarpsPerCtaM;
    static bool attr_done = false;
    if (!attr_done) {
        cudaFuncSetAttribute(map_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_done = true;
    }
    int ctas = (N + kWarpsPerCtaM - 1) / kWarpsPerCtaM;
    const int cap = 148 * 5;
    if (ctas > cap) ctas = cap;
    map_update_kernel<<<ctas, kWarpsPerCtaM * 32, smem, (cudaStream_t)stream>>>(
        pred, tgt, N, conf_thr, nms_thr, th, n_keep, rec_key, rec_flags, rec_k, (float4 *)rec_raw, n_gt, gt_cell,
        gt_cls, (float4 *)gt_raw, gt_count);
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
