// lists.cu -- list-form entry points: the reference's helpers accept arbitrary
// Python lists of fp64 boxes (tests/test_metrics.py:98-117, tests/test_yolo.py:229-313
// build them by hand; compute() runs on whatever all_predictions holds).  These
// kernels take CSR lists; one CTA per list.  They are the general (any n <= 1024
// per list) counterparts of the warp-per-image kernels and favour clarity: the
// large-batch paths are yh_decode_nms / yh_map_update.
#include "common.cuh"

namespace yh {

constexpr int kListMax = 1024;
constexpr int kListThreads = 256;

__device__ __forceinline__ Corners corners_from(const double *b) {
    return corners_of(b[0], b[1], b[2], b[3]);
}

template <int MODE>
__global__ void iou_pairs_kernel(const double *__restrict__ a, const double *__restrict__ b, int64_t n,
                                 double *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = iou_exact<MODE>(corners_from(a + 4 * i), corners_from(b + 4 * i));
}

// Greedy NMS of one list (metrics.py:258-296 / inference.py:298-317).
template <int MODE>
__global__ void __launch_bounds__(kListThreads)
nms_lists_kernel(const int64_t *__restrict__ off, const int32_t *__restrict__ cls,
                 const double *__restrict__ conf, const double *__restrict__ box, double thr, int order,
                 int32_t *__restrict__ n_keep, int32_t *__restrict__ keep_idx) {
    __shared__ short s_order[kListMax];   // rank -> element
    __shared__ short s_first[kListMax];   // rank -> rank of the first element of its class
    __shared__ unsigned char s_dead[kListMax];
    __shared__ int s_cur;
    const int64_t base = off[blockIdx.x];
    const int n = (int)(off[blockIdx.x + 1] - base);
    const int tid = threadIdx.x;
    cls += base; conf += base; box += 4 * base; keep_idx += base;
    // stable descending order by counting
    for (int i = tid; i < n; i += kListThreads) {
        const double mine = conf[i];
        int r = 0;
        for (int m = 0; m < n; ++m) {
            const double o = conf[m];
            r += (o > mine) || (o == mine && m < i);
        }
        s_order[r] = (short)i;
        s_dead[i] = 0;
    }
    __syncthreads();
    for (int r = tid; r < n; r += kListThreads) {
        const int c = cls[s_order[r]];
        int f = r;
        for (int q = 0; q < r; ++q)
            if (cls[s_order[q]] == c) { f = q; break; }
        s_first[r] = (short)f;
    }
    __syncthreads();
    // greedy pass in rank order; s_dead is indexed by rank
    for (int r = 0; r < n; ++r) {
        if (s_dead[r]) continue;  // uniform: shared value, synchronised below
        const int cur = s_order[r];
        const int c = cls[cur];
        const Corners a = corners_from(box + 4 * cur);
        for (int q = r + 1 + tid; q < n; q += kListThreads) {
            if (s_dead[q]) continue;
            const int o = s_order[q];
            if (cls[o] != c) continue;
            if (!iou_below<MODE>(a, corners_from(box + 4 * o), thr)) s_dead[q] = 1;
        }
        __syncthreads();
    }
    // output positions
    if (tid == 0) s_cur = 0;
    __syncthreads();
    for (int r = tid; r < n; r += kListThreads) {
        if (s_dead[r]) continue;
        int pos = 0;
        if (order == YH_ORDER_GLOBAL_CONF) {
            for (int q = 0; q < r; ++q) pos += !s_dead[q];
        } else {
            const int g = s_first[r];
            for (int q = 0; q < n; ++q)
                if (!s_dead[q]) pos += (s_first[q] < g) || (s_first[q] == g && q < r);
        }
        keep_idx[pos] = s_order[r];
        atomicAdd(&s_cur, 1);
    }
    __syncthreads();
    if (tid == 0) n_keep[blockIdx.x] = s_cur;
}

struct ListThresholds {
    double t[YH_MAX_THRESHOLDS];
    int n;
};

// TP flags of one image's list-form detections (metrics.py:391-413, 465-486, 613-633).
__global__ void __launch_bounds__(kListThreads)
match_lists_kernel(const int64_t *__restrict__ det_off, const int32_t *__restrict__ det_cls,
                   const double *__restrict__ det_conf, const double *__restrict__ det_box,
                   const int64_t *__restrict__ gt_off, const int32_t *__restrict__ gt_cls,
                   const double *__restrict__ gt_box, int C, ListThresholds thr,
                   uint64_t *__restrict__ rec_key, uint64_t *__restrict__ rec_flags,
                   int32_t *__restrict__ gt_count) {
    __shared__ double s_biou[4][kListMax];
    __shared__ short s_bgt[4][kListMax];
    const int64_t d0 = det_off[blockIdx.x], g0 = gt_off[blockIdx.x];
    const int nd = (int)(det_off[blockIdx.x + 1] - d0), ng = (int)(gt_off[blockIdx.x + 1] - g0);
    const int tid = threadIdx.x;
    const double small_a = __dmul_rn(__ddiv_rn(32.0, 448.0), __ddiv_rn(32.0, 448.0));
    const double medium_a = __dmul_rn(__ddiv_rn(96.0, 448.0), __ddiv_rn(96.0, 448.0));
    for (int g = tid; g < ng; g += kListThreads) {
        const double *b = gt_box + 4 * (g0 + g);
        const double area = __dmul_rn(b[2], b[3]);
        const int bucket = (area < small_a) ? 1 : ((area < medium_a) ? 2 : 3);
        const int c = gt_cls[g0 + g];
        if (c >= 0 && c < C) {
            atomicAdd(&gt_count[c], 1);
            atomicAdd(&gt_count[bucket * C + c], 1);
        }
    }
    for (int d = tid; d < nd; d += kListThreads) {
        const int c = det_cls[d0 + d];
        const Corners dc = corners_from(det_box + 4 * (d0 + d));
        double bi[4] = {0.0, 0.0, 0.0, 0.0};
        int bg[4] = {-1, -1, -1, -1};
        for (int g = 0; g < ng; ++g) {
            if (gt_cls[g0 + g] != c) continue;
            const double *b = gt_box + 4 * (g0 + g);
            const Corners gc = corners_from(b);
            const double iou = iou_exact<YH_IOU_METRICS>(dc, gc);
            const int bucket = (gc.area < small_a) ? 1 : ((gc.area < medium_a) ? 2 : 3);
            if (iou > bi[0]) { bi[0] = iou; bg[0] = g; }
#pragma unroll
            for (int q = 1; q < 4; ++q)
                if (bucket == q && iou > bi[q]) { bi[q] = iou; bg[q] = g; }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            s_biou[q][d] = bi[q];
            s_bgt[q][d] = (short)bg[q];
        }
    }
    __syncthreads();
    for (int d = tid; d < nd; d += kListThreads) {
        const int c = det_cls[d0 + d];
        const double cf = det_conf[d0 + d];
        uint64_t flags = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int g = s_bgt[q][d];
            if (g < 0) continue;
            const double mine = s_biou[q][d];
            double rival = 0.0;
            for (int e = 0; e < nd; ++e) {
                if (e == d || det_cls[d0 + e] != c || s_bgt[q][e] != g) continue;
                const double ce = det_conf[d0 + e];
                // processed before me in the global stable sort: higher confidence, or equal
                // confidence and earlier in the list
                if ((ce > cf) || (ce == cf && e < d)) rival = fmax(rival, s_biou[q][e]);
            }
            for (int t = 0; t < thr.n; ++t)
                if (mine >= thr.t[t] && !(rival >= thr.t[t])) flags |= 1ull << (q * thr.n + t);
            if (q == 0 && mine >= 0.5 && !(rival >= 0.5)) flags |= 1ull << (4 * thr.n);
        }
        rec_key[d0 + d] = ~orderable(cf == 0.0 ? 0.0 : cf);
        rec_flags[d0 + d] = flags | ((uint64_t)(c & 0xff) << 56);
    }
}

}  // namespace yh

using namespace yh;

extern "C" int yh_iou_pairs(const double *box_a, const double *box_b, int64_t n, int iou_mode,
                            double *out, void *stream) {
    if (n < 0 || (n > 0 && (!box_a || !box_b || !out))) {
        set_error("yh_iou_pairs: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (n == 0) return 0;
    const unsigned blocks = (unsigned)((n + 127) / 128);
    if (iou_mode == YH_IOU_METRICS)
        iou_pairs_kernel<YH_IOU_METRICS><<<blocks, 128, 0, (cudaStream_t)stream>>>(box_a, box_b, n, out);
    else if (iou_mode == YH_IOU_INFER)
        iou_pairs_kernel<YH_IOU_INFER><<<blocks, 128, 0, (cudaStream_t)stream>>>(box_a, box_b, n, out);
    else {
        set_error("yh_iou_pairs: unknown iou_mode %d", iou_mode);
        return YH_ERR_BAD_ARG;
    }
    return check_launch("yh_iou_pairs");
}

extern "C" int yh_nms_lists(const int64_t *off, int n_lists, const int32_t *cls, const double *conf,
                            const double *box, double nms_thr, int iou_mode, int order,
                            int32_t *n_keep, int32_t *keep_idx, int max_len, void *stream) {
    if (n_lists < 0 || !off || !n_keep || (max_len > 0 && (!cls || !conf || !box || !keep_idx))) {
        set_error("yh_nms_lists: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (max_len > kListMax) {
        set_error("yh_nms_lists: a list holds %d boxes; this build carries at most %d per list", max_len,
                  kListMax);
        return YH_ERR_UNSUPPORTED;
    }
    if (order != YH_ORDER_CLASS_GROUPED && order != YH_ORDER_GLOBAL_CONF) {
        set_error("yh_nms_lists: unknown order %d", order);
        return YH_ERR_BAD_ARG;
    }
    if (n_lists == 0) return 0;
    if (iou_mode == YH_IOU_METRICS)
        nms_lists_kernel<YH_IOU_METRICS><<<n_lists, kListThreads, 0, (cudaStream_t)stream>>>(
            off, cls, conf, box, nms_thr, order, n_keep, keep_idx);
    else if (iou_mode == YH_IOU_INFER)
        nms_lists_kernel<YH_IOU_INFER><<<n_lists, kListThreads, 0, (cudaStream_t)stream>>>(
            off, cls, conf, box, nms_thr, order, n_keep, keep_idx);
    else {
        set_error("yh_nms_lists: unknown iou_mode %d", iou_mode);
        return YH_ERR_BAD_ARG;
    }
    return check_launch("yh_nms_lists");
}

extern "C" int yh_match_lists(const int64_t *det_off, const int32_t *det_cls, const double *det_conf,
                              const double *det_box, const int64_t *gt_off, const int32_t *gt_cls,
                              const double *gt_box, int n_lists, int C, const double *thresholds_host,
                              int T, uint64_t *rec_key, uint64_t *rec_flags, int32_t *gt_count,
                              int max_dets, void *stream) {
    if (C < 1 || C > 32 || T < 1 || T > YH_MAX_THRESHOLDS) {
        set_error("yh_match_lists: C=%d / T=%d outside what this build carries (C<=32, T<=%d)", C, T,
                  YH_MAX_THRESHOLDS);
        return YH_ERR_UNSUPPORTED;
    }
    if (max_dets > kListMax) {
        set_error("yh_match_lists: an image holds %d detections; this build carries at most %d", max_dets,
                  kListMax);
        return YH_ERR_UNSUPPORTED;
    }
    if (n_lists < 0 || !det_off || !gt_off || !thresholds_host || !gt_count) {
        set_error("yh_match_lists: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (n_lists == 0) return 0;
    ListThresholds th;
    th.n = T;
    for (int i = 0; i < YH_MAX_THRESHOLDS; ++i) th.t[i] = i < T ? thresholds_host[i] : 2.0;
    match_lists_kernel<<<n_lists, kListThreads, 0, (cudaStream_t)stream>>>(
        det_off, det_cls, det_conf, det_box, gt_off, gt_cls, gt_box, C, th, rec_key, rec_flags, gt_count);
    return check_launch("yh_match_lists");
}
