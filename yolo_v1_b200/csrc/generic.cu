// generic.cu -- the head for shapes other than S=7, B=2, C=20.
//
// The reference's classes take the grid size, the boxes per cell and the class count as constructor
// arguments (YOLOLoss(S, B, C, ...) loss.py:28-53; mAPMetric(num_classes, ..., S, B) metrics.py:43-71;
// YOLOInference reads model.S / model.B, inference.py:170-171).  The kernels in loss.cu, decode_nms.cu and
// map_update.cu are specialised for the VOC head the reference trains (7 x 7 x 30) -- warp-per-image layouts,
// unrolled 20-class loops, a 98-candidate shared-memory window.  Every other shape runs here: the same
// arithmetic, operator for operator (fp64 geometry with one rounding per operator, Python's max / min / strict
// and non-strict comparisons, torch's first-maximum argmax with NaN maximal, stable orders), written as plain
// loops with ONE THREAD PER IMAGE (per cell for the loss).  Correct on the GPU and in the same launch-only style
// (no allocation, no synchronisation), but with none of the specialised kernels' throughput: a thread walks its
// image in local memory.  Limits: S*S*B <= 255 candidates (candidate indices travel as bytes), B <= 8,
// C <= 32 for the metric / inference entry points (the sort packs the class into 5 bits), any C for the loss.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace yh {

constexpr int kGenMaxK = 255;
constexpr int kGenMaxB = 8;
constexpr int kGenMaxG = 255;   // ground truths per image (cells)

bool generic_shape_ok(int S, int B, int C, bool with_classes_in_flags) {
    if (S < 1 || B < 1 || C < 1 || B > kGenMaxB) return false;
    if ((long long)S * S * B > kGenMaxK) return false;
    if (with_classes_in_flags && C > 32) return false;
    return true;
}

struct GenShape {
    int S, B, C, D, K;
};

// x_abs = (col + x) / S in doubles (metrics.py:200-201): a true division for any S
__device__ __forceinline__ double gen_abs(int cell_coord, float v, int S) {
    return __ddiv_rn(__dadd_rn((double)cell_coord, (double)v), (double)S);
}
__device__ __forceinline__ Corners gen_corners(const float *__restrict__ img, const GenShape &sh, int k) {
    const int cell = k / sh.B, b = k - cell * sh.B;
    const float *p = img + (size_t)cell * sh.D + b * 5;
    return corners_of(gen_abs(cell % sh.S, __ldg(p), sh.S), gen_abs(cell / sh.S, __ldg(p + 1), sh.S), (double)__ldg(p + 2),
                      (double)__ldg(p + 3));
}
// torch.argmax over the C class scores of a cell: first maximum, NaN maximal (the first NaN wins)
__device__ __forceinline__ int gen_argmax(const float *__restrict__ cls, int C, float &best_v) {
    float m = __ldg(cls);
    int best = 0;
    for (int i = 1; i < C; ++i) {
        const float v = __ldg(cls + i);
        if ((v > m || v != v) && m == m) {   // strictly greater, or the first NaN
            m = v;
            best = i;
        }
    }
    best_v = m;
    return best;
}

// One image: candidates above the threshold, ordered by (class asc, confidence desc, candidate asc), greedy
// per-class NMS.  Fills key[k] / cls[k] for valid candidates, ord[0..n) (sorted position -> candidate) and
// kept[p]; returns n.  bad: a valid candidate outside [0, 1] (schemas.py:9-16,88).
template <int MODE>
__device__ int gen_decode_order_nms(const float *__restrict__ img, const GenShape &sh, double conf_thr, double nms_thr,
                                    uint64_t *key, uint8_t *cls, uint8_t *ord, uint8_t *kept, bool *bad) {
    int n = 0;
    bool any_bad = false;
    for (int cell = 0; cell < sh.S * sh.S; ++cell) {
        const float *c = img + (size_t)cell * sh.D;
        float pv;
        const int cid = gen_argmax(c + sh.B * 5, sh.C, pv);
        const double prob = (double)pv;
        for (int b = 0; b < sh.B; ++b) {
            const double conf = __dmul_rn((double)__ldg(c + b * 5 + 4), prob);   // exact: 24 x 24 bits
            if (!(conf > conf_thr)) continue;                                     // metrics.py:209, strict
            const int k = cell * sh.B + b;
            key[k] = ~orderable(conf == 0.0 ? 0.0 : conf);
            cls[k] = (uint8_t)cid;
            ord[n++] = (uint8_t)k;
            if (bad != nullptr) {
                const double xa = gen_abs(cell % sh.S, __ldg(c + b * 5), sh.S), ya = gen_abs(cell / sh.S, __ldg(c + b * 5 + 1), sh.S);
                const double ww = (double)__ldg(c + b * 5 + 2), hh = (double)__ldg(c + b * 5 + 3);
                any_bad |= !(xa >= 0.0 && xa <= 1.0 && ya >= 0.0 && ya <= 1.0 && ww >= 0.0 && ww <= 1.0 && hh >= 0.0 &&
                             hh <= 1.0 && conf >= 0.0 && conf <= 1.0);
            }
        }
    }
    if (bad != nullptr) *bad = any_bad;
    // insertion sort by (class, key, candidate): candidates arrive in ascending k, so equal (class, key) keep it
    for (int i = 1; i < n; ++i) {
        const int k = ord[i];
        const int ck = cls[k];
        const uint64_t kk = key[k];
        int j = i - 1;
        while (j >= 0) {
            const int o = ord[j];
            if (cls[o] < ck || (cls[o] == ck && key[o] <= kk)) break;
            ord[j + 1] = (uint8_t)o;
            --j;
        }
        ord[j + 1] = (uint8_t)k;
    }
    for (int p = 0; p < n; ++p) kept[p] = 1;
    if (nms_thr == nms_thr) {   // a NaN threshold means "decode only"
        for (int p = 0; p < n; ++p) {
            if (!kept[p]) continue;
            const Corners a = gen_corners(img, sh, ord[p]);
            for (int q = p + 1; q < n && cls[ord[q]] == cls[ord[p]]; ++q) {
                if (!kept[q]) continue;
                // keeps IoU < thr (metrics.py:293, inference.py:314); argument order iou(current, other)
                if (!iou_below<MODE>(a, gen_corners(img, sh, ord[q]), nms_thr)) kept[q] = 0;
            }
        }
    }
    return n;
}

// ---------------------------------------------------------------- decode + NMS
template <int MODE>
__global__ void __launch_bounds__(64)
gen_decode_nms_kernel(const float *__restrict__ pred, int N, GenShape sh, double conf_thr, double nms_thr, int order,
                      int32_t *__restrict__ n_keep, uint8_t *__restrict__ keep_k, uint8_t *__restrict__ keep_cls,
                      double *__restrict__ keep_conf, double *__restrict__ keep_box, int32_t *__restrict__ n_cand,
                      uint8_t *__restrict__ bad_range) {
    const int img = blockIdx.x * blockDim.x + threadIdx.x;
    if (img >= N) return;
    const float *pimg = pred + (size_t)img * sh.S * sh.S * sh.D;
    uint64_t key[kGenMaxK];
    uint8_t cls[kGenMaxK], ord[kGenMaxK], kept[kGenMaxK], out[kGenMaxK];
    bool bad = false;
    const int n = gen_decode_order_nms<MODE>(pimg, sh, conf_thr, nms_thr, key, cls, ord, kept, bad_range != nullptr ? &bad : nullptr);
    if (bad_range != nullptr) bad_range[img] = bad ? 1 : 0;
    if (n_cand != nullptr) n_cand[img] = n;
    // output order
    int nk = 0;
    for (int p = 0; p < n; ++p)
        if (kept[p]) out[nk++] = ord[p];   // class-grouped, class ascending: regroup below
    if (order == YH_ORDER_GLOBAL_CONF) {
        // global (confidence desc, candidate asc): insertion sort of the kept by (key, candidate)
        for (int i = 1; i < nk; ++i) {
            const int k = out[i];
            int j = i - 1;
            while (j >= 0 && (key[out[j]] > key[k] || (key[out[j]] == key[k] && out[j] > k))) {
                out[j + 1] = out[j];
                --j;
            }
            out[j + 1] = (uint8_t)k;
        }
    } else {
        // mAPMetric._apply_nms: class groups in the order their best member appears in the confidence-sorted
        // list (metrics.py:274-279); `out` holds the groups in class order, each confidence-descending.  Select
        // groups by their head, smallest (key, candidate) first.
        uint8_t tmp[kGenMaxK];
        for (int i = 0; i < nk; ++i) tmp[i] = out[i];
        int written = 0;
        for (;;) {
            int best = -1;
            for (int i = 0; i < nk; ++i) {
                if (tmp[i] == 255) continue;
                const bool head = (i == 0) || tmp[i - 1] == 255 || cls[tmp[i - 1]] != cls[tmp[i]];
                if (!head) continue;
                if (best < 0 || key[tmp[i]] < key[tmp[best]] || (key[tmp[i]] == key[tmp[best]] && tmp[i] < tmp[best])) best = i;
            }
            if (best < 0) break;
            const int c = cls[tmp[best]];
            for (int i = best; i < nk && tmp[i] != 255 && cls[tmp[i]] == c; ++i) {
                out[written++] = tmp[i];
                tmp[i] = 255;
            }
        }
    }
    n_keep[img] = nk;
    const size_t base = (size_t)img * sh.K;
    for (int q = 0; q < nk; ++q) {
        const int k = out[q];
        keep_k[base + q] = (uint8_t)k;
        if (keep_cls != nullptr) keep_cls[base + q] = cls[k];
        if (keep_conf != nullptr) keep_conf[base + q] = from_orderable(~key[k]);
        if (keep_box != nullptr) {
            const int cell = k / sh.B, b = k - cell * sh.B;
            const float *p = pimg + (size_t)cell * sh.D + b * 5;
            double *o = keep_box + 4 * (base + q);
            o[0] = gen_abs(cell % sh.S, __ldg(p), sh.S);
            o[1] = gen_abs(cell / sh.S, __ldg(p + 1), sh.S);
            o[2] = (double)__ldg(p + 2);
            o[3] = (double)__ldg(p + 3);
        }
    }
}

// ---------------------------------------------------------------- update: decode + NMS + ground truths + TP/FP
struct GenThresholds {
    double t[YH_MAX_THRESHOLDS];
    int n;
};

__global__ void __launch_bounds__(64)
gen_map_update_kernel(const float *__restrict__ pred, const float *__restrict__ tgt, int N, GenShape sh, double conf_thr,
                      double nms_thr, GenThresholds thr, int32_t *__restrict__ n_keep, uint64_t *__restrict__ rec_key,
                      uint64_t *__restrict__ rec_flags, uint8_t *__restrict__ rec_k, uint64_t *__restrict__ rec_box,
                      int32_t *__restrict__ n_gt, uint8_t *__restrict__ gt_cell, uint8_t *__restrict__ gt_cls,
                      float4 *__restrict__ gt_raw, int32_t *__restrict__ gt_count) {
    const int img = blockIdx.x * blockDim.x + threadIdx.x;
    if (img >= N) return;
    const size_t img_floats = (size_t)sh.S * sh.S * sh.D;
    const float *pimg = pred + (size_t)img * img_floats, *timg = tgt + (size_t)img * img_floats;
    uint64_t key[kGenMaxK];
    uint8_t cls[kGenMaxK], ord[kGenMaxK], kept[kGenMaxK];
    const int n = gen_decode_order_nms<YH_IOU_METRICS>(pimg, sh, conf_thr, nms_thr, key, cls, ord, kept, nullptr);
    // ground truths: slot-0 rule (metrics.py:239), in cell order
    uint8_t gcell[kGenMaxG], gcls[kGenMaxG], gbucket[kGenMaxG];
    const double small_a = __dmul_rn(__ddiv_rn(32.0, 448.0), __ddiv_rn(32.0, 448.0));   // metrics.py:511
    const double medium_a = __dmul_rn(__ddiv_rn(96.0, 448.0), __ddiv_rn(96.0, 448.0));  // metrics.py:512
    int ngt = 0;
    const int G = sh.S * sh.S;
    for (int cell = 0; cell < G; ++cell) {
        const float *c = timg + (size_t)cell * sh.D;
        if (!(__ldg(c + 4) > 0.f)) continue;
        float pv;
        const int cid = gen_argmax(c + sh.B * 5, sh.C, pv);
        const double area = __dmul_rn((double)__ldg(c + 2), (double)__ldg(c + 3));
        const int bucket = (area < small_a) ? 1 : ((area < medium_a) ? 2 : 3);
        gcell[ngt] = (uint8_t)cell;
        gcls[ngt] = (uint8_t)cid;
        gbucket[ngt] = (uint8_t)bucket;
        if (gt_raw != nullptr) {
            const size_t o = (size_t)img * G + ngt;
            gt_cell[o] = (uint8_t)cell;
            gt_cls[o] = (uint8_t)cid;
            gt_raw[o] = make_float4(__ldg(c), __ldg(c + 1), __ldg(c + 2), __ldg(c + 3));
        }
        atomicAdd(&gt_count[cid], 1);
        atomicAdd(&gt_count[bucket * sh.C + cid], 1);
        ++ngt;
    }
    n_gt[img] = ngt;
    // kept detections in sorted order; per (detection, subset): best ground truth and the thresholds its IoU
    // reaches (metrics.py:391-403, 613-623); greedy TP/FP against earlier detections of the class (:406-413)
    const int T = thr.n;
    unsigned short imask[kGenMaxK][4];
    uint8_t bgt[kGenMaxK][4], kcls[kGenMaxK];
    int nk = 0;
    const size_t base = (size_t)img * sh.K;
    for (int p = 0; p < n; ++p) {
        if (!kept[p]) continue;
        const int k = ord[p], c = cls[k];
        const Corners dc = gen_corners(pimg, sh, k);
        double best[4] = {0.0, 0.0, 0.0, 0.0};
        int bg[4] = {255, 255, 255, 255};
        for (int g = 0; g < ngt; ++g) {
            if (gcls[g] != c) continue;
            const double iou = iou_exact<YH_IOU_METRICS>(dc, gen_corners(timg, sh, gcell[g] * sh.B));
            const int b = gbucket[g];
            if (iou > best[0]) { best[0] = iou; bg[0] = g; }   // strict: first ground truth wins, IoU 0 never matches
            if (iou > best[b]) { best[b] = iou; bg[b] = g; }
        }
        uint64_t flags = 0;
        for (int q = 0; q < 4; ++q) {
            unsigned m = 0;
            if (bg[q] != 255) {
                for (int t = 0; t < T; ++t) m |= (best[q] >= thr.t[t] ? 1u : 0u) << t;
                m |= (best[q] >= 0.5 ? 1u : 0u) << T;
            }
            imask[nk][q] = (unsigned short)m;
            bgt[nk][q] = (uint8_t)bg[q];
            unsigned claimed = 0;
            if (bg[q] != 255)
                for (int e = nk - 1; e >= 0 && kcls[e] == c; --e)
                    if (bgt[e][q] == bg[q]) claimed |= imask[e][q];
            const unsigned bits = m & ~claimed;
            flags |= (uint64_t)(bits & ((1u << T) - 1u)) << (q * T);
            if (q == 0 && ((bits >> T) & 1u)) flags |= 1ull << (4 * T);
        }
        kcls[nk] = (uint8_t)c;
        rec_key[base + nk] = key[k];
        rec_flags[base + nk] = flags | ((uint64_t)c << 56);
        if (rec_k != nullptr) rec_k[base + nk] = (uint8_t)k;
        if (rec_box != nullptr) {
            const int cell = k / sh.B, b = k - cell * sh.B;
            const float *pp = pimg + (size_t)cell * sh.D + b * 5;
            uint64_t *o = rec_box + 3 * (base + nk);
            o[0] = (uint64_t)__double_as_longlong(gen_abs(cell % sh.S, __ldg(pp), sh.S));
            o[1] = (uint64_t)__double_as_longlong(gen_abs(cell / sh.S, __ldg(pp + 1), sh.S));
            o[2] = (uint64_t)__float_as_uint(__ldg(pp + 2)) | ((uint64_t)__float_as_uint(__ldg(pp + 3)) << 32);
        }
        ++nk;
    }
    n_keep[img] = nk;
}

// ground truths in list form (metrics.py:220-256)
__global__ void __launch_bounds__(64)
gen_parse_gt_kernel(const float *__restrict__ tgt, int N, GenShape sh, int32_t *__restrict__ n_gt,
                    uint8_t *__restrict__ gt_cls, double *__restrict__ gt_box) {
    const int img = blockIdx.x * blockDim.x + threadIdx.x;
    if (img >= N) return;
    const int G = sh.S * sh.S;
    const float *timg = tgt + (size_t)img * G * sh.D;
    int ngt = 0;
    for (int cell = 0; cell < G; ++cell) {
        const float *c = timg + (size_t)cell * sh.D;
        if (!(__ldg(c + 4) > 0.f)) continue;
        float pv;
        const size_t o = (size_t)img * G + ngt;
        gt_cls[o] = (uint8_t)gen_argmax(c + sh.B * 5, sh.C, pv);
        gt_box[4 * o + 0] = gen_abs(cell % sh.S, __ldg(c), sh.S);
        gt_box[4 * o + 1] = gen_abs(cell / sh.S, __ldg(c + 1), sh.S);
        gt_box[4 * o + 2] = (double)__ldg(c + 2);
        gt_box[4 * o + 3] = (double)__ldg(c + 3);
        ++ngt;
    }
    n_gt[img] = ngt;
}

// ---------------------------------------------------------------- loss forward + backward, one thread per cell
template <typename T> __device__ __forceinline__ float gen_ld(const T *p, size_t i) { return (float)p[i]; }
template <> __device__ __forceinline__ float gen_ld<float>(const float *p, size_t i) { return __ldg(p + i); }
template <typename T> __device__ __forceinline__ void gen_st(T *p, size_t i, float v) { p[i] = (T)v; }

__device__ __forceinline__ float g_max(float a, float b) { return (a > b || a != a) ? a : b; }   // torch.maximum
__device__ __forceinline__ float g_min(float a, float b) { return (a < b || a != a) ? a : b; }
__device__ __forceinline__ float g_max_mask(float self, float other) { return (self < other) ? 0.f : ((self == other) ? 0.5f : 1.f); }
__device__ __forceinline__ float g_min_mask(float self, float other) { return (self > other) ? 0.f : ((self == other) ? 0.5f : 1.f); }

constexpr int kGenLossThreads = 128;

template <typename T>
__global__ void __launch_bounds__(kGenLossThreads)
gen_loss_kernel(const T *__restrict__ pred, const float *__restrict__ tgt, T *__restrict__ grad, long long n_cells, int B,
                int C, float lambda_coord, float lambda_noobj, float inv_n, double *__restrict__ partials) {
    __shared__ double s_red[kGenLossThreads / 32][4];
    const long long cell = (long long)blockIdx.x * kGenLossThreads + threadIdx.x;
    const int D = B * 5 + C;
    float acc_coord = 0.f, acc_obj = 0.f, acc_noobj = 0.f, acc_cls = 0.f;
    if (cell < n_cells) {
        const size_t o = (size_t)cell * D;
        // loss.py:98-104: object cell = any slot filled; target box = the first filled slot (slot 0 if none)
        int slot = -1;
        for (int b = 0; b < B; ++b)
            if (slot < 0 && __ldg(tgt + o + b * 5 + 4) > 0.f) slot = b;
        const bool obj = slot >= 0;
        const float gn = inv_n * lambda_noobj;
        if (!obj) {
            for (int b = 0; b < B; ++b) {
                const float c = gen_ld(pred, o + b * 5 + 4);
                acc_noobj += c * c;
                if (grad != nullptr) {
                    for (int i = 0; i < 4; ++i) gen_st(grad, o + b * 5 + i, 0.f);
                    gen_st(grad, o + b * 5 + 4, gn * (2.f * c));
                }
            }
            if (grad != nullptr)
                for (int i = 0; i < C; ++i) gen_st(grad, o + B * 5 + i, 0.f);
        } else {
            const float tx = __ldg(tgt + o + slot * 5), ty = __ldg(tgt + o + slot * 5 + 1);
            const float tw = __ldg(tgt + o + slot * 5 + 2), th = __ldg(tgt + o + slot * 5 + 3);
            const float thw = tw * 0.5f, thh = th * 0.5f;
            const float tx1 = tx - thw, tx2 = tx + thw, ty1 = ty - thh, ty2 = ty + thh;
            const float tarea = tw * th;
            // IoU of every predicted box with the target box (loss.py:191-212), argmax = first maximum, NaN maximal
            int best = 0;
            float best_iou = 0.f;
            for (int b = 0; b < B; ++b) {
                const float x = gen_ld(pred, o + b * 5), y = gen_ld(pred, o + b * 5 + 1);
                const float w = gen_ld(pred, o + b * 5 + 2), h = gen_ld(pred, o + b * 5 + 3);
                const float hw = w * 0.5f, hh = h * 0.5f;
                const float rawx = g_min(x + hw, tx2) - g_max(x - hw, tx1), rawy = g_min(y + hh, ty2) - g_max(y - hh, ty1);
                const float iwc = (rawx < 0.f) ? 0.f : rawx, ihc = (rawy < 0.f) ? 0.f : rawy;   // clamp(min=0): NaN passes
                const float inter = iwc * ihc;
                const float iou = inter / (((w * h + tarea) - inter) + 1e-6f);
                if (b == 0 || ((iou > best_iou || iou != iou) && best_iou == best_iou)) {
                    best = b;
                    best_iou = iou;
                }
            }
            for (int b = 0; b < B; ++b) {
                const float cf = gen_ld(pred, o + b * 5 + 4);
                if (b != best) {   // the other boxes of an object cell count as "no object" (loss.py:150-153)
                    acc_noobj += cf * cf;
                    if (grad != nullptr) {
                        for (int i = 0; i < 4; ++i) gen_st(grad, o + b * 5 + i, 0.f);
                        gen_st(grad, o + b * 5 + 4, gn * (2.f * cf));
                    }
                    continue;
                }
                const float x = gen_ld(pred, o + b * 5), y = gen_ld(pred, o + b * 5 + 1);
                const float w = gen_ld(pred, o + b * 5 + 2), h = gen_ld(pred, o + b * 5 + 3);
                const float hw = w * 0.5f, hh = h * 0.5f;
                const float px1 = x - hw, px2 = x + hw, py1 = y - hh, py2 = y + hh;
                const float rawx = g_min(px2, tx2) - g_max(px1, tx1), rawy = g_min(py2, ty2) - g_max(py1, ty1);
                const float iw = (rawx < 0.f) ? 0.f : rawx, ih = (rawy < 0.f) ? 0.f : rawy;
                const float inter = iw * ih;
                const float U = ((w * h + tarea) - inter) + 1e-6f;
                const float iou = inter / U;
                // coordinate term (loss.py:126-137)
                const float dx = x - tx, dy = y - ty;
                const float wc = g_max(w, 1e-6f), hc = g_max(h, 1e-6f);
                const float sw = sqrtf(wc), shh = sqrtf(hc);
                const float dw = sw - sqrtf(g_max(tw, 1e-6f)), dh = shh - sqrtf(g_max(th, 1e-6f));
                acc_coord += (dx * dx + dy * dy) + (dw * dw + dh * dh);
                // confidence term (loss.py:142-146); the IoU target is not detached
                const float dc = cf - iou;
                acc_obj += dc * dc;
                if (grad != nullptr) {
                    const float gc = inv_n * lambda_coord;
                    float gx = gc * (2.f * dx), gy = gc * (2.f * dy);
                    float gw = (w >= 1e-6f) ? (gc * (2.f * dw)) / (2.f * sw) : 0.f;
                    float gh = (h >= 1e-6f) ? (gc * (2.f * dh)) / (2.f * shh) : 0.f;
                    const float gi = -(inv_n * (2.f * dc));
                    const float g_union = -gi * inter / (U * U);
                    const float g_inter = gi / U - g_union;
                    gw += g_union * h;
                    gh += g_union * w;
                    const float g_rawx = (rawx >= 0.f) ? g_inter * ih : 0.f;
                    const float g_rawy = (rawy >= 0.f) ? g_inter * iw : 0.f;
                    const float gpx2 = g_rawx * g_min_mask(px2, tx2), gpx1 = -g_rawx * g_max_mask(px1, tx1);
                    const float gpy2 = g_rawy * g_min_mask(py2, ty2), gpy1 = -g_rawy * g_max_mask(py1, ty1);
                    gx += gpx1 + gpx2;
                    gy += gpy1 + gpy2;
                    gw += (gpx2 - gpx1) * 0.5f;
                    gh += (gpy2 - gpy1) * 0.5f;
                    gen_st(grad, o + b * 5, gx);
                    gen_st(grad, o + b * 5 + 1, gy);
                    gen_st(grad, o + b * 5 + 2, gw);
                    gen_st(grad, o + b * 5 + 3, gh);
                    gen_st(grad, o + b * 5 + 4, inv_n * (2.f * dc));
                }
            }
            // class term (loss.py:156-157)
            for (int i = 0; i < C; ++i) {
                const float d = gen_ld(pred, o + B * 5 + i) - __ldg(tgt + o + B * 5 + i);
                acc_cls += d * d;
                if (grad != nullptr) gen_st(grad, o + B * 5 + i, inv_n * (2.f * d));
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        acc_coord += __shfl_down_sync(0xffffffffu, acc_coord, off);
        acc_obj += __shfl_down_sync(0xffffffffu, acc_obj, off);
        acc_noobj += __shfl_down_sync(0xffffffffu, acc_noobj, off);
        acc_cls += __shfl_down_sync(0xffffffffu, acc_cls, off);
    }
    if ((threadIdx.x & 31) == 0) {
        s_red[threadIdx.x >> 5][0] = acc_coord; s_red[threadIdx.x >> 5][1] = acc_obj;
        s_red[threadIdx.x >> 5][2] = acc_noobj; s_red[threadIdx.x >> 5][3] = acc_cls;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        double s = 0.0;
        for (int w = 0; w < kGenLossThreads / 32; ++w) s += s_red[w][threadIdx.x];
        partials[(size_t)blockIdx.x * 4 + threadIdx.x] = s;
    }
}

// one CTA folds the per-CTA partials in a fixed order (deterministic) and writes the five loss values
__global__ void __launch_bounds__(256)
gen_loss_fold(const double *__restrict__ partials, unsigned n_blocks, float lambda_coord, float lambda_noobj, float inv_n,
              float *__restrict__ parts) {
    __shared__ double s_fin[256][4];
    double a[4] = {0.0, 0.0, 0.0, 0.0};
    for (unsigned b = threadIdx.x; b < n_blocks; b += 256)
        for (int k = 0; k < 4; ++k) a[k] += partials[(size_t)b * 4 + k];
    for (int k = 0; k < 4; ++k) s_fin[threadIdx.x][k] = a[k];
    __syncthreads();
    for (int st = 128; st > 0; st >>= 1) {
        if ((int)threadIdx.x < st)
            for (int k = 0; k < 4; ++k) s_fin[threadIdx.x][k] += s_fin[threadIdx.x + st][k];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double coord = (double)lambda_coord * s_fin[0][0], cobj = s_fin[0][1];
        const double noobj = (double)lambda_noobj * s_fin[0][2], cls = s_fin[0][3], invn = (double)inv_n;
        parts[0] = (float)((coord + cobj + noobj + cls) * invn);
        parts[1] = (float)(coord * invn);
        parts[2] = (float)(cobj * invn);
        parts[3] = (float)(noobj * invn);
        parts[4] = (float)(cls * invn);
    }
}

// ---------------------------------------------------------------- host launchers (called from the entry points)
static GenShape gen_shape(int S, int B, int C) {
    GenShape sh;
    sh.S = S; sh.B = B; sh.C = C; sh.D = B * 5 + C; sh.K = S * S * B;
    return sh;
}

int generic_decode_nms(const float *pred, int N, int S, int B, int C, double conf_thr, double nms_thr, int iou_mode,
                       int order, int32_t *n_keep, uint8_t *keep_k, uint8_t *keep_cls, double *keep_conf,
                       double *keep_box, int32_t *n_cand, uint8_t *bad_range, cudaStream_t st) {
    const GenShape sh = gen_shape(S, B, C);
    const int grid = (N + 63) / 64;
    if (iou_mode == YH_IOU_METRICS)
        gen_decode_nms_kernel<YH_IOU_METRICS><<<grid, 64, 0, st>>>(pred, N, sh, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls,
                                                                   keep_conf, keep_box, n_cand, bad_range);
    else
        gen_decode_nms_kernel<YH_IOU_INFER><<<grid, 64, 0, st>>>(pred, N, sh, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls,
                                                                 keep_conf, keep_box, n_cand, bad_range);
    return check_launch("yh_decode_nms (generic shape)");
}

int generic_map_update(const float *pred, const float *tgt, int N, int S, int B, int C, double conf_thr, double nms_thr,
                       const double *thresholds_host, int T, int32_t *n_keep, uint64_t *rec_key, uint64_t *rec_flags,
                       uint8_t *rec_k, void *rec_box, int32_t *n_gt, uint8_t *gt_cell, uint8_t *gt_cls, float *gt_raw,
                       int32_t *gt_count, cudaStream_t st) {
    GenThresholds th;
    th.n = T;
    for (int i = 0; i < YH_MAX_THRESHOLDS; ++i) th.t[i] = i < T ? thresholds_host[i] : 2.0;
    gen_map_update_kernel<<<(N + 63) / 64, 64, 0, st>>>(pred, tgt, N, gen_shape(S, B, C), conf_thr, nms_thr, th, n_keep, rec_key,
                                                        rec_flags, rec_k, (uint64_t *)rec_box, n_gt, gt_cell, gt_cls,
                                                        (float4 *)gt_raw, gt_count);
    return check_launch("yh_map_update (generic shape)");
}

int generic_parse_gt(const float *tgt, int N, int S, int B, int C, int32_t *n_gt, uint8_t *gt_cls, double *gt_box,
                     cudaStream_t st) {
    gen_parse_gt_kernel<<<(N + 63) / 64, 64, 0, st>>>(tgt, N, gen_shape(S, B, C), n_gt, gt_cls, gt_box);
    return check_launch("yh_parse_gt (generic shape)");
}

size_t generic_loss_workspace_bytes(int N, int S) {
    const size_t cells = (size_t)N * S * S;
    return 256 + ((cells + kGenLossThreads - 1) / kGenLossThreads) * 4 * sizeof(double);
}

int generic_loss(const void *pred, int pred_dtype, const float *tgt, int N, int S, int B, int C, float lambda_coord,
                 float lambda_noobj, float *parts, void *grad, void *workspace, cudaStream_t st) {
    const long long cells = (long long)N * S * S;
    const unsigned blocks = (unsigned)((cells + kGenLossThreads - 1) / kGenLossThreads);
    char *ws = (char *)workspace;
    ws += (256 - ((uintptr_t)ws & 255)) & 255;
    double *partials = (double *)ws;
    const float inv_n = 1.0f / (float)N;
    if (pred_dtype == YH_F32)
        gen_loss_kernel<float><<<blocks, kGenLossThreads, 0, st>>>((const float *)pred, tgt, (float *)grad, cells, B, C, lambda_coord,
                                                                   lambda_noobj, inv_n, partials);
    else if (pred_dtype == YH_F16)
        gen_loss_kernel<__half><<<blocks, kGenLossThreads, 0, st>>>((const __half *)pred, tgt, (__half *)grad, cells, B, C,
                                                                    lambda_coord, lambda_noobj, inv_n, partials);
    else
        gen_loss_kernel<__nv_bfloat16><<<blocks, kGenLossThreads, 0, st>>>((const __nv_bfloat16 *)pred, tgt, (__nv_bfloat16 *)grad,
                                                                           cells, B, C, lambda_coord, lambda_noobj, inv_n, partials);
    gen_loss_fold<<<1, 256, 0, st>>>(partials, blocks, lambda_coord, lambda_noobj, inv_n, parts);
    return check_launch("yh_loss_fwd_bwd (generic shape)", 2);
}

}  // namespace yh
