// common.cuh -- shared helpers for libyolohead.so (sm_100a only).
//
// Arithmetic rules (SURVEY.md section 7 "hard parts"): the reference computes the
// decode / NMS / matching geometry in Python doubles, one rounding per operator.
// Every TU is compiled with -fmad=false and the fp64 helpers below use the
// explicit round-to-nearest intrinsics, so no multiply-add is ever contracted.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/yolohead.h"

namespace yh {

constexpr int kS = 7;
constexpr int kB = 2;
constexpr int kC = 20;
constexpr int kD = kB * 5 + kC;       // 30 floats per cell
constexpr int kCells = kS * kS;       // 49
constexpr int kCand = kCells * kB;    // 98 candidates per image
constexpr int kImgFloats = kCells * kD;  // 1470 floats = 5880 B per image

void set_error(const char *fmt, ...);
// A zeroed device counter for one launch's dynamic image schedule (api.cu): slots are handed
// out round-robin from a ring, so launches in flight on different streams never share one,
// and the slot is cleared on the launch's own stream.
// Returns nullptr (error set) when a graph capture has used up its never-recycled slots.
unsigned *fresh_ticket(cudaStream_t stream);
constexpr int kMaxDevices = 64;
// true the first time it is called for (flag array of kMaxDevices, current device)
bool first_use_on_device(std::atomic<bool> *seen);
int current_device();
int check_launch(const char *what, int kernels_launched = 1);

inline bool shape_ok(int S, int B, int C) { return S == kS && B == kB && C == kC; }

// Every other head shape runs the one-thread-per-image kernels of generic.cu (same results, no tuning).
// with_classes_in_flags: the metric / inference entry points (C <= 32); the loss takes any C.
bool generic_shape_ok(int S, int B, int C, bool with_classes_in_flags);
int generic_decode_nms(const float *pred, int N, int S, int B, int C, double conf_thr, double nms_thr, int iou_mode,
                       int order, int32_t *n_keep, uint8_t *keep_k, uint8_t *keep_cls, double *keep_conf,
                       double *keep_box, int32_t *n_cand, uint8_t *bad_range, cudaStream_t st);
int generic_map_update(const float *pred, const float *tgt, int N, int S, int B, int C, double conf_thr, double nms_thr,
                       const double *thresholds_host, int T, int32_t *n_keep, uint64_t *rec_key, uint64_t *rec_flags,
                       uint8_t *rec_k, void *rec_box, int32_t *n_gt, uint8_t *gt_cell, uint8_t *gt_cls, float *gt_raw,
                       int32_t *gt_count, cudaStream_t st);
int generic_parse_gt(const float *tgt, int N, int S, int B, int C, int32_t *n_gt, uint8_t *gt_cls, double *gt_box,
                     cudaStream_t st);
size_t generic_loss_workspace_bytes(int N, int S);
int generic_loss(const void *pred, int pred_dtype, const float *tgt, int N, int S, int B, int C, float lambda_coord,
                 float lambda_noobj, float *parts, void *grad, void *workspace, cudaStream_t st);

// ---- Python float semantics -------------------------------------------------
// builtin max(a, b) / min(a, b): first argument unless the second is strictly
// greater / smaller (so a NaN second argument never wins).
__device__ __forceinline__ double py_max(double a, double b) { return (b > a) ? b : a; }
__device__ __forceinline__ double py_min(double a, double b) { return (b < a) ? b : a; }

struct Corners {
    double x1, y1, x2, y2, area;
};

// (x, y, w, h) -> corners and area exactly as metrics.py:317-320,333-334 and
// schemas.py:25-28,55 build them (x - w / 2 etc., w * h).
__device__ __forceinline__ Corners corners_of(double x, double y, double w, double h) {
    Corners c;
    double hw = __dmul_rn(w, 0.5), hh = __dmul_rn(h, 0.5);  // w / 2 is exact either way
    c.x1 = __dsub_rn(x, hw);
    c.y1 = __dsub_rn(y, hh);
    c.x2 = __dadd_rn(x, hw);
    c.y2 = __dadd_rn(y, hh);
    c.area = __dmul_rn(w, h);
    return c;
}

// Intersection area, argument order as the reference (first box = a).
__device__ __forceinline__ double inter_area(const Corners &a, const Corners &b) {
    double ix1 = py_max(a.x1, b.x1), iy1 = py_max(a.y1, b.y1);
    double ix2 = py_min(a.x2, b.x2), iy2 = py_min(a.y2, b.y2);
    double dx = py_max(0.0, __dsub_rn(ix2, ix1));
    double dy = py_max(0.0, __dsub_rn(iy2, iy1));
    return __dmul_rn(dx, dy);
}

// Denominator of the IoU: metrics.py:335 (a1 + a2 - inter) or inference.py:248
// (a1 + a2 - inter + 1e-6).
template <int MODE>
__device__ __forceinline__ double iou_denominator(const Corners &a, const Corners &b, double inter) {
    double u = __dsub_rn(__dadd_rn(a.area, b.area), inter);
    if (MODE == YH_IOU_INFER) u = __dadd_rn(u, 1e-6);
    return u;
}

// IEEE fp64 division, kept out of line: the inlined sequence is ~150 instructions
static __device__ __noinline__ double div_exact(double a, double b) { return __ddiv_rn(a, b); }
static __device__ __forceinline__ bool quotient_below(double inter, double u, double thr) {
    return div_exact(inter, u) < thr;
}

template <int MODE>
__device__ __forceinline__ double iou_exact(const Corners &a, const Corners &b) {
    double inter = inter_area(a, b);
    double u = iou_denominator<MODE>(a, b, inter);
    if (MODE == YH_IOU_METRICS && u == 0.0) return 0.0;
    return div_exact(inter, u);
}

// "is IoU(a, b) < thr" -- the keep test of NMS (metrics.py:293, inference.py:314).
// Decides without the division whenever inter and thr*denominator are further
// apart than any rounding could bridge; falls back to the exact quotient
// otherwise, so the outcome is always the reference's.
template <int MODE>
__device__ __forceinline__ bool iou_below(const Corners &a, const Corners &b, double thr) {
    double inter = inter_area(a, b);
    double u = iou_denominator<MODE>(a, b, inter);
    if (MODE == YH_IOU_METRICS && u == 0.0) return 0.0 < thr;
    if (u > 0.0 && thr > 0.0 && inter >= 0.0 && u < 1e300 && inter < 1e300) {
        double t = __dmul_rn(thr, u);
        if (t > 1e-280) {
            double slack = __dmul_rn(t, 1e-12);
            if (inter < __dsub_rn(t, slack)) return true;
            if (inter > __dadd_rn(t, slack)) return false;
        }
    }
    return quotient_below(inter, u, thr);
}

// Monotone map double -> uint64 (ascending).  -0.0 sorts just below +0.0.
__device__ __forceinline__ uint64_t orderable(double v) {
    uint64_t b = (uint64_t)__double_as_longlong(v);
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double from_orderable(uint64_t k) {
    uint64_t b = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

// max that returns NaN when either operand is NaN (PTX max.NaN.f32, sm_80+)
__device__ __forceinline__ float max_nan(float a, float b) {
    float d;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}

// torch.argmax over the C class scores of one cell: first maximum, NaN maximal
// (the first NaN wins).  Two short passes instead of a compare-and-branch chain:
// the maximum, then the first index that equals it.
__device__ __forceinline__ int argmax_first(const float *v, int n, float &best_v) {
    float m = v[0];
#pragma unroll
    for (int i = 1; i < n; ++i) m = max_nan(m, v[i]);
    // first index that holds the maximum; a NaN maximum matches the first NaN.  Fully
    // unrolled with constant indices so v[] stays in registers.
    int best = 0;
    if (m != m) {
#pragma unroll
        for (int i = n - 1; i >= 0; --i) best = (v[i] != v[i]) ? i : best;
    } else {
#pragma unroll
        for (int i = n - 1; i >= 0; --i) best = (v[i] == m) ? i : best;
    }
    best_v = m;
    return best;
}

// Class of one staged prediction cell (cell = pointer to its kD floats in shared memory,
// 8-byte aligned; conf0 / conf1 = its two box confidences): best_v = the largest class score; the return value is its first index
// -- but only when a box of the cell can pass the confidence threshold (otherwise the
// class is never looked at and the index search is skipped; a NaN maximum makes both
// confidences NaN, which never pass).  Out of line: one copy for all call sites keeps
// the per-image kernels inside the instruction cache.
static __device__ __noinline__ int cell_class_shared(const float *cell, double conf_thr, float conf0, float conf1,
                                                     float &best_v) {
    const unsigned addr = (unsigned)__cvta_generic_to_shared(cell + kB * 5);
    float v[kC];
#pragma unroll
    for (int i = 0; i < kC / 2; ++i)
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v[2 * i]), "=f"(v[2 * i + 1]) : "r"(addr + 8 * i) : "memory");
    float m = v[0];
#pragma unroll
    for (int i = 1; i < kC; ++i) m = max_nan(m, v[i]);
    best_v = m;
    const double prob = (double)m;
    const bool need = __dmul_rn((double)conf0, prob) > conf_thr || __dmul_rn((double)conf1, prob) > conf_thr;
    int best = 0;
    if (need) {
#pragma unroll
        for (int i = kC - 1; i >= 0; --i) best = (v[i] == m) ? i : best;
    }
    return best;
}
static __device__ __noinline__ int cell_class_global(const float *cell) {
    float cls_p[kC];
#pragma unroll
    for (int i = 0; i < kC / 2; ++i) {
        const float2 v = __ldg(reinterpret_cast<const float2 *>(cell + kB * 5) + i);
        cls_p[2 * i] = v.x;
        cls_p[2 * i + 1] = v.y;
    }
    float pv;
    return argmax_first(cls_p, kC, pv);
}

}  // namespace yh
