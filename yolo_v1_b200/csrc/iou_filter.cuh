// iou_filter.cuh -- fp32 interval prefilter for the "IoU < threshold" / "IoU >= threshold" tests.
//
// The reference decides every NMS keep test and every TP/FP threshold test on an IoU computed in
// Python doubles, one rounding per operator (metrics.py:317-341, inference.py:237-248).  Repeating
// that arithmetic for every same-class pair costs ~25 dependent fp64 instructions per pair, and the
// B200's fp64 pipe issues a warp instruction every other cycle.  Almost every pair is nowhere near
// the threshold, so the test is first made on fp32 INTERVALS that provably contain the reference's
// fp64 intersection and union; only when the two intervals do not separate the quotient from the
// threshold does the caller fall back to the exact fp64 sequence.  The outcome is therefore always
// the reference's -- the filter only ever answers when every value the reference could have computed
// gives the same answer.
//
// This header is plain C++ (host + device): tests/test_iou_filter.py compiles it with g++ and checks
// the claim against an fp64 restatement of the reference on adversarial inputs.
//
// Error model (fourth mitigation).  A box is (xa, ya, w, h): w, h fp32 inputs,
// xa = (col + x) / S in the reference's doubles.  The fp32 corner table holds
//     xaf = RN32(RN32(col + x) * RN32(1/S))          |xaf - xa_ref| <= 3 * 2^-24 |xa|
//     x1f = RN32(xaf - w/2), x2f = RN32(xaf + w/2)   |x1f - x1_ref| <= 1.25 * 2^-22 * M
// with M >= every |corner| of the image.  Intersections: dx_f = RN32(min(x2f) - max(x1f)) is within
// 3 * 2^-22 * M of the reference's dx; the filter uses E = 2^-19 * M + 1e-30 (more than 2x that).
// Areas: af = RN32(w * h) (the reference's product is exact), |af - a| <= 2^-24 |a|; the union
// interval is widened by 2^-20 (|a1| + |a2| + inter_hi) (+ 1e-12 for the inference flavour's 1e-6),
// four times what the fp32 evaluation can lose.  The threshold is widened by 2^-20 either way, which
// covers the fp32 product roundings and the final rounding of the reference's quotient.
// Fast answers are only given when the image has no NaN / infinite / huge coordinate (M < 1e15), the
// union is safely positive (> 1e-30) and the threshold lies in [1e-3, 1e3]; anything else is
// "undecided" and takes the exact path.
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define YH_HD __host__ __device__ __forceinline__
#else
#define YH_HD inline
#endif

namespace yh {

struct BoxF {             // fp32 corners and area of one candidate (16 + 4 bytes in shared memory)
    float x1, y1, x2, y2, area;
};

struct FilterParams {     // per launch (threshold) and per image (E, ok)
    float thr_lo, thr_hi;  // thr * (1 -/+ 2^-20), rounded to float
    bool thr_ok;           // threshold inside the range the error model covers
};

YH_HD FilterParams filter_params(double thr) {
    FilterParams p;
    p.thr_ok = (thr >= 1e-3) && (thr <= 1e3);  // false for NaN
    p.thr_lo = (float)(thr * (1.0 - 9.5367431640625e-07));
    p.thr_hi = (float)(thr * (1.0 + 9.5367431640625e-07));
    return p;
}

// fp32 corner table entry from the raw cell values (cell_coord = column for x, row for y)
YH_HD float abs_coord_f(int cell_coord, float v, float inv_s) { return ((float)cell_coord + v) * inv_s; }

YH_HD BoxF make_box_f(float xaf, float yaf, float w, float h) {
    BoxF b;
    const float hw = 0.5f * w, hh = 0.5f * h;
    b.x1 = xaf - hw;
    b.y1 = yaf - hh;
    b.x2 = xaf + hw;
    b.y2 = yaf + hh;
    b.area = w * h;
    return b;
}

// magnitude bound of one box for the per-image M (NaN if any corner is NaN: callers compare
// `M < kFilterMaxCoord`, which is false for NaN)
YH_HD float box_magnitude(const BoxF &b) {
    // a + b propagates NaN, fmaxf would not
    return fabsf(b.x1) + fabsf(b.y1) + fabsf(b.x2) + fabsf(b.y2);
}
constexpr float kFilterMaxCoord = 1e15f;

// half-width of the interval around an fp32 intersection side, from the image's magnitude bound
YH_HD float filter_margin(float M) { return M * 1.9073486328125e-06f + 1e-30f; }  // 2^-19

// Verdict on "IoU(a, b) < thr" as the reference computes it: +1 certainly below, -1 certainly not
// below (>= thr), 0 undecided.  add_eps: the inference flavour's `+ 1e-6` in the denominator.
// Symmetric in (a, b) except through fp32 rounding, which the intervals absorb.
template <bool ADD_EPS>
YH_HD int iou_below_fast(const BoxF &a, const BoxF &b, float E, const FilterParams &p) {
    const float dx = fminf(a.x2, b.x2) - fmaxf(a.x1, b.x1);
    const float dy = fminf(a.y2, b.y2) - fmaxf(a.y1, b.y1);
    const float dxl = fmaxf(dx - E, 0.f), dxh = fmaxf(dx + E, 0.f);
    const float dyl = fmaxf(dy - E, 0.f), dyh = fmaxf(dy + E, 0.f);
    const float ih = dxh * dyh, il = dxl * dyl;
    const float S = a.area + b.area;
    float aerr = (fabsf(a.area) + fabsf(b.area) + ih) * 9.5367431640625e-07f;  // 2^-20
    float ulo = S - ih, uhi = S - il;
    if (ADD_EPS) {
        aerr += 1e-12f;
        ulo += 1e-6f;
        uhi += 1e-6f;
    }
    if (ih == 0.f) {
        // dx + E <= 0 or dy + E <= 0: the reference's intersection is exactly 0, so its IoU is 0 / u = +-0 (or the
        // metrics flavour's 0.0 for an empty union) -- below every threshold this filter accepts (>= 1e-3), whatever
        // the sign of the union (raw heads emit negative sizes).  The inference flavour divides by u + 1e-6, which
        // must not be 0 (0 / 0 is NaN, and NaN < thr is false).
        if (!ADD_EPS) return 1;
        if (fabsf(S + 1e-6f) > aerr) return 1;
    }
    ulo -= aerr;
    uhi += aerr;
    if (!(ulo > 1e-30f)) return 0;
    if (ih < p.thr_lo * ulo) return 1;
    if (il > p.thr_hi * uhi) return -1;
    return 0;
}

// Interval [lo, hi] that contains the reference's fp64 IoU (quotient rounding included), for the best-ground-truth
// search and the threshold sweep of the TP/FP matching (metrics flavour: no epsilon).  Returns 1 with the interval,
// 2 when the reference's IoU is exactly zero (empty intersection: such a pair never matches, metrics.py:398-403),
// 0 when no statement can be made.
YH_HD int iou_interval(const BoxF &a, const BoxF &b, float E, float &lo, float &hi) {
    const float dx = fminf(a.x2, b.x2) - fmaxf(a.x1, b.x1);
    const float dy = fminf(a.y2, b.y2) - fmaxf(a.y1, b.y1);
    const float dxl = fmaxf(dx - E, 0.f), dxh = fmaxf(dx + E, 0.f);
    const float dyl = fmaxf(dy - E, 0.f), dyh = fmaxf(dy + E, 0.f);
    const float ih = dxh * dyh, il = dxl * dyl;
    if (ih == 0.f) return 2;
    const float S = a.area + b.area;
    const float aerr = (fabsf(a.area) + fabsf(b.area) + ih) * 9.5367431640625e-07f;
    const float ulo = (S - ih) - aerr, uhi = (S - il) + aerr;
    if (!(ulo > 1e-30f)) return 0;
    lo = (il / uhi) * (1.0f - 9.5367431640625e-07f);
    hi = (ih / ulo) * (1.0f + 9.5367431640625e-07f) + 1e-37f;
    return 1;
}

}  // namespace yh
