// Host-side check of yolo_v1_b200/csrc/iou_filter.cuh (test infrastructure).
//
// The prefilter may only answer when the reference's own fp64 arithmetic (metrics.py:317-341,
// inference.py:237-248, restated below with one rounding per operator) gives the same answer.
// This program draws boxes the way the kernels see them (cell column/row + fp32 offsets, fp32
// sizes), including pairs steered onto the threshold, degenerate / negative sizes and huge or
// non-finite values, and counts wrong answers (must be 0) and the share of pairs decided.
//
//   g++ -O2 -ffp-contract=off -o iou_filter_check iou_filter_check.cpp && ./iou_filter_check [pairs] [seed]
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "../../yolo_v1_b200/csrc/iou_filter.cuh"

using namespace yh;

static uint64_t s_state = 0x9E3779B97F4A7C15ull;
static inline uint64_t rnd64() {
    uint64_t z = (s_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline float u01() { return (float)(rnd64() >> 40) * (1.0f / 16777216.0f); }
static inline double py_max(double a, double b) { return (b > a) ? b : a; }
static inline double py_min(double a, double b) { return (b < a) ? b : a; }

struct Raw {
    int col, row;
    float x, y, w, h;
};

// the reference's doubles
struct Ref {
    double x1, y1, x2, y2, area;
};
static Ref ref_box(const Raw &r, int S) {
    volatile double xa = ((double)r.col + (double)r.x) / (double)S;
    volatile double ya = ((double)r.row + (double)r.y) / (double)S;
    Ref b;
    b.x1 = xa - (double)r.w / 2;
    b.y1 = ya - (double)r.h / 2;
    b.x2 = xa + (double)r.w / 2;
    b.y2 = ya + (double)r.h / 2;
    b.area = (double)r.w * (double)r.h;
    return b;
}
// IoU exactly as metrics.py:322-341 (eps = false) or inference.py:237-248 (eps = true)
static double ref_iou(const Ref &a, const Ref &b, bool eps) {
    double ix1 = py_max(a.x1, b.x1), iy1 = py_max(a.y1, b.y1);
    double ix2 = py_min(a.x2, b.x2), iy2 = py_min(a.y2, b.y2);
    double inter = py_max(0.0, ix2 - ix1) * py_max(0.0, iy2 - iy1);
    double u = a.area + b.area - inter;
    if (eps) return inter / (u + 1e-6);
    return u == 0.0 ? 0.0 : inter / u;
}

static BoxF fast_box(const Raw &r, int S) {
    const float inv_s = 1.0f / (float)S;
    return make_box_f(abs_coord_f(r.col, r.x, inv_s), abs_coord_f(r.row, r.y, inv_s), r.w, r.h);
}

static Raw draw(int S, int kind) {
    Raw r;
    r.col = (int)(rnd64() % S);
    r.row = (int)(rnd64() % S);
    switch (kind) {
    case 0:  // uniform head output
        r.x = u01(); r.y = u01(); r.w = u01(); r.h = u01();
        break;
    case 1:  // raw head: signed, beyond [0,1]
        r.x = 3.f * (u01() - 0.5f); r.y = 3.f * (u01() - 0.5f); r.w = 2.f * (u01() - 0.3f); r.h = 2.f * (u01() - 0.3f);
        break;
    case 2:  // quantised (ties)
        r.x = floorf(u01() * 9) / 8; r.y = floorf(u01() * 9) / 8; r.w = floorf(u01() * 9) / 8; r.h = floorf(u01() * 9) / 8;
        break;
    case 3:  // tiny sizes
        r.x = u01(); r.y = u01(); r.w = u01() * 1e-4f; r.h = u01() * 1e-4f;
        break;
    case 4: {  // wild magnitudes
        const float sc[6] = {1e-20f, 1e-6f, 1.f, 1e4f, 1e12f, 1e25f};
        r.x = u01() * sc[rnd64() % 6]; r.y = u01() * sc[rnd64() % 6]; r.w = u01() * sc[rnd64() % 6]; r.h = u01() * sc[rnd64() % 6];
        break;
    }
    default:
        r.x = u01(); r.y = u01(); r.w = 0.05f + 0.75f * u01(); r.h = 0.05f + 0.75f * u01();
    }
    return r;
}

int main(int argc, char **argv) {
    const long pairs = argc > 1 ? atol(argv[1]) : 2000000;
    if (argc > 2) s_state ^= (uint64_t)atoll(argv[2]) * 0xD1342543DE82EF95ull;
    const double thrs[] = {0.4, 0.5, 0.45, 0.3, 0.75, 0.05, 0.95, 1.0, 0.001, 3.0};
    long wrong = 0, decided = 0, total = 0, near_decided = 0, near_total = 0, ival_bad = 0, ival_total = 0;
    for (long it = 0; it < pairs; ++it) {
        const int S = 7;
        const int kind = (int)(rnd64() % 6);
        Raw ra = draw(S, kind), rb = draw(S, kind);
        const double thr = thrs[rnd64() % 10];
        const bool eps = rnd64() & 1;
        bool steered = false;
        if ((rnd64() & 3) == 0 && kind != 4) {
            // steer the pair onto the threshold: make b a shifted copy of a and bisect the shift
            rb = ra;
            rb.col = ra.col; rb.row = ra.row;
            float lo = 0.f, hi = fabsf(ra.w) * (float)S + 1e-3f;
            const Ref A = ref_box(ra, S);
            for (int k = 0; k < 40; ++k) {
                const float mid = 0.5f * (lo + hi);
                rb.x = ra.x + mid;
                const double v = ref_iou(A, ref_box(rb, S), eps);
                if (v >= thr) lo = mid; else hi = mid;
            }
            rb.x = ra.x + ((rnd64() & 1) ? lo : hi);
            // a few ulps either way
            int ul = (int)(rnd64() % 9) - 4;
            for (; ul > 0; --ul) rb.x = nextafterf(rb.x, 1e30f);
            for (; ul < 0; ++ul) rb.x = nextafterf(rb.x, -1e30f);
            steered = true;
        }
        const Ref A = ref_box(ra, S), B = ref_box(rb, S);
        const double v = ref_iou(A, B, eps);
        const bool below = v < thr;
        const BoxF fa = fast_box(ra, S), fb = fast_box(rb, S);
        float M = box_magnitude(fa);
        const float mb = box_magnitude(fb);
        M = (mb > M || mb != mb) ? mb : M;
        // an image-wide bound is at least this pair's bound; sometimes take a larger one
        if (rnd64() & 1) M = M * 2.f + 1.f;
        const FilterParams p = filter_params(thr);
        int verdict = 0;
        if (p.thr_ok && M < kFilterMaxCoord) {
            const float E = filter_margin(M);
            verdict = eps ? iou_below_fast<true>(fa, fb, E, p) : iou_below_fast<false>(fa, fb, E, p);
            if (!eps) {
                float lo = 0.f, hi = 0.f;
                const int ok = iou_interval(fa, fb, E, lo, hi);
                if (ok) {
                    ++ival_total;
                    const bool good = ok == 2 ? (v == 0.0) : ((double)lo <= v && v <= (double)hi);
                    if (!good) {
                        if (ival_bad < 10) fprintf(stderr, "interval(%d): v=%.17g not in [%.9g, %.9g] kind=%d\n", ok, v, lo, hi, kind);
                        ++ival_bad;
                    }
                }
            }
        }
        ++total;
        if (steered) ++near_total;
        if (verdict != 0) {
            ++decided;
            if (steered) ++near_decided;
            if ((verdict > 0) != below) {
                if (wrong < 10)
                    fprintf(stderr, "WRONG verdict=%d iou=%.17g thr=%.17g eps=%d kind=%d\n", verdict, v, thr, (int)eps, kind);
                ++wrong;
            }
        }
    }
    printf("{\"pairs\": %ld, \"decided\": %ld, \"wrong\": %ld, \"steered\": %ld, \"steered_decided\": %ld, "
           "\"intervals\": %ld, \"interval_misses\": %ld}\n",
           total, decided, wrong, near_total, near_decided, ival_total, ival_bad);
    return (wrong || ival_bad) ? 1 : 0;
}
