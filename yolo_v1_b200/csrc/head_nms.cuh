// head_nms.cuh -- ordering, bitmask NMS and output ordering on top of the
// candidate table built by head_image.cuh (one warp per image).
#pragma once

#include "head_image.cuh"

namespace yh {

__device__ __forceinline__ Corners cand_corners(const WarpWindow &w, int k) {
    Corners c;
    c.x1 = w.cand.x1[k]; c.y1 = w.cand.y1[k]; c.x2 = w.cand.x2[k]; c.y2 = w.cand.y2[k];
    c.area = w.cand.area[k];
    return c;
}

// ---- ordering --------------------------------------------------------------------
// Sorted position p = (class asc, confidence desc, candidate asc): fills w.order[p] = k
// and w.cls_start[c]; returns n.  Lane owns candidates k = lane + 32*s.
__device__ __forceinline__ int order_candidates(WarpWindow &w, const Mask128 &valid, int lane) {
    const int n = valid.count();
    if (lane < kC) w.count[lane] = 0;
    __syncwarp();
#pragma unroll 1
    for (int s = 0; s < kSlots; ++s)
        if ((valid.word(s) >> lane) & 1u) atomicAdd(&w.count[w.cand.cls[lane + 32 * s]], 1);
    __syncwarp();
    // exclusive scan of the class histogram (20 bins, in-warp)
    const int cnt = (lane < kC) ? w.count[lane] : 0;
    int inc = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        int v = __shfl_up_sync(kFull, inc, off);
        if (lane >= off) inc += v;
    }
    __syncwarp();
    if (lane < kC) {
        w.cls_start[lane] = (uint8_t)(inc - cnt);
        w.count[lane] = inc - cnt;  // becomes the running cursor
    }
    if (lane == 0) w.cls_start[kC] = (uint8_t)n;
    __syncwarp();
#pragma unroll 1
    for (int s = 0; s < kSlots; ++s)
        if ((valid.word(s) >> lane) & 1u) {
            const int k = lane + 32 * s;
            const int at = atomicAdd(&w.count[w.cand.cls[k]], 1);
            w.member[at] = (uint8_t)k;
        }
    __syncwarp();
    // rank inside the class: members that beat me (higher confidence, or equal confidence
    // and earlier candidate -- the stable order of sorted(..., reverse=True))
#pragma unroll 1
    for (int s = 0; s < kSlots; ++s)
        if ((valid.word(s) >> lane) & 1u) {
            const int k = lane + 32 * s;
            const int c = w.cand.cls[k];
            const int lo = w.cls_start[c], hi = w.cls_start[c + 1];
            const double mine = w.cand.conf[k];
            int r = 0;
            for (int q = lo; q < hi; ++q) {
                const int m = w.member[q];
                const double other = w.cand.conf[m];
                r += (other > mine) || (other == mine && m < k);
            }
            w.order[lo + r] = (uint8_t)k;
        }
    __syncwarp();
    return n;
}

// ---- NMS ---------------------------------------------------------------------------
// Greedy per-class NMS over the sorted candidates.  Each round, every class's best
// still-undecided candidate is kept and all later undecided candidates of that
// class are tested against it; rounds = max kept per class.  Returns the keep mask
// over sorted positions.  A NaN threshold means "decode only": keep everything.
template <int MODE>
static __device__ __noinline__ Mask128 nms_rounds(const WarpWindow &w, int n, double nms_thr, int lane) {
    Mask128 rem, keep = {0, 0, 0, 0};
#pragma unroll
    for (int s = 0; s < kSlots; ++s) {
        const int left = n - 32 * s;
        rem.set_word(s, left >= 32 ? kFull : (left > 0 ? ((1u << left) - 1u) : 0u));
    }
    if (nms_thr != nms_thr) return rem;
    while (rem.any()) {
        Mask128 still = {0, 0, 0, 0};
#pragma unroll 1
        for (int s = 0; s < kSlots; ++s) {
            const unsigned wd = rem.word(s);
            if (wd == 0) continue;  // warp-uniform
            const int p = lane + 32 * s;
            bool mine = (wd >> lane) & 1u, kept = false;
            if (mine) {
                const int k = w.order[p];
                const int head = first_from(rem, w.cls_start[w.cand.cls[k]]);
                if (head == p) {
                    kept = true;
                    mine = false;
                } else {
                    // reference argument order: iou(current, other)
                    mine = iou_below<MODE>(cand_corners(w, w.order[head]), cand_corners(w, k), nms_thr);
                }
            }
            still.set_word(s, __ballot_sync(kFull, mine));
            keep.set_word(s, keep.word(s) | __ballot_sync(kFull, kept));
        }
        rem = still;
    }
    return keep;
}

// Overlay of `stage` between decode and the next staging: NMS scratch.
struct NmsOverlay {
    unsigned long long sup[kCand];  // bit i: the i-th member of my class (i < my rank) overlaps me
    uint8_t kept[kPad];
};
static_assert(sizeof(NmsOverlay) <= sizeof(float2) * (kStagePairs + 1), "overlay too large");

// Same result as nms_rounds, organised for the common case of many small classes:
//   1. every candidate tests itself against ALL earlier members of its class
//      (independent work, no rounds) and records the overlaps as a bitmask;
//   2. one lane per class replays the greedy scan over those bitmasks.
// Classes with more than 64 members fall back to nms_rounds.
template <int MODE>
__device__ __forceinline__ Mask128 nms_pairwise(WarpWindow &w, int n, double nms_thr, int lane) {
    const int cc = lane < kC ? lane : 0;
    const int clo = w.cls_start[cc], chi = w.cls_start[cc + 1];
    const bool big = (lane < kC) && (chi - clo > 64);
    if (__any_sync(kFull, big) || nms_thr != nms_thr) return nms_rounds<MODE>(w, n, nms_thr, lane);
    NmsOverlay &no = *reinterpret_cast<NmsOverlay *>(w.stage);
#pragma unroll 1
    for (int p = lane; p < n; p += 32) {
        const int k = w.order[p];
        const int lo = w.cls_start[w.cand.cls[k]];
        const Corners me = cand_corners(w, k);
        unsigned long long sup = 0;
        for (int e = lo; e < p; ++e)  // reference argument order: iou(current, other)
            if (!iou_below<MODE>(cand_corners(w, w.order[e]), me, nms_thr)) sup |= 1ull << (e - lo);
        no.sup[p] = sup;
    }
    __syncwarp();
    if (lane < kC) {
        unsigned long long km = 0;
        for (int i = 0; i < chi - clo; ++i) {
            const bool kp = (no.sup[clo + i] & km) == 0;
            km |= (unsigned long long)kp << i;
            no.kept[clo + i] = kp;
        }
    }
    __syncwarp();
    Mask128 keep;
    keep.w0 = __ballot_sync(kFull, lane < n && no.kept[lane]);
    keep.w1 = __ballot_sync(kFull, lane + 32 < n && no.kept[lane + 32]);
    keep.w2 = __ballot_sync(kFull, lane + 64 < n && no.kept[lane + 64]);
    keep.w3 = __ballot_sync(kFull, lane + 96 < n && no.kept[lane + 96]);
    __syncwarp();
    return keep;
}

// Output offset of every class group for the class-grouped order of
// mAPMetric._apply_nms: groups appear in the order of their best member in the
// global (confidence desc, candidate asc) order.  Fills w.cls_off.
__device__ __forceinline__ void class_group_offsets(WarpWindow &w, const Mask128 &keep, int lane) {
    const int c = lane < kC ? lane : 0;
    const int lo = w.cls_start[c], hi = w.cls_start[c + 1];
    const bool present = (lane < kC) && hi > lo;
    const int hk = present ? w.order[lo] : 0;
    const double hc = present ? w.cand.conf[hk] : 0.0;
    const int kbase = present ? count_below(keep, lo) : 0;
    const int kc = present ? count_below(keep, hi) - kbase : 0;
    const unsigned have = __ballot_sync(kFull, present);
    int off = 0;
    for (unsigned m = have; m; m &= m - 1) {
        const int o = __ffs(m) - 1;
        const double oc = __shfl_sync(kFull, hc, o);
        const int ok = __shfl_sync(kFull, hk, o);
        const int on = __shfl_sync(kFull, kc, o);
        if ((oc > hc) || (oc == hc && ok < hk)) off += on;
    }
    if (lane < kC) {
        w.cls_off[lane] = (uint8_t)off;
        w.cls_kbase[lane] = (uint8_t)kbase;
    }
    __syncwarp();
}

}  // namespace yh
