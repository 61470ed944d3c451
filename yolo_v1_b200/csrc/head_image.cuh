// head_image.cuh -- per-image building blocks of the decode / NMS / matching
// kernels: ONE WARP OWNS ONE IMAGE.
//
// The 98 candidate boxes of an image never leave the SM: the 5,880-byte image is
// staged into the warp's shared-memory window with coalesced 8-byte loads (an
// image start is only 8-byte aligned), decoded into registers (each lane owns
// cells `lane` and `lane+32`, i.e. up to four candidates), ordered by
// (class, confidence desc, candidate asc) with a per-class counting rank, and
// suppressed by a bitmask NMS whose rounds advance every class at once.
//
// Reference semantics reproduced here (paths relative to the reference repo):
//   decode            src/yolo/metrics.py:187-216, src/yolo/inference.py:174-208
//   NMS               src/yolo/metrics.py:274-294, src/yolo/inference.py:302-315
//   ground truths     src/yolo/metrics.py:234-254
#pragma once

#include "common.cuh"

namespace yh {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kStagePairs = kImgFloats / 2;  // 735 float2 per image
constexpr int kSlots = 4;                    // candidates a lane can own

// Shared-memory window of one warp.  `stage` holds the raw image while it is
// being decoded; afterwards the same bytes are reused (see the overlay structs).
struct __align__(16) WarpWindow {
    float2 stage[kStagePairs + 1];  // 5,888 B
    // candidates in sorted order p: class asc, confidence desc, candidate asc
    double sx1[kCand], sy1[kCand], sx2[kCand], sy2[kCand], sarea[kCand], sconf[kCand];
    uint8_t scls[kCand + 6], sk[kCand + 6], sidx[kCand + 6];
    uint8_t cls_start[kC + 4];  // first sorted position of each class; [kC] = n
    uint8_t cls_off[kC + 4];    // output offset of each class group (class-grouped order)
};

// Overlay of `stage` once every lane holds its decoded candidates in registers.
struct RankOverlay {
    double conf[kCand];        // by compact index (= ascending candidate index)
    uint8_t cls[kCand + 6];
    uint8_t member[kCand + 6];  // class-segmented member lists (compact indices)
    int cursor[kC];
    int count[kC];
};
static_assert(sizeof(RankOverlay) <= sizeof(float2) * (kStagePairs + 1), "overlay too large");

// Overlay of `stage` after the target image has been parsed into registers.
struct MatchOverlay {
    double gx1[kCells], gy1[kCells], gx2[kCells], gy2[kCells], garea[kCells];  // 1,960 B
    uint8_t gcls[kCells + 7], gbucket[kCells + 7];                              // 112 B
    double biou[4][kCand];                                                      // 3,136 B
    uint8_t bgt[4][kCand + 6];                                                  // 416 B
};
static_assert(sizeof(MatchOverlay) <= sizeof(float2) * (kStagePairs + 1), "overlay too large");

// ---- staging ------------------------------------------------------------------
__device__ __forceinline__ void stage_image(WarpWindow &w, const float *__restrict__ img, int lane) {
    const float2 *src = reinterpret_cast<const float2 *>(img);
#pragma unroll
    for (int i = 0; i < (kStagePairs + 31) / 32; ++i) {
        int q = lane + 32 * i;
        if (q < kStagePairs) w.stage[q] = __ldg(src + q);
    }
    __syncwarp();
}

// ---- decode ---------------------------------------------------------------------
struct LaneCands {      // the (up to) four candidates a lane decodes: slot = cellslot*2 + b
    double conf[kSlots];
    float x[kSlots], y[kSlots], w[kSlots], h[kSlots];
    int cls[2];         // per owned cell
    unsigned valid;     // bit per slot: conf > conf_thr
};

__device__ __forceinline__ void decode_cells(const WarpWindow &w, int lane, double conf_thr,
                                             LaneCands &lc) {
    const float *st = reinterpret_cast<const float *>(w.stage);
    lc.valid = 0;
#pragma unroll
    for (int cs = 0; cs < 2; ++cs) {
        const int cell = lane + 32 * cs;
        const bool live = cell < kCells;
        const float *c = st + (live ? cell : 0) * kD;
        float cls_p[kC];
#pragma unroll
        for (int i = 0; i < kC / 2; ++i) {
            float2 v = *reinterpret_cast<const float2 *>(c + kB * 5 + 2 * i);
            cls_p[2 * i] = v.x;
            cls_p[2 * i + 1] = v.y;
        }
        float pv;
        lc.cls[cs] = argmax_first(cls_p, kC, pv);
        const double prob = (double)pv;
#pragma unroll
        for (int b = 0; b < kB; ++b) {
            const int s = cs * 2 + b;
            lc.x[s] = c[b * 5 + 0];
            lc.y[s] = c[b * 5 + 1];
            lc.w[s] = c[b * 5 + 2];
            lc.h[s] = c[b * 5 + 3];
            lc.conf[s] = __dmul_rn((double)c[b * 5 + 4], prob);  // exact: 24b x 24b
            if (live && lc.conf[s] > conf_thr) lc.valid |= 1u << s;
        }
    }
}

// Compact index (ascending candidate index k = 2*cell + b) of each valid slot.
// Returns n, the number of candidates above the threshold.
__device__ __forceinline__ int compact_slots(unsigned valid, int lane, int idx[kSlots]) {
    const unsigned lt = (1u << lane) - 1u;
    unsigned m[kSlots];
#pragma unroll
    for (int s = 0; s < kSlots; ++s) m[s] = __ballot_sync(kFull, (valid >> s) & 1u);
    const int first_half = __popc(m[0]) + __popc(m[1]);
    const int below0 = __popc(m[0] & lt) + __popc(m[1] & lt);
    const int below1 = __popc(m[2] & lt) + __popc(m[3] & lt);
    idx[0] = below0;
    idx[1] = below0 + (int)(valid & 1u);
    idx[2] = first_half + below1;
    idx[3] = first_half + below1 + (int)((valid >> 2) & 1u);
    return first_half + __popc(m[2]) + __popc(m[3]);
}

// x_abs = (col + x) / S, y_abs = (row + y) / S in doubles (metrics.py:200-201).
__device__ __forceinline__ double abs_coord(int cell_coord, float v) {
    return __ddiv_rn(__dadd_rn((double)cell_coord, (double)v), (double)kS);
}

// ---- ordering --------------------------------------------------------------------
// Writes the candidates into the sorted arrays of `w` and returns n.  After the
// call: w.scls/sk/sidx/sconf/sx1.. hold position p = (class asc, conf desc,
// candidate asc); w.cls_start[c] is the first position of class c.
__device__ __forceinline__ int order_candidates(WarpWindow &w, const LaneCands &lc, int lane) {
    int idx[kSlots];
    const int n = compact_slots(lc.valid, lane, idx);
    __syncwarp();  // every lane is done reading `stage`
    RankOverlay &ro = *reinterpret_cast<RankOverlay *>(w.stage);
    if (lane < kC) ro.count[lane] = 0;
    __syncwarp();
#pragma unroll
    for (int s = 0; s < kSlots; ++s)
        if ((lc.valid >> s) & 1u) {
            ro.conf[idx[s]] = lc.conf[s];
            ro.cls[idx[s]] = (uint8_t)lc.cls[s >> 1];
            atomicAdd(&ro.count[lc.cls[s >> 1]], 1);
        }
    __syncwarp();
    // exclusive scan of the class histogram (20 bins, in-warp)
    int cnt = (lane < kC) ? ro.count[lane] : 0;
    int inc = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        int v = __shfl_up_sync(kFull, inc, off);
        if (lane >= off) inc += v;
    }
    if (lane < kC) {
        w.cls_start[lane] = (uint8_t)(inc - cnt);
        ro.cursor[lane] = inc - cnt;
    }
    if (lane == 0) w.cls_start[kC] = (uint8_t)n;
    __syncwarp();
#pragma unroll
    for (int s = 0; s < kSlots; ++s)
        if ((lc.valid >> s) & 1u) {
            int at = atomicAdd(&ro.cursor[lc.cls[s >> 1]], 1);
            ro.member[at] = (uint8_t)idx[s];
        }
    __syncwarp();
    // rank inside the class: members that beat me (higher conf, or equal conf and
    // earlier candidate -- the stable order of sorted(..., reverse=True))
#pragma unroll
    for (int s = 0; s < kSlots; ++s)
        if ((lc.valid >> s) & 1u) {
            const int c = lc.cls[s >> 1];
            const int lo = w.cls_start[c], hi = lo + ro.count[c];
            const double mine = lc.conf[s];
            int r = 0;
            for (int q = lo; q < hi; ++q) {
                const int m = ro.member[q];
                const double other = ro.conf[m];
                r += (other > mine) || (other == mine && m < idx[s]);
            }
            const int p = lo + r;
            const int cell = lane + 32 * (s >> 1);
            const double xa = abs_coord(cell % kS, lc.x[s]);
            const double ya = abs_coord(cell / kS, lc.y[s]);
            const Corners cr = corners_of(xa, ya, (double)lc.w[s], (double)lc.h[s]);
            w.sx1[p] = cr.x1; w.sy1[p] = cr.y1; w.sx2[p] = cr.x2; w.sy2[p] = cr.y2;
            w.sarea[p] = cr.area;
            w.sconf[p] = mine;
            w.scls[p] = (uint8_t)c;
            w.sk[p] = (uint8_t)(cell * kB + (s & 1));
            w.sidx[p] = (uint8_t)idx[s];
        }
    __syncwarp();
    return n;
}

__device__ __forceinline__ Corners load_corners(const WarpWindow &w, int p) {
    Corners c;
    c.x1 = w.sx1[p]; c.y1 = w.sy1[p]; c.x2 = w.sx2[p]; c.y2 = w.sy2[p]; c.area = w.sarea[p];
    return c;
}

// 128-bit position masks: word s, bit l <-> sorted position p = 32*s + l.
struct Mask128 {
    unsigned w[kSlots];
};

// first set bit at position >= from (caller guarantees one exists)
__device__ __forceinline__ int first_from(const Mask128 &m, int from) {
    int wi = from >> 5;
    unsigned cur = 0;
#pragma unroll
    for (int s = 0; s < kSlots; ++s)
        if (s == wi) cur = m.w[s] & (kFull << (from & 31));
#pragma unroll
    for (int s = 0; s < kSlots; ++s) {
        if (s > wi && cur == 0) {
            cur = m.w[s];
            if (cur) wi = s;
        }
    }
    return (wi << 5) + __ffs(cur) - 1;
}

__device__ __forceinline__ bool test_bit(const Mask128 &m, int p) {  // select, no indexed array
    const unsigned wsel = p < 32 ? m.w[0] : (p < 64 ? m.w[1] : (p < 96 ? m.w[2] : m.w[3]));
    return (wsel >> (p & 31)) & 1u;
}

__device__ __forceinline__ int count_below(const Mask128 &m, int p) {  // set bits at positions < p
    int c = 0;
#pragma unroll
    for (int s = 0; s < kSlots; ++s) {
        if (32 * s + 32 <= p) c += __popc(m.w[s]);
        else if (32 * s < p) c += __popc(m.w[s] & ((1u << (p - 32 * s)) - 1u));
    }
    return c;
}

// ---- NMS ---------------------------------------------------------------------------
// Greedy per-class NMS over the sorted candidates.  Each round, every class's best
// still-undecided candidate is kept and all later undecided candidates of that
// class are tested against it; rounds = max kept per class.  Returns the keep mask.
template <int MODE>
__device__ __forceinline__ Mask128 nms_rounds(const WarpWindow &w, int n, double nms_thr, int lane) {
    Mask128 rem, keep;
#pragma unroll
    for (int s = 0; s < kSlots; ++s) {
        int left = n - 32 * s;
        rem.w[s] = left >= 32 ? kFull : (left > 0 ? ((1u << left) - 1u) : 0u);
        keep.w[s] = 0;
    }
    if (nms_thr != nms_thr) return rem;  // NaN threshold = decode only, keep every candidate
    while (rem.w[0] | rem.w[1] | rem.w[2] | rem.w[3]) {
        unsigned still[kSlots], kept[kSlots];
#pragma unroll
        for (int s = 0; s < kSlots; ++s) {
            const int p = lane + 32 * s;
            bool mine = (rem.w[s] >> lane) & 1u, k = false;
            if (mine) {
                const int head = first_from(rem, w.cls_start[w.scls[p]]);
                if (head == p) {
                    k = true;
                    mine = false;
                } else {
                    // reference argument order: iou(current, other)
                    mine = iou_below<MODE>(load_corners(w, head), load_corners(w, p), nms_thr);
                }
            }
            still[s] = __ballot_sync(kFull, mine);
            kept[s] = __ballot_sync(kFull, k);
        }
#pragma unroll
        for (int s = 0; s < kSlots; ++s) {
            rem.w[s] = still[s];
            keep.w[s] |= kept[s];
        }
    }
    return keep;
}

// Output offset of every class group for the class-grouped order of
// mAPMetric._apply_nms: groups appear in the order of their best member in the
// global (confidence desc, candidate asc) order.  Fills w.cls_off.
__device__ __forceinline__ void class_group_offsets(WarpWindow &w, const Mask128 &keep, int lane) {
    const int c = lane < kC ? lane : 0;
    const int lo = w.cls_start[c], hi = w.cls_start[c + 1];
    const bool present = (lane < kC) && hi > lo;
    const double hc = present ? w.sconf[lo] : 0.0;
    const int hidx = present ? w.sidx[lo] : 0;
    const int kc = present ? count_below(keep, hi) - count_below(keep, lo) : 0;
    int off = 0;
    for (int o = 0; o < kC; ++o) {
        const double oc = __shfl_sync(kFull, hc, o);
        const int oi = __shfl_sync(kFull, hidx, o);
        const int ok = __shfl_sync(kFull, kc, o);
        const bool op = __shfl_sync(kFull, (int)present, o);
        if (op && ((oc > hc) || (oc == hc && oi < hidx))) off += ok;
    }
    if (lane < kC) w.cls_off[lane] = (uint8_t)off;
    __syncwarp();
}

}  // namespace yh
