// head_nms.cuh -- ordering, bitmask NMS and output ordering on top of the
// candidate table built by head_image.cuh (one warp per image).
#pragma once

#include "head_image.cuh"

namespace yh {

// fp32 corners / area of candidate k from the table
__device__ __forceinline__ BoxF cand_box(const WarpWindow &w, int k) {
    const float4 b = w.cand.box[k];
    BoxF f;
    f.x1 = b.x; f.y1 = b.y; f.x2 = b.z; f.y2 = b.w;
    f.area = w.cand.area[k];
    return f;
}

// "IoU(first, second) < thr" with the reference's fp64 arithmetic and argument order, the two boxes
// rebuilt from the raw image in global memory.  The rare path behind the prefilter.
template <int MODE>
static __device__ __noinline__ bool pair_below_exact(const float *__restrict__ img, int k_first, int k_second,
                                                     double thr) {
    return iou_below<MODE>(exact_corners(img, k_first), exact_corners(img, k_second), thr);
}

// The keep test of NMS for two candidates of one image: the fp32 interval verdict when there is one,
// else the exact sequence.
template <int MODE>
__device__ __forceinline__ bool pair_below(const WarpWindow &w, const float *__restrict__ img, int k_first,
                                           int k_second, double thr, const FilterParams &fp, const ImageFilter &filt) {
    int verdict = 0;
    if (filt.ok) verdict = iou_below_fast<MODE == YH_IOU_INFER>(cand_box(w, k_first), cand_box(w, k_second), filt.E, fp);
    if (verdict == 0) return pair_below_exact<MODE>(img, k_first, k_second, thr);
    return verdict > 0;
}

// ---- class segments -------------------------------------------------------------------
// Sorted position p = (class asc, confidence desc, candidate asc).  build_members groups
// the valid candidates by class (w.member, unordered inside a class; w.cls_start) and
// returns n.  Lane owns candidates k = lane + 32*s.
__device__ __forceinline__ int build_members(WarpWindow &w, const Mask128 &valid, int lane) {
    const int n = valid.count();
    if (lane < kC) w.count[lane] = 0;
    if (lane == 0) {
        w.validw[0] = valid.w0; w.validw[1] = valid.w1; w.validw[2] = valid.w2; w.validw[3] = valid.w3;
    }
    __syncwarp();
#pragma unroll
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
#pragma unroll
    for (int s = 0; s < kSlots; ++s)
        if ((valid.word(s) >> lane) & 1u) {
            const int k = lane + 32 * s;
            const int at = atomicAdd(&w.count[w.cand.cls[k]], 1);
            w.member[at] = (uint8_t)k;
        }
    __syncwarp();
    return n;
}

// Fills w.order[p] = k by counting, for every candidate, the members of its class that
// beat it (higher confidence, or equal confidence and earlier candidate -- the stable
// order of sorted(..., reverse=True)).  General path; the common case gets its ranks
// from the pair loop of order_and_nms.
static __device__ __noinline__ void rank_members(WarpWindow &w, int lane) {
#pragma unroll 1
    for (int s = 0; s < kSlots; ++s)
        if ((w.validw[s] >> lane) & 1u) {
            const int k = lane + 32 * s;
            const int c = w.cand.cls[k];
            const int lo = w.cls_start[c], hi = w.cls_start[c + 1];
            const uint64_t mine = w.cand.key[k];
            int r = 0;
            for (int q = lo; q < hi; ++q) {
                const int m = w.member[q];
                const uint64_t other = w.cand.key[m];
                r += (other < mine) || (other == mine && m < k);
            }
            w.order[lo + r] = (uint8_t)k;
        }
    __syncwarp();
}

// ---- NMS ---------------------------------------------------------------------------
// Greedy per-class NMS over the sorted candidates, round-based: every round keeps each
// class's best undecided candidate and tests all later undecided ones against it.
// Rounds = max kept per class.  Handles any class size; used as the fallback of
// nms_pairwise (a class with more than 64 members) and for "decode only" (NaN
// threshold: keep everything).
template <int MODE>
static __device__ __noinline__ Mask128 nms_rounds(const WarpWindow &w, const float *__restrict__ img, int n,
                                                  double nms_thr, const FilterParams &fp, const ImageFilter &filt,
                                                  int lane) {
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
                    mine = pair_below<MODE>(w, img, w.order[head], k, nms_thr, fp, filt);
                }
            }
            still.set_word(s, __ballot_sync(kFull, mine));
            keep.set_word(s, keep.word(s) | __ballot_sync(kFull, kept));
        }
        rem = still;
    }
    return keep;
}

constexpr int kPairTable = 256;

// Overlay of `stage` between decode and the next staging: NMS scratch.
struct NmsOverlay {
    unsigned sup[kCand][2];         // by member position: bit i = member i of my class is earlier and overlaps me
    int beaten[kPad];               // by member position: members of my class that come before me
    uint8_t kept[kPad];             // by sorted position
    uint8_t ordm[kPad];             // sorted position -> member index inside its class
    unsigned short pair_base[kC + 4];  // pairs of all classes before c
    uint8_t pair_mark[kPairTable];     // class + 1 at the first pair of every class that has pairs, else 0
};
static_assert(sizeof(NmsOverlay) <= sizeof(float2) * kHalfPairs, "overlay too large");

// Ordering and greedy per-class NMS in one sweep over the same-class pairs:
//   1. all unordered pairs of a class are enumerated flat across the warp, so every
//      lane handles one pair per step however the classes are sized.  The pair is
//      ordered by (confidence desc, candidate asc); the later box gets one more
//      `beaten` count and, if the two overlap, the earlier one's bit in its `sup` mask;
//   2. beaten[] is the rank inside the class: scatter to w.order / ordm;
//   3. one lane per class replays the greedy scan over the bitmasks in rank order.
// Returns n and the keep mask by sorted position.  A class with more than 64 members,
// or a NaN threshold ("decode only": keep everything), takes rank_members + nms_rounds.
template <int MODE>
__device__ __forceinline__ Mask128 order_and_nms(WarpWindow &w, const float *__restrict__ img, const Mask128 &valid,
                                                 double nms_thr, const FilterParams &fp, const ImageFilter &filt,
                                                 int lane, int &n_out) {
    const int n = build_members(w, valid, lane);
    n_out = n;
    const int cc = lane < kC ? lane : 0;
    const int clo = w.cls_start[cc], chi = w.cls_start[cc + 1];
    const int m = lane < kC ? chi - clo : 0;
    if (__any_sync(kFull, m > 64) || nms_thr != nms_thr) {
        rank_members(w, lane);
        return nms_rounds<MODE>(w, img, n, nms_thr, fp, filt, lane);
    }
    NmsOverlay &no = *reinterpret_cast<NmsOverlay *>(w.stage);
    const int np = m * (m - 1) / 2;
    int inc = np;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        int v = __shfl_up_sync(kFull, inc, off);
        if (lane >= off) inc += v;
    }
    const int total = __shfl_sync(kFull, inc, 31);
    if (lane <= kC) no.pair_base[lane] = (unsigned short)(inc - np);  // [kC] = total
    for (int p = lane; p < n; p += 32) {
        no.sup[p][0] = 0;
        no.sup[p][1] = 0;
        no.beaten[p] = 0;
    }
    // Class of every pair.  With few pairs (the usual case) every class that has pairs marks its first one, and a
    // lane reads off "the last mark at or before my pair" with a max-scan across the warp (marks ascend with the
    // pair index); otherwise a lane walks the class boundaries as its pair index grows.
    const bool table = total <= kPairTable;
    if (table) {
        reinterpret_cast<uint2 *>(no.pair_mark)[lane] = make_uint2(0u, 0u);   // 32 x 8 B = kPairTable
        __syncwarp();
        if (np > 0) no.pair_mark[inc - np] = (uint8_t)(lane + 1);
    }
    __syncwarp();
    int c = 0, carry = 0;
#pragma unroll 1
    for (int t0 = 0; t0 < total; t0 += 32) {
        const int t = t0 + lane;
        const bool act = t < total;
        if (table) {
            int v = act ? no.pair_mark[t] : 0;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_up_sync(kFull, v, off);
                if (lane >= off) v = max(v, o);
            }
            v = max(v, carry);
            carry = __shfl_sync(kFull, v, 31);
            c = v - 1;
        }
        if (!act) continue;
        if (!table)
            while (t >= no.pair_base[c + 1]) ++c;
        const int u = t - no.pair_base[c];
        // row j of the strict lower triangle that holds pair u: an approximate root, put right by one step either way
        float root;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(root) : "f"((float)(8 * u + 1)));
        int j = (int)((1.0f + root) * 0.5f);
        j -= (j * (j - 1) / 2 > u);
        j += ((j + 1) * j / 2 <= u);
        const int i = u - j * (j - 1) / 2;                           // i < j, member indices
        const int lo = w.cls_start[c];
        const int ki = w.member[lo + i], kj = w.member[lo + j];
        // ascending key = descending confidence.  The leading words decide almost every pair; equal ones read the
        // whole keys (little-endian: word 1 of a key is its upper half)
        const unsigned *kw32 = reinterpret_cast<const unsigned *>(w.cand.key);
        const unsigned hi_i = kw32[2 * ki + 1], hi_j = kw32[2 * kj + 1];
        bool i_first = hi_i < hi_j;
        if (hi_i == hi_j) {
            const uint64_t ci = w.cand.key[ki], cj = w.cand.key[kj];
            i_first = (ci < cj) || (ci == cj && ki < kj);
        }
        const int e = i_first ? i : j, l = i_first ? j : i;          // earlier / later member
        atomicAdd(&no.beaten[lo + l], 1);
        // reference argument order: iou(current = earlier, other = later)
        if (!pair_below<MODE>(w, img, i_first ? ki : kj, i_first ? kj : ki, nms_thr, fp, filt))
            atomicOr(&no.sup[lo + l][e >> 5], 1u << (e & 31));
    }
    __syncwarp();
    for (int p = lane; p < n; p += 32) {
        const int k = w.member[p];
        const int lo = w.cls_start[w.cand.cls[k]];
        const int r = lo + no.beaten[p];
        w.order[r] = (uint8_t)k;
        no.ordm[r] = (uint8_t)(p - lo);
    }
    __syncwarp();
    if (lane < kC) {
        unsigned km0 = 0, km1 = 0;   // kept members of my class, by member index
        for (int r = 0; r < m; ++r) {
            const int mi = no.ordm[clo + r];
            const uint2 sp = *reinterpret_cast<const uint2 *>(no.sup[clo + mi]);
            const bool kp = ((sp.x & km0) | (sp.y & km1)) == 0;
            const unsigned bit = (kp ? 1u : 0u) << (mi & 31);
            if (mi < 32) km0 |= bit; else km1 |= bit;
            no.kept[clo + r] = kp;
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

// Publishes the keep mask and its per-word prefix in shared memory (rolled loops read
// them with a runtime word index).
__device__ __forceinline__ void publish_keep(WarpWindow &w, const Mask128 &keep, int lane) {
    if (lane == 0) {
        w.keepw[0] = keep.w0; w.keepw[1] = keep.w1; w.keepw[2] = keep.w2; w.keepw[3] = keep.w3;
        const int k1 = __popc(keep.w0), k2 = k1 + __popc(keep.w1), k3 = k2 + __popc(keep.w2);
        w.keepbase[0] = 0; w.keepbase[1] = (uint8_t)k1; w.keepbase[2] = (uint8_t)k2; w.keepbase[3] = (uint8_t)k3;
    }
    __syncwarp();
}
__device__ __forceinline__ bool kept_at(const WarpWindow &w, int p) { return (w.keepw[p >> 5] >> (p & 31)) & 1u; }

// Output offset of every class group for the class-grouped order of
// mAPMetric._apply_nms: groups appear in the order of their best member in the
// global (confidence desc, candidate asc) order.  Fills w.cls_off / w.cls_kbase.
__device__ __forceinline__ void class_group_offsets(WarpWindow &w, const Mask128 &keep, int lane) {
    const int c = lane < kC ? lane : 0;
    const int lo = w.cls_start[c], hi = w.cls_start[c + 1];
    const bool present = (lane < kC) && hi > lo;
    const int hk = present ? w.order[lo] : 0;
    const uint64_t hc = present ? w.cand.key[hk] : 0ull;
    const int kbase = present ? count_below(keep, lo) : 0;
    const int kc = present ? count_below(keep, hi) - kbase : 0;
    const unsigned have = __ballot_sync(kFull, present);
    int off = 0;
    for (unsigned m = have; m; m &= m - 1) {
        const int o = __ffs(m) - 1;
        const uint64_t oc = __shfl_sync(kFull, hc, o);
        const int ok = __shfl_sync(kFull, hk, o);
        const int on = __shfl_sync(kFull, kc, o);
        if ((oc < hc) || (oc == hc && ok < hk)) off += on;
    }
    if (lane < kC) {
        w.cls_off[lane] = (uint8_t)off;
        w.cls_kbase[lane] = (uint8_t)kbase;
    }
    __syncwarp();
}

}  // namespace yh
