// head_image.cuh -- per-image building blocks of the decode / NMS / matching
// kernels: ONE WARP OWNS ONE IMAGE.
//
// The 98 candidate boxes of an image never leave the SM: the 5,880-byte image is
// staged into the warp's shared-memory window with coalesced 8-byte loads (an
// image start is only 8-byte aligned), decoded into a per-candidate table, ordered
// by (class, confidence desc, candidate asc) with a per-class counting rank, and
// suppressed by a bitmask NMS.
//
// Round 2: the candidate table is fp32.  It holds the corners and the area of every
// candidate rounded to float, plus the exact confidence as an orderable 64-bit key.
// Every "IoU < threshold" test is first made on fp32 intervals (iou_filter.cuh) that
// provably contain the reference's fp64 values; the exact fp64 sequence (one rounding per
// operator, as the reference's Python doubles) runs only for the few pairs the intervals
// cannot separate from the threshold, on the raw values re-read from global memory.  The
// absolute centre (col + x) / S is no longer computed in fp64 per candidate -- only for
// those rare pairs and for box outputs.
//
// Code-size discipline: the first version of this file unrolled every 4-slot loop
// and inlined every fp64 division; the kernels grew past 64 KB of SASS and ncu
// showed 85% of stall samples in `no_inst` (instruction-cache misses with ~20
// unsynchronised warps per SM; profiles/r01_map_update_v1_lines.txt).  Slot loops
// are therefore kept rolled and the rare exact paths are calls.
//
// Reference semantics reproduced here (paths relative to the reference repo):
//   decode            src/yolo/metrics.py:187-216, src/yolo/inference.py:174-208
//   NMS               src/yolo/metrics.py:274-294, src/yolo/inference.py:302-315
//   ground truths     src/yolo/metrics.py:234-254
#pragma once

#include "common.cuh"
#include "iou_filter.cuh"

namespace yh {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kStagePairs = kImgFloats / 2;   // 735 float2 per image
constexpr int kHalfCells = 32;                // cells staged per half: 0..31, then 32..48
constexpr int kHalfPairs = kHalfCells * kD / 2;  // 480 float2 = 3,840 B
constexpr int kSlots = 4;                     // candidates a lane can own
constexpr int kPad = 104;                     // 98 rounded up to a multiple of 8

// Candidate table, indexed by candidate k = 2*cell + b.  Only entries whose
// confidence passed the threshold are written / read.
struct CandTable {
    float4 box[kCand];      // fp32 corners (x1, y1, x2, y2), see iou_filter.cuh for their error bound
    uint64_t key[kCand];    // ~orderable(conf): ascending key = descending fp64 confidence (exact, invertible)
    float area[kCand];      // RN32(w * h)
    uint8_t cls[kPad];
};

// Shared-memory window of one warp: 7,088 B, so 32 warps (4 CTAs of 8) fit an SM.
struct __align__(16) WarpWindow {
    float2 stage[kHalfPairs];       // 3,840 B: half an image; reused as NmsOverlay / MatchOverlay
    CandTable cand;                 // 2,848 B
    uint8_t order[kPad];            // sorted position p -> candidate k
    uint8_t member[kPad];           // class-segmented member lists (scratch of the ordering)
    int count[kC];                  // class histogram, then running cursor
    uint8_t cls_start[kC + 4];      // first sorted position of each class; [kC] = n
    uint8_t cls_off[kC + 4];        // output offset of each class group (class-grouped order)
    uint8_t cls_kbase[kC + 4];      // kept detections at sorted positions before the class
    unsigned validw[4];             // candidates above the threshold, by candidate index
    unsigned keepw[4];              // kept after NMS, by sorted position
    uint8_t keepbase[4];            // kept detections in the words before word s
};

// What the fp32 prefilter needs to know about the image in hand (warp-uniform).
struct ImageFilter {
    float M;       // bound on the magnitude of every fp32 corner of the image's candidates (NaN if one is NaN)
    float E;       // half-width of the interval around an fp32 intersection side
    bool ok;       // no NaN / infinite / huge coordinate among the candidates (and the threshold in range)
};

// ---- exact (j + v) / 7 ----------------------------------------------------------
// x / 7 correctly rounded without the generic division sequence: q = x * RN(1/7),
// one exact remainder, one correction (Markstein).  Checked bit-for-bit against
// x / 7.0 on 5e8 random inputs of the form int + float on the CPU.
__device__ __forceinline__ double div_by_S(double x) {
    static_assert(kS == 7, "the reciprocal below is 1/7");
    const double r = 0x1.2492492492492p-3;  // RN(1/7)
    const double q = __dmul_rn(x, r);
    const double rem = __fma_rn(-q, 7.0, x);
    const double q1 = __fma_rn(rem, r, q);
    return (fabs(x) == __longlong_as_double(0x7ff0000000000000ll)) ? x : q1;
}
// x_abs = (col + x) / S, y_abs = (row + y) / S in doubles (metrics.py:200-201).
__device__ __forceinline__ double abs_coord(int cell_coord, float v) {
    return div_by_S(__dadd_rn((double)cell_coord, (double)v));
}

// The reference's fp64 corners of candidate k (or of the box in slot 0 of cell `k >> 1` of a
// target image when k is even), rebuilt from the raw values in global memory.  Rare path.
static __device__ __noinline__ Corners exact_corners(const float *__restrict__ img, int k) {
    const int cell = k >> 1;
    const float *b = img + cell * kD + (k & 1) * 5;
    const float x = __ldg(b), y = __ldg(b + 1), w = __ldg(b + 2), h = __ldg(b + 3);
    return corners_of(abs_coord(cell % kS, x), abs_coord(cell / kS, y), (double)w, (double)h);
}

// ---- staging ------------------------------------------------------------------
// Half h of an image -> w.stage: cells [32h, 32h + 32) (half 1 holds 17 cells).  An
// image start is only 8-byte aligned, so the widest coalesced access is 8 B.
__device__ __forceinline__ void stage_half(WarpWindow &w, const float *__restrict__ img, int half, int lane) {
    const float2 *src = reinterpret_cast<const float2 *>(img) + half * kHalfPairs;
    const int count = half ? (kStagePairs - kHalfPairs) : kHalfPairs;  // 255 or 480
    __syncwarp();  // every lane is done with what the buffer held before
#pragma unroll
    for (int i = 0; i < kHalfPairs / 32; ++i) {
        int q = lane + 32 * i;
        if (q < count) w.stage[q] = __ldg(src + q);
    }
    __syncwarp();
}

// Pull one image (5,880 B, 47 lines of 128 B) towards L2 ahead of its staging loads.
__device__ __forceinline__ void prefetch_image(const float *__restrict__ img, int lane) {
    const char *p = reinterpret_cast<const char *>(img);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int off = (lane + 32 * i) * 128;
        if (off < kImgFloats * 4) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + off));
    }
}

// 128-bit masks over candidates or sorted positions: word s, bit l <-> 32*s + l.
struct Mask128 {
    unsigned w0, w1, w2, w3;
    __device__ __forceinline__ unsigned word(int s) const { return s == 0 ? w0 : (s == 1 ? w1 : (s == 2 ? w2 : w3)); }
    __device__ __forceinline__ void set_word(int s, unsigned v) {
        if (s == 0) w0 = v; else if (s == 1) w1 = v; else if (s == 2) w2 = v; else w3 = v;
    }
    __device__ __forceinline__ bool any() const { return (w0 | w1 | w2 | w3) != 0; }
    __device__ __forceinline__ bool test(int p) const { return (word(p >> 5) >> (p & 31)) & 1u; }
    __device__ __forceinline__ int count() const { return __popc(w0) + __popc(w1) + __popc(w2) + __popc(w3); }
};

// set bits at positions < p
static __device__ __noinline__ int count_below(Mask128 m, int p) {
    int c = 0;
#pragma unroll
    for (int s = 0; s < kSlots; ++s) {
        const int left = p - 32 * s;
        const unsigned wd = m.word(s);
        c += left >= 32 ? __popc(wd) : (left > 0 ? __popc(wd & ((1u << left) - 1u)) : 0);
    }
    return c;
}

// first set bit at position >= from (caller guarantees one exists)
static __device__ __noinline__ int first_from(Mask128 m, int from) {
    int wi = from >> 5;
    unsigned cur = m.word(wi) & (kFull << (from & 31));
    while (cur == 0 && wi < 3) {
        ++wi;
        cur = m.word(wi);
    }
    return (wi << 5) + __ffs(cur) - 1;
}

// bit i of a 16-bit value -> bit 2i
__device__ __forceinline__ unsigned spread16(unsigned x) {
    x = (x | (x << 8)) & 0x00ff00ffu;
    x = (x | (x << 4)) & 0x0f0f0f0fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x;
}

// ---- decode ---------------------------------------------------------------------
// Each lane decodes cells `lane` and `lane + 32` (one per staged half); candidate
// k = 2*cell + b.  Fills w.cand for every candidate above conf_thr and returns the
// validity mask (indexed by k).  bad (optional) reports a kept candidate outside
// [0,1] (schemas.py:9-16,88).  filt receives the image's prefilter margin.
// tgt_img / tconf0 / tconf1 (optional): the target image and the slot-0 confidences of this
// lane's two target cells; object cells get their lines prefetched towards L2 between the
// two halves, a good while before the caller parses them.
// hook.after_half(w, cs, lane, cell, live, c): called once per half after the lane's cell has been decoded and
// before the buffer is staged again -- the fused loss of map_update.cu works on the staged cell there.
struct NoHalfHook {
    __device__ __forceinline__ void after_half(WarpWindow &, int, int, int, bool, float *) {}
};
template <class Hook>
__device__ __forceinline__ Mask128 decode_image(WarpWindow &w, const float *__restrict__ img, int lane,
                                                double conf_thr, bool *bad, ImageFilter &filt,
                                                const float *__restrict__ tgt_img, float tconf0, float tconf1,
                                                Hook &hook) {
    const float *st = reinterpret_cast<const float *>(w.stage);
    Mask128 valid = {0, 0, 0, 0};
    bool any_bad = false;
    float mag = 0.f;
    const float inv_s = 1.0f / (float)kS;
#pragma unroll 1
    for (int cs = 0; cs < 2; ++cs) {
        stage_half(w, img, cs, lane);
        if (cs == 0 && tgt_img != nullptr) {
            // the staging loads have landed, so the target confidences issued before them have too: pull the object
            // cells of BOTH halves towards L2 now, a decode (or more) ahead of whoever parses them
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if (lane + 32 * h < kCells && (h ? tconf1 : tconf0) > 0.f) {
                    const float *tc = tgt_img + (lane + 32 * h) * kD;  // 120 B: may straddle two lines
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(tc));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(tc + kD - 1));
                }
        }
        const int cell = lane + 32 * cs;
        const bool live = cell < kCells;
        const float *c = st + (live ? lane : 0) * kD;
        // the ten box values of the cell as five 8-byte loads (conflict-free at the 120-byte cell stride; scalar
        // loads of the same words collide two ways)
        const float2 *c2 = reinterpret_cast<const float2 *>(c);
        const float2 q0 = c2[0], q1 = c2[1], q2 = c2[2], q3 = c2[3], q4 = c2[4];
        const float bxs[kB][5] = {{q0.x, q0.y, q1.x, q1.y, q2.x}, {q2.y, q3.x, q3.y, q4.x, q4.y}};
        static_assert(kB == 2, "the unpacking above is for two boxes per cell");
        float pv;
        const int cls = cell_class_shared(c, conf_thr, bxs[0][4], bxs[1][4], pv);
        const double prob = (double)pv;
        const int col = cell % kS, row = cell / kS;
        const float colf = (float)col, rowf = (float)row;
        bool ok[kB];
#pragma unroll
        for (int b = 0; b < kB; ++b) {
            const double conf = __dmul_rn((double)bxs[b][4], prob);  // exact: 24b x 24b
            ok[b] = live && conf > conf_thr;
            if (ok[b]) {
                const int k = cell * kB + b;
                const float x = bxs[b][0], y = bxs[b][1], bw = bxs[b][2], bh = bxs[b][3];
                const BoxF f = make_box_f((colf + x) * inv_s, (rowf + y) * inv_s, bw, bh);
                w.cand.box[k] = make_float4(f.x1, f.y1, f.x2, f.y2);
                w.cand.area[k] = f.area;
                w.cand.key[k] = ~orderable(conf == 0.0 ? 0.0 : conf);
                w.cand.cls[k] = (uint8_t)cls;
                const float m = box_magnitude(f);
                mag = (m > mag || m != m) ? m : mag;   // NaN sticks
                if (bad != nullptr) {
                    // (col + x) / S in [0, 1] <=> col + x in [0, S] <=> x in [-col, S - col]: exact in fp32
                    // (the reference's doubles: col + x is exact, the quotient is monotone, S / S == 1)
                    any_bad |= !(x >= -colf && x <= (float)kS - colf && y >= -rowf && y <= (float)kS - rowf &&
                                 bw >= 0.f && bw <= 1.f && bh >= 0.f && bh <= 1.f && conf >= 0.0 && conf <= 1.0);
                }
            }
        }
        // k = 2*cell + b = 2*lane + b + 64*cs: interleave the two ballots into words 2cs, 2cs+1
        const unsigned m0 = __ballot_sync(kFull, ok[0]), m1 = __ballot_sync(kFull, ok[1]);
        valid.set_word(2 * cs, spread16(m0 & 0xffffu) | (spread16(m1 & 0xffffu) << 1));
        valid.set_word(2 * cs + 1, spread16(m0 >> 16) | (spread16(m1 >> 16) << 1));
        hook.after_half(w, cs, lane, cell, live, const_cast<float *>(c));
    }
    if (bad != nullptr) *bad = any_bad;
    // image-wide magnitude bound: non-negative floats (and NaN, whose pattern is the largest) order like
    // their bit patterns, so one integer max reduction does it
    const float M = __uint_as_float(__reduce_max_sync(kFull, __float_as_uint(mag)));
    filt.M = M;
    filt.ok = M < kFilterMaxCoord;
    filt.E = filter_margin(M);
    __syncwarp();
    return valid;
}
__device__ __forceinline__ Mask128 decode_image(WarpWindow &w, const float *__restrict__ img, int lane,
                                                double conf_thr, bool *bad, ImageFilter &filt,
                                                const float *__restrict__ tgt_img = nullptr,
                                                float tconf0 = 0.f, float tconf1 = 0.f) {
    NoHalfHook none;
    return decode_image(w, img, lane, conf_thr, bad, filt, tgt_img, tconf0, tconf1, none);
}

}  // namespace yh
