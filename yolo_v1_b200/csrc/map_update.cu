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

#ifndef YH_WARPS_PER_CTA
// 8 warps x 4 CTAs = 32 images in flight per SM (64 registers per thread; 4 x (8 x 7,088 B + static + 1 KB) of
// the SM's 228 KB).  4 x 7 = 28 warps measured 1.5-2 % slower (r02: 0.319 vs 0.314 ms for map_update).
#define YH_WARPS_PER_CTA 8
#define YH_CTAS_PER_SM 4
#endif
constexpr int kWarpsPerCtaM = YH_WARPS_PER_CTA;   // warps (= images in flight) per CTA
constexpr int kCtasPerSmM = YH_CTAS_PER_SM;     // resident CTAs per SM the kernel is built for (registers, shared memory)

struct Thresholds {
    double t[YH_MAX_THRESHOLDS];
    int n;
};

// Overlay of `stage` once the predictions have been decoded: ground truths, the (detection, ground truth)
// pairs of equal class with their IoU intervals, and the per-item match results.
constexpr int kPairCap = 96;           // images with more such pairs compute the intervals inside the item loop
struct MatchOverlay {
    float4 gbox[kCells];               // fp32 corners of the ground truths, compact, in cell order      784 B
    float garea[kCells];               //                                                                 196 B
    uint8_t gcell[kCells + 7], gcls[kCells + 7], gbucket[kCells + 7];          //                         168 B
    unsigned short pbase[kCells + 7];  // first pair of ground truth g (pairs: its class's sorted range)   112 B
    float plo[kPairCap], phi[kPairCap];  // IoU interval of pair t; (0, 0) = no match possible, (-1, -1) = unknown  768 B
    uint8_t pmark[kPairCap];           // g + 1 at pbase[g] for every ground truth that has pairs, else 0     96 B
    // by item and ground-truth subset (0 all, 1 small, 2 medium, 3 large):
    unsigned short imask[kCand][4];    // bit t: best IoU >= threshold t; bit T (subset 0): best IoU >= 0.5  784 B
    uint8_t bgt[kCand][4];             // best ground truth (255 = none)                                  392 B
    // by item: candidate, class, sorted position, record slot within the image                            416 B
    uint8_t ik[kPad], icls[kPad], ipos[kPad], irec[kPad];
};
static_assert(sizeof(MatchOverlay) <= sizeof(float2) * kHalfPairs, "overlay too large");

// Thresholds in shared memory: the doubles the reference compares with, and for each the largest float not above
// and the smallest float not below it (the fp32 interval tests).  Entry T is the 0.5 of the overall
// precision / recall pass (metrics.py:477).
struct ThresholdTable {
    double t[YH_MAX_THRESHOLDS + 1];
    float lo[16], hi[16];     // entries >= T are +inf (the sorted fast path counts with a 16-entry binary search)
    float half_lo, half_hi;   // the overall pass's 0.5
    int ascending;            // thresholds are non-decreasing: "IoU >= t" holds for a prefix of them
};

// number of entries of the ascending 16-entry table `tab` that are <= v
__device__ __forceinline__ int count_le(const float *tab, float v) {
    int c = 0;
#pragma unroll
    for (int step = 8; step >= 1; step >>= 1)
        if (v >= tab[c + step - 1]) c += step;
    return c;
}

// Ground truths of one target image -> MatchOverlay (compact, in cell order).
// The target tensor is almost empty (an object in about 1 cell of 18), so it is not
// staged: tconf0/tconf1 are the slot-0 confidences of cells `lane` and `lane + 32`
// (loaded by the caller long before, 49 sectors per image), and only object cells
// fetch their box and class scores, one lane per ground truth.  Returns the number of
// ground truths; cls_mask gets bit c for every class present; gmag the largest fp32
// corner magnitude (for the prefilter margin); n_pairs the number of (sorted position,
// ground truth) pairs of equal class, whose layout goes to pbase / pmark.  gt_hist
// (shared, 4*kC ints) accumulates the per-(subset, class) counts of this CTA.
__device__ __forceinline__ int parse_targets(WarpWindow &w, const float *__restrict__ img, int lane,
                                             float tconf0, float tconf1, int *gt_hist, unsigned &cls_mask,
                                             float &gmag, int &n_pairs, uint8_t *__restrict__ out_cell,
                                             uint8_t *__restrict__ out_cls, float4 *__restrict__ out_raw) {
    const bool gt0 = tconf0 > 0.f, gt1 = (lane + 32 < kCells) && tconf1 > 0.f;  // metrics.py:239
    const unsigned m0 = __ballot_sync(kFull, gt0), m1 = __ballot_sync(kFull, gt1);
    const unsigned lt = (1u << lane) - 1u;
    const int ngt = __popc(m0) + __popc(m1);
    MatchOverlay &mo = *reinterpret_cast<MatchOverlay *>(w.stage);
    const double small_a = __dmul_rn(__ddiv_rn(32.0, 448.0), __ddiv_rn(32.0, 448.0));   // metrics.py:511
    const double medium_a = __dmul_rn(__ddiv_rn(96.0, 448.0), __ddiv_rn(96.0, 448.0));  // metrics.py:512
    const float inv_s = 1.0f / (float)kS;
    __syncwarp();  // every lane is done with what the overlay region held before
    if (gt0) mo.gcell[__popc(m0 & lt)] = (uint8_t)lane;
    if (gt1) mo.gcell[__popc(m0) + __popc(m1 & lt)] = (uint8_t)(lane + 32);
    if (lane < kPairCap / 4) reinterpret_cast<unsigned *>(mo.pmark)[lane] = 0u;
    __syncwarp();
    unsigned classes = 0;
    float mag = 0.f;
    int carry = 0;
#pragma unroll 1
    for (int g0 = 0; g0 < ngt; g0 += 32) {
        const int g = g0 + lane;
        int cnt = 0;
        if (g < ngt) {
            const int cell = mo.gcell[g];
            const float *c = img + cell * kD;
            const float2 xy = __ldg(reinterpret_cast<const float2 *>(c));
            const float2 wh = __ldg(reinterpret_cast<const float2 *>(c) + 1);
            const int cls = cell_class_global(c);
            const BoxF f = make_box_f(((float)(cell % kS) + xy.x) * inv_s, ((float)(cell / kS) + xy.y) * inv_s, wh.x, wh.y);
            mo.gbox[g] = make_float4(f.x1, f.y1, f.x2, f.y2);
            mo.garea[g] = f.area;
            const float m = box_magnitude(f);
            mag = (m > mag || m != m) ? m : mag;
            mo.gcls[g] = (uint8_t)cls;
            const double area = __dmul_rn((double)wh.x, (double)wh.y);            // exact, as metrics.py:519
            const int bucket = (area < small_a) ? 1 : ((area < medium_a) ? 2 : 3);
            mo.gbucket[g] = (uint8_t)bucket;
            if (out_raw != nullptr) {  // enough to rebuild the reference's tuple later (yh_list_form)
                out_cell[g] = (uint8_t)cell;
                out_cls[g] = (uint8_t)cls;
                out_raw[g] = make_float4(xy.x, xy.y, wh.x, wh.y);
            }
            atomicAdd(&gt_hist[cls], 1);
            atomicAdd(&gt_hist[bucket * kC + cls], 1);
            classes |= 1u << cls;
            cnt = w.cls_start[cls + 1] - w.cls_start[cls];   // candidates of its class, kept or not
        }
        int inc = cnt;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int v = __shfl_up_sync(kFull, inc, off);
            if (lane >= off) inc += v;
        }
        const int first = carry + inc - cnt;
        if (g < ngt) {
            mo.pbase[g] = (unsigned short)first;
            if (cnt > 0 && first < kPairCap) mo.pmark[first] = (uint8_t)(g + 1);
        }
        carry += __shfl_sync(kFull, inc, 31);
    }
    n_pairs = carry;
    cls_mask = __reduce_or_sync(kFull, classes);
    gmag = __uint_as_float(__reduce_max_sync(kFull, __float_as_uint(mag)));
    __syncwarp();
    return ngt;
}

// Best ground truth of detection k within subset q and the thresholds its IoU reaches, with the reference's
// own fp64 arithmetic (metrics.py:391-403, 613-623): both boxes rebuilt from the raw tensors.  The rare path
// behind the fp32 intervals.  Returns the ground truth (255 = none); mask as MatchOverlay::imask.
static __device__ __noinline__ int match_exact(const MatchOverlay &mo, const float *__restrict__ pimg,
                                               const float *__restrict__ timg, int k, int c, int q, int ngt,
                                               const ThresholdTable &th, int T, unsigned &mask) {
    const Corners dc = exact_corners(pimg, k);
    double best = 0.0;
    int bg = 255;
    for (int g = 0; g < ngt; ++g) {
        if (mo.gcls[g] != c || (q != 0 && mo.gbucket[g] != q)) continue;
        const double iou = iou_exact<YH_IOU_METRICS>(dc, exact_corners(timg, 2 * mo.gcell[g]));
        if (iou > best) {  // strict: the first ground truth wins ties, IoU 0 never matches
            best = iou;
            bg = g;
        }
    }
    unsigned m = 0;
    if (bg != 255)
        for (int t = 0; t <= T; ++t) m |= (best >= th.t[t] ? 1u : 0u) << t;
    mask = m;
    return bg;
}

// ------------------------------------------------------------------ fused loss (yh_map_update_loss)
// A validation / monitoring step wants the loss AND the metric of the same predictions; run separately the two
// kernels read the prediction and target tensors twice (loss.cu: 17,640 B per image on its own).  Here the loss
// rides on the update kernel: while a lane has its prediction cell staged in shared memory (decode_image's
// per-half hook) it adds the cell's no-object term and overwrites the cell with d loss / d prediction; the warp
// then stores the staged half as the gradient, coalesced, before the buffer is staged again.
// OBJECT cells (about 2.7 of an image's 49) are not handled there: their arithmetic (IoU of both boxes, square
// roots, the gradient through the IoU) is ~800 instructions that would sit in the middle of a kernel already at
// the edge of the instruction cache -- measured: 0.51 ms against 0.30 ms, `no_instruction` the top stall.  A second,
// small kernel (loss_object_cells, one thread per object cell) takes them from a 49-bit mask per image that the hook
// leaves behind, does loss.cu's arithmetic operator for operator (loss.py:98-159 in fp32, gradient through the IoU,
// torch's maximum / minimum / clamp conventions) and overwrites their 30 gradient values.  loss_rows_fold adds the
// per-image no-object sums and the object-cell sums up in a fixed order (deterministic) and applies the lambdas and
// 1 / N (loss.py:162-170).
struct LossArgs {
    float lambda_coord, lambda_noobj, inv_n;
    float2 *grad;        // (N, 7, 7, 30) fp32 as float2, nullable (forward only)
    float *noobj;        // [N] per image: sum of squared confidences over its no-object cells (raw)
    uint4 *objmask;      // [N] per image: bits 0..48 of (x, y) = cell holds an object; of (z, w) = its target box is in slot 1
    double *obj_sums;    // [ceil(N / kObjImages)][4] raw sums (coord, conf_obj, conf_noobj, class) of the object cells
};
constexpr int kObjImages = 64;    // images per CTA of loss_object_cells: ~170 object cells, at most one per thread as a rule

__device__ __forceinline__ float lt_max(float a, float b) { return (a > b || a != a) ? a : b; }   // torch.maximum
__device__ __forceinline__ float lt_min(float a, float b) { return (a < b || a != a) ? a : b; }
__device__ __forceinline__ float lt_relu(float v) { return (v < 0.f) ? 0.f : v; }
__device__ __forceinline__ float lt_max_mask(float self, float other) { return (self < other) ? 0.f : ((self == other) ? 0.5f : 1.f); }
__device__ __forceinline__ float lt_min_mask(float self, float other) { return (self > other) ? 0.f : ((self == other) ? 0.5f : 1.f); }

struct LtIou {
    float iou, inter, U, iw, ih, rawx, rawy, px1, px2, py1, py2;
};
__device__ __forceinline__ LtIou lt_iou(float x, float y, float w, float h, float tx1, float tx2, float ty1, float ty2,
                                        float tarea) {   // loss.py:191-212
    LtIou c;
    const float hw = w * 0.5f, hh = h * 0.5f;
    c.px1 = x - hw; c.px2 = x + hw;
    c.py1 = y - hh; c.py2 = y + hh;
    c.rawx = lt_min(c.px2, tx2) - lt_max(c.px1, tx1);
    c.rawy = lt_min(c.py2, ty2) - lt_max(c.py1, ty1);
    c.iw = lt_relu(c.rawx);
    c.ih = lt_relu(c.rawy);
    c.inter = c.iw * c.ih;
    c.U = ((w * h + tarea) - c.inter) + 1e-6f;
    c.iou = c.inter / c.U;
    return c;
}

// An object cell (rare: about one in eighteen): the full loss.cu arithmetic on the staged prediction cell `p`
// (overwritten with the gradient when want_grad) against the target cell `t` in global memory.
static __device__ __noinline__ void loss_object_cell(float *p, const float *__restrict__ t, int slot, float lc, float ln,
                                                     float inv_n, bool want_grad, float &a_coord, float &a_obj,
                                                     float &a_noobj, float &a_cls) {
    const float tx = __ldg(t + slot * 5), ty = __ldg(t + slot * 5 + 1), tw = __ldg(t + slot * 5 + 2), th = __ldg(t + slot * 5 + 3);
    const float thw = tw * 0.5f, thh = th * 0.5f;
    const float tx1 = tx - thw, tx2 = tx + thw, ty1 = ty - thh, ty2 = ty + thh, tarea = tw * th;
    const LtIou c0 = lt_iou(p[0], p[1], p[2], p[3], tx1, tx2, ty1, ty2, tarea);
    const LtIou c1 = lt_iou(p[5], p[6], p[7], p[8], tx1, tx2, ty1, ty2, tarea);
    const bool pick1 = (c1.iou > c0.iou) || (c1.iou != c1.iou && c0.iou == c0.iou);   // first maximum, NaN maximal
    const LtIou c = pick1 ? c1 : c0;
    const float x = pick1 ? p[5] : p[0], y = pick1 ? p[6] : p[1], w = pick1 ? p[7] : p[2], h = pick1 ? p[8] : p[3];
    const float cf = pick1 ? p[9] : p[4], co = pick1 ? p[4] : p[9];
    const float dx = x - tx, dy = y - ty;
    const float wc = lt_max(w, 1e-6f), hc = lt_max(h, 1e-6f);
    const float sw = sqrtf(wc), sh = sqrtf(hc);
    const float dw = sw - sqrtf(lt_max(tw, 1e-6f)), dh = sh - sqrtf(lt_max(th, 1e-6f));
    a_coord += (dx * dx + dy * dy) + (dw * dw + dh * dh);
    const float dc = cf - c.iou;
    a_obj += dc * dc;
    a_noobj += co * co;
    float cs = 0.f;
    {   // the 20 class targets as ten 8-byte loads, all in flight before the first use; summed in class order
        float2 tv[kC / 2];
#pragma unroll
        for (int i = 0; i < kC / 2; ++i) tv[i] = __ldg(reinterpret_cast<const float2 *>(t + kB * 5) + i);
#pragma unroll
        for (int i = 0; i < kC / 2; ++i) {
            float2 *pp = reinterpret_cast<float2 *>(p + kB * 5) + i;
            const float2 pv = *pp;
            const float d0 = pv.x - tv[i].x, d1 = pv.y - tv[i].y;
            cs += d0 * d0;
            cs += d1 * d1;
            if (want_grad) *pp = make_float2(inv_n * (2.f * d0), inv_n * (2.f * d1));
        }
    }
    a_cls += cs;
    if (!want_grad) return;
    const float gc = inv_n * lc, gn = inv_n * ln;
    const float g_other = gn * (2.f * co), g_resp = inv_n * (2.f * dc);
    float gx = gc * (2.f * dx), gy = gc * (2.f * dy);
    float gw = (w >= 1e-6f) ? (gc * (2.f * dw)) / (2.f * sw) : 0.f;
    float gh = (h >= 1e-6f) ? (gc * (2.f * dh)) / (2.f * sh) : 0.f;
    const float gi = -(inv_n * (2.f * dc));
    const float g_union = -gi * c.inter / (c.U * c.U);
    const float g_inter = gi / c.U - g_union;
    gw += g_union * h;
    gh += g_union * w;
    const float g_rawx = (c.rawx >= 0.f) ? g_inter * c.ih : 0.f;
    const float g_rawy = (c.rawy >= 0.f) ? g_inter * c.iw : 0.f;
    const float gpx2 = g_rawx * lt_min_mask(c.px2, tx2), gpx1 = -g_rawx * lt_max_mask(c.px1, tx1);
    const float gpy2 = g_rawy * lt_min_mask(c.py2, ty2), gpy1 = -g_rawy * lt_max_mask(c.py1, ty1);
    gx += gpx1 + gpx2;
    gy += gpy1 + gpy2;
    gw += (gpx2 - gpx1) * 0.5f;
    gh += (gpy2 - gpy1) * 0.5f;
    p[0] = pick1 ? 0.f : gx; p[1] = pick1 ? 0.f : gy; p[2] = pick1 ? 0.f : gw; p[3] = pick1 ? 0.f : gh;
    p[4] = pick1 ? g_other : g_resp;
    p[5] = pick1 ? gx : 0.f; p[6] = pick1 ? gy : 0.f; p[7] = pick1 ? gw : 0.f; p[8] = pick1 ? gh : 0.f;
    p[9] = pick1 ? g_resp : g_other;
}

struct LossHalfHook {
    LossArgs la;
    float2 *gimg;           // this image's gradient (nullable)
    float tconf0, tconf1;   // slot-0 target confidences of cells lane / lane + 32
    float tsec0, tsec1;     // slot-1 confidences of the same cells (loaded up front: no global load in the decode chain)
    float a_noobj;          // the lane's running sum of no-object confidences squared
    unsigned obj_lo, obj_hi, s1_lo, s1_hi;   // the image's object-cell mask and slot-1 mask (warp-uniform)

    __device__ __forceinline__ void after_half(WarpWindow &w, int cs, int lane, int cell, bool live, float *c) {
        const bool want_grad = gimg != nullptr;
        {
            const bool h0 = (cs ? tconf1 : tconf0) > 0.f, h1 = (cs ? tsec1 : tsec0) > 0.f;
            const unsigned ob = __ballot_sync(kFull, live && (h0 || h1)), s1 = __ballot_sync(kFull, live && !h0 && h1);
            if (cs) { obj_hi = ob; s1_hi = s1; } else { obj_lo = ob; s1_lo = s1; }
        }
        if (live) {
            const bool has0 = (cs ? tconf1 : tconf0) > 0.f, has1 = (cs ? tsec1 : tsec0) > 0.f;   // loss.py:98-99: any slot
            // every cell gets the no-object treatment here (both confidences are "noobj" terms, loss.py:150-153, and
            // nothing else has a gradient); object cells are left out of the sum and their gradient cell is
            // overwritten by loss_object_cells afterwards
            const float2 q2 = *reinterpret_cast<const float2 *>(c + 4), q4 = *reinterpret_cast<const float2 *>(c + 8);
            const float c0 = q2.x, c1 = q4.y;
            if (!(has0 || has1)) a_noobj += c0 * c0 + c1 * c1;
            if (want_grad) {
                const float gn = la.inv_n * la.lambda_noobj;
                float2 *c2 = reinterpret_cast<float2 *>(c);
                const float2 z = make_float2(0.f, 0.f);
#pragma unroll
                for (int i = 0; i < kD / 2; ++i) c2[i] = z;
                c2[2] = make_float2(gn * (2.f * c0), 0.f);
                c2[4] = make_float2(0.f, gn * (2.f * c1));
            }
        }
        if (want_grad) {
            // the staged half now holds the gradient of its cells: store it, coalesced, before the next staging
            __syncwarp();
            const int count = cs ? (kStagePairs - kHalfPairs) : kHalfPairs;
            float2 *dst = gimg + cs * kHalfPairs;
#pragma unroll
            for (int i = 0; i < kHalfPairs / 32; ++i) {
                const int q = lane + 32 * i;
                if (q < count) dst[q] = w.stage[q];
            }
        }
    }
};

// The object cells (loss.py:98-104: a cell with any slot filled) of kObjImages images per CTA, ONE THREAD PER OBJECT
// CELL: the update kernel's hook left a 49-bit mask per image, from which the CTA builds the list of its object
// cells in (image, cell) order (a prefix sum over the per-image counts); thread t then takes entries t, t + 256, ...
// -- copies the prediction cell from global memory, runs it through loss_object_cell, and replaces the 30 gradient
// values the hook stored for the cell.  A lane per object inside a warp per image measured 0.12 ms (a warp ran the
// ~800 instructions for one or two active lanes); compacted, the same work is ~20 x fewer warp instructions.
// Sums: per thread in entry order, then a fixed tree over the CTA, in doubles -> obj_sums[cta] (deterministic).
__global__ void __launch_bounds__(256)
loss_object_cells(const float *__restrict__ pred, const float *__restrict__ tgt, int N, const LossArgs la) {
    __shared__ unsigned short s_list[kObjImages * kCells];   // (image in CTA) << 7 | slot << 6 | cell
    __shared__ int s_warp[8];
    __shared__ double s_sum[256][4];
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    const int img = blockIdx.x * kObjImages + tid;
    uint4 m = make_uint4(0u, 0u, 0u, 0u);
    if (tid < kObjImages && img < N) m = la.objmask[img];
    const int cnt = __popc(m.x) + __popc(m.y);
    int inc = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(kFull, inc, off);
        if (lane >= off) inc += v;
    }
    if (lane == 31) s_warp[wp] = inc;
    __syncthreads();
    int base = inc - cnt, total = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        if (q < wp) base += s_warp[q];
        total += s_warp[q];
    }
    for (unsigned long long bits = ((unsigned long long)m.y << 32) | m.x, s1 = ((unsigned long long)m.w << 32) | m.z; bits;
         bits &= bits - 1) {
        const int cell = __ffsll((long long)bits) - 1;
        s_list[base++] = (unsigned short)((tid << 7) | ((int)((s1 >> cell) & 1ull) << 6) | cell);
    }
    __syncthreads();
    const bool want_grad = la.grad != nullptr;
    float a_coord = 0.f, a_obj = 0.f, a_noobj = 0.f, a_cls = 0.f;
    double d[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 1
    for (int e = tid; e < total; e += 256) {
        const int ent = s_list[e], cell = ent & 63, slot = (ent >> 6) & 1;
        const size_t im = (size_t)blockIdx.x * kObjImages + (ent >> 7);
        const float *t = tgt + im * kImgFloats + cell * kD;
        float p[kD];
        const float2 *src = reinterpret_cast<const float2 *>(pred + im * kImgFloats + cell * kD);
#pragma unroll
        for (int i = 0; i < kD / 2; ++i) {
            const float2 v = __ldg(src + i);
            p[2 * i] = v.x;
            p[2 * i + 1] = v.y;
        }
        a_coord = a_obj = a_noobj = a_cls = 0.f;
        loss_object_cell(p, t, slot, la.lambda_coord, la.lambda_noobj, la.inv_n, want_grad, a_coord, a_obj, a_noobj, a_cls);
        d[0] += (double)a_coord; d[1] += (double)a_obj; d[2] += (double)a_noobj; d[3] += (double)a_cls;
        if (want_grad) {
            float2 *dst = la.grad + im * kStagePairs + cell * (kD / 2);
#pragma unroll
            for (int i = 0; i < kD / 2; ++i) dst[i] = make_float2(p[2 * i], p[2 * i + 1]);
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) s_sum[tid][k] = d[k];
    __syncthreads();
    for (int st = 128; st > 0; st >>= 1) {
        if (tid < st)
#pragma unroll
            for (int k = 0; k < 4; ++k) s_sum[tid][k] += s_sum[tid + st][k];
        __syncthreads();
    }
    if (tid < 4) la.obj_sums[(size_t)blockIdx.x * 4 + tid] = s_sum[0][tid];
}

// noobj[N] + obj_sums[n_obj_ctas][4] -> the five loss values, summed in a fixed order: CTA b adds a contiguous slice
// of the per-image no-object sums (doubles); the last CTA to finish adds the CTA sums and the object-cell sums in a
// fixed order (strided per thread, then a tree) and applies the lambdas and 1 / N.
constexpr int kFoldThreads = 256;
__global__ void __launch_bounds__(kFoldThreads)
loss_rows_fold(const float *__restrict__ noobj, int N, const double *__restrict__ obj_sums, int n_obj_ctas,
               float lambda_coord, float lambda_noobj, float inv_n, double *__restrict__ partials,
               unsigned *__restrict__ ticket, float *__restrict__ parts) {
    __shared__ double s_sum[kFoldThreads][4];
    __shared__ bool s_last;
    const int tid = threadIdx.x;
    const int per = (N + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo = min(N, (int)blockIdx.x * per), hi = min(N, lo + per);
    double a = 0.0;
    for (int i = lo + tid; i < hi; i += kFoldThreads) a += (double)__ldcg(noobj + i);   // written just before
    s_sum[tid][0] = a;
    __syncthreads();
    for (int st = kFoldThreads / 2; st > 0; st >>= 1) {
        if (tid < st) s_sum[tid][0] += s_sum[tid + st][0];
        __syncthreads();
    }
    if (tid == 0) {
        partials[blockIdx.x] = s_sum[0][0];
        __threadfence();
        s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double t[4] = {0.0, 0.0, 0.0, 0.0};
    for (int b = tid; b < n_obj_ctas; b += kFoldThreads)
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] += __ldcg(obj_sums + (size_t)b * 4 + k);
    for (int b = tid; b < (int)gridDim.x; b += kFoldThreads) t[2] += __ldcg(partials + b);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; ++k) s_sum[tid][k] = t[k];
    __syncthreads();
    for (int st = kFoldThreads / 2; st > 0; st >>= 1) {
        if (tid < st)
#pragma unroll
            for (int k = 0; k < 4; ++k) s_sum[tid][k] += s_sum[tid + st][k];
        __syncthreads();
    }
    if (tid == 0) {
        const double coord = (double)lambda_coord * s_sum[0][0], cobj = s_sum[0][1];
        const double noobj_t = (double)lambda_noobj * s_sum[0][2], cls = s_sum[0][3], invn = (double)inv_n;
        parts[0] = (float)((coord + cobj + noobj_t + cls) * invn);
        parts[1] = (float)(coord * invn);
        parts[2] = (float)(cobj * invn);
        parts[3] = (float)(noobj_t * invn);
        parts[4] = (float)(cls * invn);
        *ticket = 0u;   // ready for the next call
    }
}

template <bool LOSS>
__global__ void __launch_bounds__(kWarpsPerCtaM * 32, kCtasPerSmM)
map_update_kernel(const float *__restrict__ pred, const float *__restrict__ tgt, int N,
                  double conf_thr, double nms_thr, const __grid_constant__ Thresholds thr,
                  int32_t *__restrict__ n_keep, uint64_t *__restrict__ rec_key,
                  uint64_t *__restrict__ rec_flags, uint8_t *__restrict__ rec_k,
                  uint64_t *__restrict__ rec_box, int32_t *__restrict__ n_gt, uint8_t *__restrict__ gt_cell,
                  uint8_t *__restrict__ gt_cls, float4 *__restrict__ gt_raw,
                  int32_t *__restrict__ gt_count, unsigned *__restrict__ ticket, const LossArgs la) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    WarpWindow *windows = reinterpret_cast<WarpWindow *>(smem_raw);
    __shared__ int gt_hist[4 * kC];
    __shared__ ThresholdTable th;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    WarpWindow &w = windows[wib];
    const int warps_total = gridDim.x * kWarpsPerCtaM;
    const int T = thr.n;
    const unsigned lt = (1u << lane) - 1u;
    const FilterParams fp = filter_params(nms_thr);
    for (int i = threadIdx.x; i < 4 * kC; i += blockDim.x) gt_hist[i] = 0;
    if (threadIdx.x < 16) {
        const double v = threadIdx.x < T ? thr.t[threadIdx.x] : 0.0;
        if (threadIdx.x <= YH_MAX_THRESHOLDS) th.t[threadIdx.x] = threadIdx.x < T ? v : (threadIdx.x == T ? 0.5 : 2.0);
        const float inf = __uint_as_float(0x7f800000u);
        th.lo[threadIdx.x] = threadIdx.x < T ? __double2float_rd(v) : inf;
        th.hi[threadIdx.x] = threadIdx.x < T ? __double2float_ru(v) : inf;
    }
    if (threadIdx.x == 32) {
        th.half_lo = th.half_hi = 0.5f;
        bool asc = true;
        for (int t = 1; t < T; ++t) asc = asc && thr.t[t] >= thr.t[t - 1];   // false if a NaN is involved
        th.ascending = asc ? 1 : 0;
    }
    __syncthreads();

    // A warp's first image is fixed by its position; every further one is taken from a
    // global ticket counter, so warps that drew cheap images take more of them and the
    // grid drains evenly.  The ticket for the next round is drawn a whole image ahead.
    for (int img = blockIdx.x * kWarpsPerCtaM + wib; img < N;) {
        unsigned next_ticket = 0;
        if (lane == 0) next_ticket = atomicAdd(ticket, 1u);
        const float *pimg = pred + (size_t)img * kImgFloats, *timg = tgt + (size_t)img * kImgFloats;
#ifdef YH_PREFETCH_AHEAD
        // Experiment (off): tickets are handed out in order, so image img + warps_total is drawn by some warp about
        // one image time from now; pulling it towards L2 measured no gain for decode_nms (0.229 vs 0.233 ms) and a
        // loss here (0.343 vs 0.325 ms: the target image is read sparsely, prefetching all of it adds DRAM traffic).
        if (img + YH_PREFETCH_AHEAD * warps_total < N) {
            prefetch_image(pimg + (size_t)YH_PREFETCH_AHEAD * warps_total * kImgFloats, lane);
            prefetch_image(timg + (size_t)YH_PREFETCH_AHEAD * warps_total * kImgFloats, lane);
        }
#endif
        // slot-0 confidence of my two target cells: in flight while the predictions are decoded
        const float tconf0 = __ldg(timg + lane * kD + 4);
        const float tconf1 = __ldg(timg + (lane + 32 < kCells ? lane + 32 : lane) * kD + 4);
        float tsec0 = 0.f, tsec1 = 0.f;
        if (LOSS) {   // the loss's object rule looks at every slot (loss.py:98-99)
            tsec0 = __ldg(timg + lane * kD + 9);
            tsec1 = __ldg(timg + (lane + 32 < kCells ? lane + 32 : lane) * kD + 9);
        }
        // ---- predictions: decode, order, NMS
        ImageFilter filt;
        Mask128 valid;
        if (LOSS) {
            LossHalfHook hook;
            hook.la = la;
            hook.gimg = la.grad != nullptr ? la.grad + (size_t)img * kStagePairs : nullptr;
            hook.tconf0 = tconf0;
            hook.tconf1 = tconf1;
            hook.tsec0 = tsec0;
            hook.tsec1 = tsec1;
            hook.a_noobj = 0.f;
            hook.obj_lo = hook.obj_hi = hook.s1_lo = hook.s1_hi = 0u;
            valid = decode_image(w, pimg, lane, conf_thr, nullptr, filt, timg, tconf0, tconf1, hook);
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) hook.a_noobj += __shfl_down_sync(kFull, hook.a_noobj, off);
            if (lane == 0) {
                la.noobj[img] = hook.a_noobj;
                la.objmask[img] = make_uint4(hook.obj_lo, hook.obj_hi, hook.s1_lo, hook.s1_hi);
            }
        } else {
            valid = decode_image(w, pimg, lane, conf_thr, nullptr, filt, timg, tconf0, tconf1);
        }
        ImageFilter nfilt = filt;
        nfilt.ok = filt.ok && fp.thr_ok;
        int n;
        const Mask128 keep = order_and_nms<YH_IOU_METRICS>(w, pimg, valid, nms_thr, fp, nfilt, lane, n);
        publish_keep(w, keep, lane);
        const bool lists = gt_raw != nullptr;
        if (rec_box != nullptr) {
            // the kept boxes as the reference stores them: (x_abs, y_abs) fp64, (w, h) fp32 (optional list data)
#pragma unroll 1
            for (int s = 0; s < kSlots; ++s) {
                if (!((w.keepw[s] >> lane) & 1u)) continue;
                const int k = w.order[lane + 32 * s];
                uint64_t *o = rec_box + 3 * ((size_t)img * kCand + w.keepbase[s] + __popc(w.keepw[s] & lt));
                const int cell = k >> 1;
                const float *b = pimg + cell * kD + (k & 1) * 5;
                o[0] = (uint64_t)__double_as_longlong(abs_coord(cell % kS, __ldg(b)));
                o[1] = (uint64_t)__double_as_longlong(abs_coord(cell / kS, __ldg(b + 1)));
                o[2] = (uint64_t)__float_as_uint(__ldg(b + 2)) | ((uint64_t)__float_as_uint(__ldg(b + 3)) << 32);
            }
        }

        // ---- targets: parse into the overlay
        unsigned cls_mask;
        float gmag;
        int n_pairs;
        const int ngt = parse_targets(w, timg, lane, tconf0, tconf1, gt_hist, cls_mask, gmag, n_pairs,
                                      lists ? gt_cell + (size_t)img * kCells : nullptr,
                                      lists ? gt_cls + (size_t)img * kCells : nullptr,
                                      lists ? gt_raw + (size_t)img * kCells : nullptr);
        MatchOverlay &mo = *reinterpret_cast<MatchOverlay *>(w.stage);
        if (lane == 0) {
            n_keep[img] = keep.count();
            n_gt[img] = ngt;
        }
        // prefilter margin for detection-vs-ground-truth pairs: both sets of corners count
        const float mM = fmaxf(filt.M, gmag);
        const bool mok = filt.ok && gmag < kFilterMaxCoord;
        const float mE = filter_margin(mM);
        const bool pairs_ok = mok && n_pairs <= kPairCap;

        // ---- records, in sorted order among the kept (flags = class only: true positives are patched in below),
        // and the items: kept detections whose class has a ground truth
        int n_items = 0;
        const size_t rec0 = (size_t)img * kCand;
#pragma unroll 1
        for (int s = 0; s < kSlots; ++s) {
            const unsigned kw = w.keepw[s];
            if (kw == 0) continue;   // warp-uniform
            const int p = lane + 32 * s;
            const bool kept = (kw >> lane) & 1u;
            int k = 0, c = 0, slot = 0;
            if (kept) {
                k = w.order[p];
                c = w.cand.cls[k];
                slot = w.keepbase[s] + __popc(kw & lt);
                rec_key[rec0 + slot] = w.cand.key[k];
                rec_flags[rec0 + slot] = (uint64_t)c << 56;
                if (rec_k != nullptr) rec_k[rec0 + slot] = (uint8_t)k;
            }
            const bool is_item = kept && ((cls_mask >> c) & 1u);
            const unsigned im = __ballot_sync(kFull, is_item);
            if (is_item) {
                const int it = n_items + __popc(im & lt);
                mo.ik[it] = (uint8_t)k;
                mo.icls[it] = (uint8_t)c;
                mo.ipos[it] = (uint8_t)p;
                mo.irec[it] = (uint8_t)slot;
            }
            n_items += __popc(im);
        }

        // ---- IoU intervals of all (detection, ground truth) pairs of equal class, one lane per pair: pair t of
        // ground truth g is sorted position cls_start[class] + (t - pbase[g]).  An image has a handful of them.
        if (pairs_ok && n_items > 0) {
            int carry = 0;
#pragma unroll 1
            for (int t0 = 0; t0 < n_pairs; t0 += 32) {
                const int t = t0 + lane;
                const bool act = t < n_pairs;
                int v = act ? mo.pmark[t] : 0;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const int o = __shfl_up_sync(kFull, v, off);
                    if (lane >= off) v = max(v, o);
                }
                v = max(v, carry);
                carry = __shfl_sync(kFull, v, 31);
                if (!act) continue;
                const int g = v - 1;
                const int p = w.cls_start[mo.gcls[g]] + (t - mo.pbase[g]);
                float lo = 0.f, hi = 0.f;
                if ((w.keepw[p >> 5] >> (p & 31)) & 1u) {
                    const float4 gb4 = mo.gbox[g];
                    BoxF gb;
                    gb.x1 = gb4.x; gb.y1 = gb4.y; gb.x2 = gb4.z; gb.y2 = gb4.w;
                    gb.area = mo.garea[g];
                    const int r = iou_interval(cand_box(w, w.order[p]), gb, mE, lo, hi);
                    if (r == 2) lo = hi = 0.f;                          // IoU exactly 0: never matches
                    else if (r == 0 || !(lo > 0.f)) lo = hi = -1.f;     // cannot tell (or cannot tell from 0)
                }
                mo.plo[t] = lo;
                mo.phi[t] = hi;
            }
        }
        __syncwarp();

        // ---- best ground truth and threshold mask of every (item, subset): one lane each, the reference's loop
        // per subset (metrics.py:391-403 over all ground truths of the class, :613-623 over one size bucket).
        // IoUs are fp32 intervals; a comparison the intervals cannot settle sends that (item, subset) to the
        // exact fp64 loop.
        const int n_work = 4 * n_items;
#pragma unroll 1
        for (int base = 0; base < n_work; base += 32) {
            const int wq = base + lane, q = lane & 3;
            if (wq >= n_work) continue;
            const int it = wq >> 2;
            const int k = mo.ik[it], c = mo.icls[it];
            const int rel = mo.ipos[it] - w.cls_start[c];     // my place inside every pair run of my class
            float blo = 0.f, bhi = 0.f;
            int bg = 255;
            bool amb = !mok;
            for (int g = 0; g < ngt && !amb; ++g) {
                if (mo.gcls[g] != c || (q != 0 && mo.gbucket[g] != q)) continue;
                float lo, hi;
                if (pairs_ok) {
                    lo = mo.plo[mo.pbase[g] + rel];
                    hi = mo.phi[mo.pbase[g] + rel];
                    if (hi == 0.f) continue;
                    if (lo < 0.f) { amb = true; break; }
                } else {
                    const float4 gb4 = mo.gbox[g];
                    BoxF gb;
                    gb.x1 = gb4.x; gb.y1 = gb4.y; gb.x2 = gb4.z; gb.y2 = gb4.w;
                    gb.area = mo.garea[g];
                    const int r = iou_interval(cand_box(w, k), gb, mE, lo, hi);
                    if (r == 2) continue;                         // IoU exactly 0: never matches
                    if (r == 0 || !(lo > 0.f)) { amb = true; break; }   // cannot tell IoU from 0
                }
                if (bg == 255 || lo > bhi) { blo = lo; bhi = hi; bg = g; }   // certainly greater (strict)
                else if (!(hi < blo)) amb = true;             // cannot order it against the best so far
            }
            unsigned mask = 0;
            if (!amb && bg != 255) {
                const bool ge5 = blo >= th.half_hi;              // the overall pass's IoU >= 0.5 (bit T, read for subset 0)
                amb = !ge5 && !(bhi < th.half_lo);
                if (th.ascending) {
                    // "IoU >= t" holds for a prefix of the thresholds: count it from below and from above
                    const int c_lo = count_le(th.hi, blo), c_hi = count_le(th.lo, bhi);
                    amb |= c_lo != c_hi;
                    mask = (1u << c_lo) - 1u;
                } else {
#pragma unroll 1
                    for (int t = 0; t < T; ++t) {
                        const bool ge = blo >= th.hi[t];
                        amb |= !ge && !(bhi < th.lo[t]);
                        mask |= (ge ? 1u : 0u) << t;
                    }
                }
                mask |= (ge5 ? 1u : 0u) << T;
            }
            if (amb) bg = match_exact(mo, pimg, timg, k, c, q, ngt, th, T, mask);
            mo.bgt[it][q] = (uint8_t)bg;
            mo.imask[it][q] = (unsigned short)mask;
        }
        __syncwarp();

        // ---- greedy TP/FP: an item is a TP at threshold t iff its best IoU >= t and no earlier detection of
        // its class claimed the same ground truth with IoU >= t (metrics.py:406-413).  "Earlier" = lower sorted
        // position = lower item index.  The flag word of an item that is a true positive somewhere replaces the
        // class-only word written above (the __syncwarp orders the two stores).
#pragma unroll 1
        for (int base = 0; base < n_work; base += 32) {
            const int wq = base + lane, q = lane & 3;
            const bool act = wq < n_work;
            const int it = act ? wq >> 2 : 0;
            uint64_t flags = 0;
            const unsigned g = act ? mo.bgt[it][q] : 255u;  // my best ground truth in subset q
            const int c = mo.icls[it];
            if (g != 255u) {
                unsigned claimed = 0;   // thresholds at which an earlier item took the same ground truth
                for (int e = it - 1; e >= 0 && mo.icls[e] == c; --e)
                    if (mo.bgt[e][q] == g) claimed |= mo.imask[e][q];
                const unsigned bits = mo.imask[it][q] & ~claimed;
                flags = (uint64_t)(bits & ((1u << T) - 1u)) << (q * T);
                if (q == 0 && ((bits >> T) & 1u)) flags |= 1ull << (4 * T);
            }
            // the four subsets of an item sit on four neighbouring lanes
            flags |= __shfl_xor_sync(kFull, flags, 1);
            flags |= __shfl_xor_sync(kFull, flags, 2);
            if (act && q == 0 && flags != 0) rec_flags[rec0 + mo.irec[it]] = flags | ((uint64_t)c << 56);
        }
        __syncwarp();
        asm volatile("" : "+r"(next_ticket)::"memory");  // keep the (slow) ticket read down here
        const unsigned drawn = __shfl_sync(kFull, next_ticket, 0);
        img = drawn < (unsigned)N ? warps_total + (int)drawn : N;
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
                             void *rec_box, int32_t *n_gt, uint8_t *gt_cell, uint8_t *gt_cls,
                             float *gt_raw, int32_t *gt_count, void *stream) {
    const bool fast = shape_ok(S, B, C);
    if (!fast && !generic_shape_ok(S, B, C, true)) {
        set_error("yh_map_update: shape S=%d,B=%d,C=%d outside this build (S*S*B <= 255, B <= 8, C <= 32)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (T < 1 || T > YH_MAX_THRESHOLDS) {
        set_error("yh_map_update: T=%d thresholds; this build carries 1..%d", T, YH_MAX_THRESHOLDS);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !pred || !tgt || !thresholds_host || !n_keep || !rec_key || !rec_flags ||
        !n_gt || !gt_count || ((gt_raw != nullptr) != (gt_cell != nullptr)) ||
        ((gt_raw != nullptr) != (gt_cls != nullptr))) {
        set_error("yh_map_update: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (fast ? (((uintptr_t)pred % 8) || ((uintptr_t)tgt % 8) || ((uintptr_t)rec_box % 8) || ((uintptr_t)gt_raw % 16))
             : (((uintptr_t)pred % 4) || ((uintptr_t)tgt % 4) || ((uintptr_t)rec_box % 8) || ((uintptr_t)gt_raw % 16))) {
        set_error("yh_map_update: pred/tgt/rec_box must be 8-byte aligned, gt_raw 16-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    if (N == 0) return 0;
    if (!fast)   // any other head shape: one thread per image (generic.cu)
        return generic_map_update(pred, tgt, N, S, B, C, conf_thr, nms_thr, thresholds_host, T, n_keep, rec_key, rec_flags,
                                  rec_k, rec_box, n_gt, gt_cell, gt_cls, gt_raw, gt_count, (cudaStream_t)stream);
    Thresholds th;
    th.n = T;
    for (int i = 0; i < YH_MAX_THRESHOLDS; ++i) th.t[i] = i < T ? thresholds_host[i] : 2.0;
    const size_t smem = sizeof(WarpWindow) * kWarpsPerCtaM;
    static std::atomic<bool> attr_done[kMaxDevices];
    if (first_use_on_device(attr_done)) {
        cudaFuncSetAttribute(map_update_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    int ctas = (N + kWarpsPerCtaM - 1) / kWarpsPerCtaM;
    const int cap = 148 * kCtasPerSmM;
    if (ctas > cap) ctas = cap;
    unsigned *ticket = fresh_ticket((cudaStream_t)stream);
    if (ticket == nullptr) return YH_ERR_UNSUPPORTED;
    LossArgs la = {0.f, 0.f, 0.f, nullptr, nullptr, nullptr, nullptr};
    map_update_kernel<false><<<ctas, kWarpsPerCtaM * 32, smem, (cudaStream_t)stream>>>(
        pred, tgt, N, conf_thr, nms_thr, th, n_keep, rec_key, rec_flags, rec_k, (uint64_t *)rec_box, n_gt, gt_cell,
        gt_cls, (float4 *)gt_raw, gt_count, ticket, la);
    return check_launch("yh_map_update");
}

constexpr int kFoldCtas = 64;
static inline size_t mul_up256(size_t v) { return (v + 255) & ~(size_t)255; }
// workspace: [fold ticket 256 B][fold partials][object-cell masks N x 16 B][no-object sums N x 4 B][object-cell CTA sums]
extern "C" size_t yh_map_update_loss_workspace_bytes(int N) {
    if (N < 0) return 0;
    const size_t n_obj = ((size_t)N + kObjImages - 1) / kObjImages;
    return 256 + mul_up256(sizeof(double) * kFoldCtas) + mul_up256(sizeof(uint4) * (size_t)N) +
           mul_up256(sizeof(float) * (size_t)N) + mul_up256(sizeof(double) * 4 * n_obj);
}

extern "C" int yh_map_update_loss(const float *pred, const float *tgt, int N, int S, int B, int C,
                                  double conf_thr, double nms_thr, const double *thresholds_host, int T,
                                  int32_t *n_keep, uint64_t *rec_key, uint64_t *rec_flags, uint8_t *rec_k,
                                  int32_t *n_gt, int32_t *gt_count, float lambda_coord, float lambda_noobj,
                                  float *parts, float *grad_pred, void *workspace, size_t workspace_bytes,
                                  int phases, void *stream) {
    if (phases < 1 || phases > 3) {
        set_error("yh_map_update_loss: phases must be 1 (update kernel), 2 (loss tail) or 3 (both), got %d", phases);
        return YH_ERR_BAD_ARG;
    }
    if (!shape_ok(S, B, C)) {
        set_error("yh_map_update_loss: the fused loss is built for S=7,B=2,C=20 (got %d,%d,%d); call yh_loss_fwd_bwd and "
                  "yh_map_update separately", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (T < 1 || T > YH_MAX_THRESHOLDS) {
        set_error("yh_map_update_loss: T=%d thresholds; this build carries 1..%d", T, YH_MAX_THRESHOLDS);
        return YH_ERR_UNSUPPORTED;
    }
    if (N <= 0 || !pred || !tgt || !thresholds_host || !n_keep || !rec_key || !rec_flags || !n_gt || !gt_count ||
        !parts || !workspace) {
        set_error("yh_map_update_loss: bad argument (N=%d)", N);
        return YH_ERR_BAD_ARG;
    }
    if (((uintptr_t)pred % 8) || ((uintptr_t)tgt % 8) || ((uintptr_t)grad_pred % 8) || ((uintptr_t)workspace % 256)) {
        set_error("yh_map_update_loss: pred / tgt / grad_pred must be 8-byte aligned (workspace: 256 B)");
        return YH_ERR_ALIGNMENT;
    }
    if (workspace_bytes < yh_map_update_loss_workspace_bytes(N)) {
        set_error("yh_map_update_loss: workspace too small");
        return YH_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    Thresholds th;
    th.n = T;
    for (int i = 0; i < YH_MAX_THRESHOLDS; ++i) th.t[i] = i < T ? thresholds_host[i] : 2.0;
    const size_t smem = sizeof(WarpWindow) * kWarpsPerCtaM;
    static std::atomic<bool> attr_done[kMaxDevices];
    if (first_use_on_device(attr_done))
        cudaFuncSetAttribute(map_update_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int ctas = (N + kWarpsPerCtaM - 1) / kWarpsPerCtaM;
    const int cap = 148 * kCtasPerSmM;
    if (ctas > cap) ctas = cap;
    unsigned *ticket = nullptr;
    if (phases & 1) {
        ticket = fresh_ticket(st);
        if (ticket == nullptr) return YH_ERR_UNSUPPORTED;
    }
    char *ws = (char *)workspace;
    unsigned *fold_ticket = (unsigned *)ws;                       // zero between calls (the fold resets it)
    double *partials = (double *)(ws + 256);
    char *q = ws + 256 + mul_up256(sizeof(double) * kFoldCtas);
    const int n_obj = (N + kObjImages - 1) / kObjImages;
    LossArgs la;
    la.lambda_coord = lambda_coord;
    la.lambda_noobj = lambda_noobj;
    la.inv_n = 1.0f / (float)N;
    la.grad = (float2 *)grad_pred;
    la.objmask = (uint4 *)q;
    q += mul_up256(sizeof(uint4) * (size_t)N);
    la.noobj = (float *)q;
    q += mul_up256(sizeof(float) * (size_t)N);
    la.obj_sums = (double *)q;
    int launches = 0;
    if (phases & 1) {
        map_update_kernel<true><<<ctas, kWarpsPerCtaM * 32, smem, st>>>(
            pred, tgt, N, conf_thr, nms_thr, th, n_keep, rec_key, rec_flags, rec_k, nullptr, n_gt, nullptr, nullptr, nullptr,
            gt_count, ticket, la);
        ++launches;
    }
    if (phases & 2) {   // after the update kernel (stream order, or the caller's event): overwrites gradient cells
        loss_object_cells<<<n_obj, 256, 0, st>>>(pred, tgt, N, la);
        const int fold = N < 1024 * kFoldCtas ? (N + 1023) / 1024 : kFoldCtas;
        loss_rows_fold<<<fold, kFoldThreads, 0, st>>>(la.noobj, N, la.obj_sums, n_obj, lambda_coord, lambda_noobj, la.inv_n,
                                                      partials, fold_ticket, parts);
        launches += 2;
    }
    return check_launch("yh_map_update_loss", launches);
}

extern "C" int yh_parse_gt(const float *tgt, int N, int S, int B, int C, int32_t *n_gt,
                           uint8_t *gt_cls, double *gt_box, void *stream) {
    const bool fast = shape_ok(S, B, C);
    if (!fast && !generic_shape_ok(S, B, C, true)) {
        set_error("yh_parse_gt: shape S=%d,B=%d,C=%d outside this build (S*S*B <= 255, B <= 8, C <= 32)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !tgt || !n_gt || !gt_cls || !gt_box) {
        set_error("yh_parse_gt: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (N == 0) return 0;
    if (!fast) return generic_parse_gt(tgt, N, S, B, C, n_gt, gt_cls, gt_box, (cudaStream_t)stream);
    parse_gt_kernel<<<(N + 3) / 4, 128, 0, (cudaStream_t)stream>>>(tgt, N, n_gt, gt_cls, gt_box);
    return check_launch("yh_parse_gt");
}
