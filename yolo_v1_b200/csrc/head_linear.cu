// head_linear.cu -- the detection head's last Linear(4096 -> S*S*(B*5+C)) + view, fused with the YOLOv1 loss
// forward + backward (sm_100a: tcgen05.mma with TMEM accumulators, operands brought in by TMA).
//
// Replaces, for S=7, B=2, C=20 and fp16 / bf16 operands (what `autocast("cuda")` makes of the layer,
// reference src/yolo/training/trainer.py:70, 165):
//     x = self.fc_layers[-1](x); x = x.view(-1, S, S, B*5 + C)          (src/yolo/models.py:326-331, 345-348)
//     loss, loss_dict = criterion(x, targets); loss.backward() as far as d loss / d x   (src/yolo/loss.py:55-172)
// in ONE kernel: the [M, 1470] prediction tensor never goes to HBM (optionally it is also written out, for the
// metric).  SURVEY 8(f) row 4.
//
// Shape of the kernel (one persistent CTA per SM, 320 threads, roles per warp):
//   warp 0      TMA producer: A = activations [M, 4096] and B = weight [1470, 4096], both K-major, 128-byte
//               swizzled boxes of 64 K-elements into a 4-stage shared-memory ring (mbarrier full / empty).
//   warp 1      allocates TMEM; one elected thread issues tcgen05.mma (M=128, N=224, K=16 per instruction) into
//               one of two 224-column fp32 accumulators, committing each stage back to the producer and the
//               finished accumulator to the epilogue.
//   warps 2-9   epilogue, two warps per TMEM lane quarter (a lone warp on a scheduler cannot hide its own instruction
//               latency), each taking half of the tile's cells: a thread owns one image row of the 128 x 224 tile.  An N tile is 210 useful columns =
//               7 whole cells (7 tiles x 7 cells = 49 cells); the thread reads a cell's 30 accumulators with one
//               tcgen05.ld, adds the bias, evaluates the cell's loss terms and gradient in registers (the same
//               arithmetic as loss.cu), writes the 30 gradient values and adds to its running sums.  The MMA
//               of the next tile runs meanwhile in the other accumulator.
// Tiles: (m_blk, n_blk) = (t / 7, t % 7), t = blockIdx.x + i * gridDim.x -- the 7 N tiles of an M block run at the
// same time on neighbouring SMs and share its A tile in L2; the weight (12 MB) stays L2-resident.
//
// Roofline: tensor.  Useful flops 2 * M * 4096 * 1470; issued 2 * M * 4096 * 1568 (224 / 210: 14 columns of the
// next cell are computed and dropped so that cells never straddle tiles).
#include <cuda.h>
#include <stdlib.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace yh {
namespace hl {

constexpr int kK = 4096;              // in_features
constexpr int kNOut = kCells * kD;    // 1470 out_features
constexpr int kBM = 128, kBN = 224;
constexpr int kRowBytes = 128;        // one K block of a tile row: 64 fp16 / bf16 or 32 fp32 (the 128-byte swizzle span)
constexpr int kCellsPerTile = 7;      // 210 useful columns
constexpr int kNUse = kCellsPerTile * kD;
constexpr int kNTiles = kCells / kCellsPerTile;  // 7
constexpr int kABytes = kBM * kRowBytes;   // 16,384
constexpr int kAccStride = 256;       // TMEM columns between the two accumulators
constexpr int kTmemCols = 512;
constexpr int kEpiGroups = 2;          // epilogue warps per TMEM lane quarter: each takes a share of the tile's 7 cells
constexpr int kEpiWarps = 4 * kEpiGroups;
constexpr int kThreads = 64 + 32 * kEpiWarps;   // TMA warp, MMA warp, epilogue warps
#ifndef YH_HL_CONV_WARPS
#define YH_HL_CONV_WARPS 2   // 2 / 4 / 6 warps: 3.61 / 3.63 / 3.74 ms at 65,536 rows (12 warps keep 168 registers per thread; 14 - 16 leave 128)
#endif
constexpr int kConvWarps = YH_HL_CONV_WARPS;   // 3xTF32 only: warps that split a landed fp32 stage into its hi and lo parts
static_assert(kNTiles * kCellsPerTile == kCells, "cells per tile");

// ------------------------------------------------------------------ PTX
__device__ __forceinline__ unsigned s32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(unsigned long long *b, unsigned n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory");
}
__device__ __forceinline__ void bar_expect(unsigned long long *b, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_arrive(unsigned long long *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory");
}
__device__ __forceinline__ void bar_wait(unsigned long long *b, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra D_%=;\n"
        "bra W_%=;\n"
        "D_%=:\n"
        "}\n" ::"r"(s32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int c_inner, int c_outer, unsigned long long *b) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                     s32(dst)),
                 "l"(map), "r"(s32(b)), "r"(c_inner), "r"(c_outer)
                 : "memory");
}
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
// tcgen05.commit: the mbarrier gets one arrival once every MMA issued so far by this thread has finished
__device__ __forceinline__ void mma_commit(unsigned long long *b) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(b)) : "memory");
}
__device__ __forceinline__ void mma_f16(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc, unsigned idesc,
                                        unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Shared-memory matrix descriptor of a K-major tile whose rows are 128 bytes (64 fp16 / bf16), 128-byte swizzle:
// start address / 16 in bits [0,14), stride between 8-row groups (1,024 B) / 16 in bits [32,46), descriptor version 1
// in bits [46,48), layout type 2 (SWIZZLE_128B) in bits [61,64).  The leading-dimension offset is not used by this layout.
__device__ __forceinline__ unsigned long long smem_desc(unsigned addr) {
    return (unsigned long long)((addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((unsigned long long)(1024 >> 4) << 32) | (1ull << 46) |
           (2ull << 61);
}
// Instruction descriptor, kind::f16: D = fp32 (bits [4,6) = 1), A / B format in bits [7,10) / [10,13) (0 = fp16, 1 = bf16),
// both K-major (bits 15, 16 = 0), N / 8 in bits [17,23), M / 16 in bits [24,29).
__host__ __device__ constexpr unsigned instr_desc(int fmt, int m) {
    return (1u << 4) | ((unsigned)fmt << 7) | ((unsigned)fmt << 10) | ((unsigned)(kBN >> 3) << 17) | ((unsigned)(m >> 4) << 24);
}
// 32 consecutive fp32 columns of this thread's TMEM lane (the warp reads its own 32-lane quarter)
__device__ __forceinline__ void tmem_ld32(unsigned taddr, float (&v)[32]) {
    unsigned r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        "tcgen05.wait::ld.sync.aligned;\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// ------------------------------------------------------------------ element types
template <typename T> struct Elem;
template <> struct Elem<__half> {
    static constexpr int fmt = 0;
    static __device__ __forceinline__ float round(float v) { return __half2float(__float2half_rn(v)); }
    static __device__ __forceinline__ unsigned pack(float a, float b) {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<unsigned *>(&h);
    }
};
template <> struct Elem<__nv_bfloat16> {
    static constexpr int fmt = 1;
    static __device__ __forceinline__ float round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
    static __device__ __forceinline__ unsigned pack(float a, float b) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<unsigned *>(&h);
    }
};

template <> struct Elem<float> {   // fp32 operands: the tensor core reads them as TF32 (kind::tf32); nothing to round, fp32 gradient
    static constexpr int fmt = 2;
    static __device__ __forceinline__ float round(float v) { return v; }
    static __device__ __forceinline__ unsigned pack(float a, float) { return __float_as_uint(a); }
};

// ------------------------------------------------------------------ the loss of one cell, in registers
// torch.maximum / torch.minimum / clamp(min=0): NaN propagates (as in loss.cu).
__device__ __forceinline__ float t_max(float a, float b) { return (a > b || a != a) ? a : b; }
__device__ __forceinline__ float t_min(float a, float b) { return (a < b || a != a) ? a : b; }
__device__ __forceinline__ float max_mask(float self, float other) { return (self < other) ? 0.f : ((self == other) ? 0.5f : 1.f); }
__device__ __forceinline__ float min_mask(float self, float other) { return (self > other) ? 0.f : ((self == other) ? 0.5f : 1.f); }

struct Iou {   // loss.py:191-212 in fp32, one rounding per operator, with what the backward needs
    float iou, inter, U, iw, ih, rawx, rawy, px1, px2, py1, py2;
};
__device__ __forceinline__ Iou iou_fwd(float x, float y, float w, float h, float tx1, float tx2, float ty1, float ty2, float tarea) {
    Iou c;
    const float hw = w * 0.5f, hh = h * 0.5f;
    c.px1 = x - hw; c.px2 = x + hw;
    c.py1 = y - hh; c.py2 = y + hh;
    c.rawx = t_min(c.px2, tx2) - t_max(c.px1, tx1);
    c.rawy = t_min(c.py2, ty2) - t_max(c.py1, ty1);
    c.iw = c.rawx < 0.f ? 0.f : c.rawx;
    c.ih = c.rawy < 0.f ? 0.f : c.rawy;
    c.inter = c.iw * c.ih;
    const float uni = (w * h + tarea) - c.inter;
    c.U = uni + 1e-6f;
    c.iou = c.inter / c.U;
    return c;
}

struct LossSums {   // one tile's share (at most 4 cells) in fp32; the thread's running sums over all its tiles are fp64
    float coord, obj, noobj, cls;
};

// p[0..29]: the cell's predictions; tg: its 30 target floats (8-byte aligned).  Writes d total / d p into g[0..29]
// (not yet multiplied by the caller's gradient scale) and adds the cell's four raw sums (loss.py:126-157).
__device__ __forceinline__ void cell_loss(const float (&p)[32], const float2 *__restrict__ tg, bool has0, bool has1, float inv_n,
                                          float lambda_coord, float lambda_noobj, float (&g)[30], LossSums &acc) {
#pragma unroll
    for (int i = 0; i < 30; ++i) g[i] = 0.f;
    const float gn = inv_n * lambda_noobj;
    if (!(has0 || has1)) {                                  // loss.py:150-153: both confidences are no-object terms
        acc.noobj += p[4] * p[4] + p[9] * p[9];
        g[4] = gn * (2.f * p[4]);
        g[9] = gn * (2.f * p[9]);
        return;
    }
    float t[30];
#pragma unroll
    for (int i = 0; i < 15; ++i) {
        const float2 v = __ldg(tg + i);
        t[2 * i] = v.x;
        t[2 * i + 1] = v.y;
    }
    const int slot = (!has0 && has1) ? 1 : 0;              // loss.py:102: the first filled slot
    const float tx = slot ? t[5] : t[0], ty = slot ? t[6] : t[1], tw = slot ? t[7] : t[2], th = slot ? t[8] : t[3];
    const float thw = tw * 0.5f, thh = th * 0.5f;
    const float tx1 = tx - thw, tx2 = tx + thw, ty1 = ty - thh, ty2 = ty + thh, tarea = tw * th;
    const Iou c0 = iou_fwd(p[0], p[1], p[2], p[3], tx1, tx2, ty1, ty2, tarea);
    const Iou c1 = iou_fwd(p[5], p[6], p[7], p[8], tx1, tx2, ty1, ty2, tarea);
    const bool pick1 = (c1.iou > c0.iou) || (c1.iou != c1.iou && c0.iou == c0.iou);   // first maximum, NaN maximal
    const Iou c = pick1 ? c1 : c0;
    const float x = pick1 ? p[5] : p[0], y = pick1 ? p[6] : p[1], w = pick1 ? p[7] : p[2], h = pick1 ? p[8] : p[3];
    const float cf = pick1 ? p[9] : p[4], co = pick1 ? p[4] : p[9];
    const float dx = x - tx, dy = y - ty;
    const float wc = t_max(w, 1e-6f), hc = t_max(h, 1e-6f);
    const float sw = sqrtf(wc), sh = sqrtf(hc);
    const float dw = sw - sqrtf(t_max(tw, 1e-6f)), dh = sh - sqrtf(t_max(th, 1e-6f));
    acc.coord += (dx * dx + dy * dy) + (dw * dw + dh * dh);
    const float dc = cf - c.iou;                            // the IoU target is not detached (loss.py:142-147)
    acc.obj += dc * dc;
    acc.noobj += co * co;
    const float gc = inv_n * lambda_coord;
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
    const float gpx2 = g_rawx * min_mask(c.px2, tx2), gpx1 = -g_rawx * max_mask(c.px1, tx1);
    const float gpy2 = g_rawy * min_mask(c.py2, ty2), gpy1 = -g_rawy * max_mask(c.py1, ty1);
    gx += gpx1 + gpx2;
    gy += gpy1 + gpy2;
    gw += (gpx2 - gpx1) * 0.5f;
    gh += (gpy2 - gpy1) * 0.5f;
    g[0] = pick1 ? 0.f : gx; g[1] = pick1 ? 0.f : gy; g[2] = pick1 ? 0.f : gw; g[3] = pick1 ? 0.f : gh;
    g[4] = pick1 ? g_other : g_resp;
    g[5] = pick1 ? gx : 0.f; g[6] = pick1 ? gy : 0.f; g[7] = pick1 ? gw : 0.f; g[8] = pick1 ? gh : 0.f;
    g[9] = pick1 ? g_resp : g_other;
    float cs = 0.f;                                         // loss.py:156-157, summed in class order
#pragma unroll
    for (int i = 10; i < 30; ++i) {
        const float d = p[i] - t[i];
        cs += d * d;
        g[i] = inv_n * (2.f * d);
    }
    acc.cls += cs;
}

// ------------------------------------------------------------------ the kernel
struct Args {
    const float *bias;        // [1470] or nullptr
    const float2 *tgt;        // [M, 7, 7, 30] fp32, or nullptr: plain Linear (pred_out only)
    int M;
    int round_pred;           // round the prediction to the operand type first, as the autocast Linear's output is
    float lambda_coord, lambda_noobj, inv_n, grad_scale;
    void *grad;               // [M, 1470] in GT, or nullptr
    float *pred_out;          // [M, 1470] fp32, or nullptr
    double *partials;         // [gridDim.x][4]
    unsigned *ticket;
    float *parts;             // total, coord, conf_obj, conf_noobj, class
};

// CTAS = 2: two CTAs of a cluster (one SM pair) share a 256 x 224 tile.  Each loads its own 128 rows of A and HALF of
// the B tile (112 rows); the leader's thread issues tcgen05.mma.cta_group::2 (M = 256), which reads A from each CTA's
// own shared memory and the two halves of B from both, and accumulates into each CTA's own TMEM.  Per SM that is
// 7.5 KB of shared-memory reads per K = 16 step instead of 11 KB, and 30 KB of TMA writes per stage instead of 44 KB --
// the single-CTA form needs ~200 B / clk of a 128 B / clk shared memory to keep the tensor core busy.
// PASSES = 3 (fp32 operands only): 3xTF32.  a = hi + lo with hi = a's top 19 bits (what TF32 keeps) and lo = a - hi (exact in
// fp32); x.w ~ hi.hi + hi.lo + lo.hi leaves a relative error of ~2^-21 per product, fp32's own.  Two converter warps
// turn every landed stage into (hi in place, lo in a second tile) before the MMA warp may read it.
template <int CTAS, int PASSES>
struct Cfg {
    static constexpr int kBRows = kBN / CTAS;                    // rows of the B tile one CTA holds
    static constexpr int kBBytes1 = kBRows * kRowBytes;
    static constexpr int kStageBytes1 = kABytes + kBBytes1;      // 45,056 / 30,720
    static constexpr int kStages1 = PASSES == 3 ? 3 : (CTAS == 1 ? 4 : 6);
    static constexpr int kThreads1 = kThreads + (PASSES == 3 ? 32 * kConvWarps : 0);
};
template <int CTAS, int PASSES>
struct SmemT {
    alignas(1024) unsigned char tiles[Cfg<CTAS, PASSES>::kStages1][Cfg<CTAS, PASSES>::kStageBytes1];
    alignas(1024) unsigned char lo[PASSES == 3 ? Cfg<CTAS, PASSES>::kStages1 : 1][PASSES == 3 ? Cfg<CTAS, PASSES>::kStageBytes1 : 16];
    alignas(8) unsigned long long full[Cfg<CTAS, PASSES>::kStages1], empty[Cfg<CTAS, PASSES>::kStages1],
        conv[Cfg<CTAS, PASSES>::kStages1], acc_full[2], acc_empty[2];
    unsigned tmem_base;
    float bias[kNTiles * kBN];        // bias of every N tile, padded to 224 (zeros)
    float2 out[kEpiWarps][32 * (kD / 2)];   // one cell of a warp's 32 rows on its way out (3,840 B per warp)
    double red[kEpiWarps][4];
    bool last;
};

__device__ __forceinline__ unsigned cluster_rank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// the address, in the cluster's shared window, of `local` in CTA `rank`
__device__ __forceinline__ unsigned map_to_rank(unsigned local, unsigned rank) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
__device__ __forceinline__ void bar_arrive_cluster(unsigned cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(void *dst, const CUtensorMap *map, int c_inner, int c_outer, unsigned bar_cluster_addr) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            s32(dst)),
        "l"(map), "r"(bar_cluster_addr), "r"(c_inner), "r"(c_outer)
        : "memory");
}
__device__ __forceinline__ void mma_commit_pair(unsigned long long *b) {   // arrives at this offset in both CTAs
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(s32(b)),
                 "h"((unsigned short)3)
                 : "memory");
}
__device__ __forceinline__ void mma_f16_pair(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc, unsigned idesc,
                                             unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void mma_tf32(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc, unsigned idesc,
                                         unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mma_tf32_pair(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc, unsigned idesc,
                                              unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
template <typename T, int CTAS>
__device__ __forceinline__ void mma_any(unsigned d, unsigned long long ad, unsigned long long bd, unsigned idesc, unsigned acc) {
    if (sizeof(T) == 4) {
        if (CTAS == 2) mma_tf32_pair(d, ad, bd, idesc, acc); else mma_tf32(d, ad, bd, idesc, acc);
    } else {
        if (CTAS == 2) mma_f16_pair(d, ad, bd, idesc, acc); else mma_f16(d, ad, bd, idesc, acc);
    }
}
// arrival that publishes this thread's earlier writes to a waiter in another CTA of the cluster
__device__ __forceinline__ void bar_arrive_cluster_release(unsigned cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void bar_wait_cluster_acquire(unsigned long long *b, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W_%=:\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
        "@p bra D_%=;\n"
        "bra W_%=;\n"
        "D_%=:\n"
        "}\n" ::"r"(s32(b)), "r"(parity) : "memory");
}

template <typename T, typename GT, int CTAS, int PASSES>
__global__ void __launch_bounds__(Cfg<CTAS, PASSES>::kThreads1, 1)   // 10 (12) warps are allocated as 12: 168 registers per thread
head_linear_loss_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                        const __grid_constant__ CUtensorMap map_blo, const Args a) {
    using C = Cfg<CTAS, PASSES>;
    constexpr int kSt = C::kStages1;
    constexpr int kBKe = kRowBytes / (int)sizeof(T);            // K elements per block: 64 (fp16 / bf16) or 32 (fp32)
    constexpr int kKBlocks = kK / kBKe;
    constexpr int kNThreads = C::kThreads1;
    constexpr bool kOwnFull = CTAS == 1 || PASSES == 3;          // every CTA waits for its own loads (the converter needs to)
    // No static shared memory in this kernel: the dynamic window starts at offset 0 of the CTA's shared memory, which is
    // what gives the tiles the 1,024-byte alignment the 128-byte swizzle needs (checked below).  Used in place, so that
    // the compiler keeps the shared address space (an integer round-up of the pointer turns every access generic).
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    SmemT<CTAS, PASSES> &sm = *reinterpret_cast<SmemT<CTAS, PASSES> *>(smem_raw);
    if ((s32(smem_raw) & 1023u) != 0u) __trap();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned rank = CTAS == 2 ? cluster_rank() : 0u;       // 0 = the leader: issues the MMAs, owns full / acc_empty
    const int unit = blockIdx.x / CTAS, units = gridDim.x / CTAS;
    const int m_units = (a.M + kBM * CTAS - 1) / (kBM * CTAS);
    const int total = m_units * kNTiles;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kSt; ++s) {
            bar_init(&sm.full[s], 1);
            bar_init(&sm.empty[s], 1);
            bar_init(&sm.conv[s], kConvWarps * CTAS);
        }
        for (int s = 0; s < 2; ++s) {
            bar_init(&sm.acc_full[s], 1);
            bar_init(&sm.acc_empty[s], kEpiWarps * CTAS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    for (int i = threadIdx.x; i < kNTiles * kBN; i += kNThreads) {
        const int nb = i / kBN, j = i - nb * kBN;
        sm.bias[i] = (a.bias != nullptr && j < kNUse) ? __ldg(a.bias + nb * kNUse + j) : 0.f;
    }
    if (warp == 1) {   // the warp that allocates TMEM also frees it (the same warp in both CTAs of a pair)
        if (CTAS == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&sm.tmem_base)), "n"(kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&sm.tmem_base)), "n"(kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    fence_before();
    if (CTAS == 2) cluster_sync(); else __syncthreads();          // barriers of both CTAs initialised before anyone signals them
    fence_after();
    const unsigned tmem = sm.tmem_base;

    if (warp == 0) {
        if (lane == 0) {                       // ---- TMA producer (in every CTA: its A rows, its share of the B rows)
            int s = 0, ph = 0;
            for (int t = unit; t < total; t += units) {
                const int m_unit = t / kNTiles, n_blk = t - m_unit * kNTiles;
                const int a_row = (m_unit * CTAS + (int)rank) * kBM, b_row = n_blk * kNUse + (int)rank * C::kBRows;
                for (int kb = 0; kb < kKBlocks; ++kb) {
                    bar_wait(&sm.empty[s], ph ^ 1);
                    if (!kOwnFull) {
                        // all bytes of both CTAs complete on the leader's barrier; the leader arms it for the pair
                        const unsigned full0 = map_to_rank(s32(&sm.full[s]), 0);
                        if (rank == 0) bar_expect(&sm.full[s], 2 * C::kStageBytes1);
                        tma_load_2d_pair(sm.tiles[s], &map_a, kb * kBKe, a_row, full0);
                        tma_load_2d_pair(sm.tiles[s] + kABytes, &map_b, kb * kBKe, b_row, full0);
                    } else {
                        // 3xTF32: the weight's lo part was split off once per call (split_lo_kernel) and arrives by TMA as well
                        bar_expect(&sm.full[s], C::kStageBytes1 + (PASSES == 3 ? C::kBBytes1 : 0));
                        tma_load_2d(sm.tiles[s], &map_a, kb * kBKe, a_row, &sm.full[s]);
                        tma_load_2d(sm.tiles[s] + kABytes, &map_b, kb * kBKe, b_row, &sm.full[s]);
                        if constexpr (PASSES == 3) tma_load_2d(sm.lo[s] + kABytes, &map_blo, kb * kBKe, b_row, &sm.full[s]);
                    }
                    if (++s == kSt) { s = 0; ph ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {          // ---- MMA issuer (the leader's, for the pair)
            const unsigned idesc = instr_desc(Elem<T>::fmt, kBM * CTAS);
            int s = 0, ph = 0, as = 0, aph = 0;
            for (int t = unit; t < total; t += units) {
                bar_wait(&sm.acc_empty[as], aph ^ 1);          // the epilogue (of both CTAs) has drained this accumulator
                fence_after();
                const unsigned d = tmem + as * kAccStride;
                // 3xTF32: the two small cross terms get an accumulator of their own.  The tensor core truncates on every
                // accumulate (~2^-24 of the accumulator per MMA): in one accumulator the 1,024 extra steps of the cross terms
                // would cost more accuracy (2e-5 absolute on a prediction) than the terms add; apart, they cost nothing.
                const unsigned d_small = tmem + kAccStride;
                for (int kb = 0; kb < kKBlocks; ++kb) {
                    if (PASSES == 3) bar_wait_cluster_acquire(&sm.conv[s], ph);   // both CTAs' stages are split into hi and lo
                    else bar_wait(&sm.full[s], ph);
                    fence_after();
                    const unsigned long long ad = smem_desc(s32(sm.tiles[s])), bd = smem_desc(s32(sm.tiles[s] + kABytes));
#pragma unroll
                    for (int k = 0; k < kRowBytes / 32; ++k) {      // 32 bytes along K per instruction: + 2 in the address field
                        mma_any<T, CTAS>(d, ad + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
                        if constexpr (PASSES == 3) {
                            const unsigned long long al = smem_desc(s32(sm.lo[s])), bl = smem_desc(s32(sm.lo[s] + kABytes));
                            mma_any<T, CTAS>(d_small, ad + 2 * k, bl + 2 * k, idesc, (kb | k) != 0);     // hi . lo
                            mma_any<T, CTAS>(d_small, al + 2 * k, bd + 2 * k, idesc, 1u);                // lo . hi
                        }
                    }
                    if (CTAS == 2) mma_commit_pair(&sm.empty[s]); else mma_commit(&sm.empty[s]);   // frees the stage
                    if (++s == kSt) { s = 0; ph ^= 1; }
                }
                if (CTAS == 2) mma_commit_pair(&sm.acc_full[as]); else mma_commit(&sm.acc_full[as]);
                if (PASSES == 3) aph ^= 1;                      // one accumulator pair: the next tile waits for the epilogue
                else { as ^= 1; if (as == 0) aph ^= 1; }
            }
        }
        __syncwarp();
    } else if (PASSES == 3 && warp >= 2 + kEpiWarps) {
      if constexpr (PASSES == 3) {
        // ---- converters (3xTF32): hi = the top 19 bits in place, lo = value - hi in the second tile; elementwise, so the
        // swizzled layout does not matter.  Then the writes are made visible to the tensor core (async proxy) and the
        // leader's MMA warp is told, across the cluster.
        const int ct = (warp - 2 - kEpiWarps) * 32 + lane;
        int s = 0, ph = 0;
        const unsigned conv0 = CTAS == 2 ? map_to_rank(s32(&sm.conv[0]), 0) : s32(&sm.conv[0]);
        for (int t = unit; t < total; t += units) {
            for (int kb = 0; kb < kKBlocks; ++kb) {
                bar_wait(&sm.full[s], ph);
                uint4 *hi4 = reinterpret_cast<uint4 *>(sm.tiles[s]);
                float4 *lo4 = reinterpret_cast<float4 *>(sm.lo[s]);
#pragma unroll
                for (int i = ct; i < kABytes / 16; i += 32 * kConvWarps) {      // the activations' rows; the weight's lo came by TMA
                    const uint4 v = hi4[i];
                    const uint4 h = make_uint4(v.x & 0xFFFFE000u, v.y & 0xFFFFE000u, v.z & 0xFFFFE000u, v.w & 0xFFFFE000u);
#ifdef YH_HL_HI_MASK   // not needed: the tensor core ignores the low 13 bits of a TF32 operand (same results bit for bit, 4 % faster)
                    hi4[i] = h;
#endif
                    lo4[i] = make_float4(__uint_as_float(v.x) - __uint_as_float(h.x), __uint_as_float(v.y) - __uint_as_float(h.y),
                                         __uint_as_float(v.z) - __uint_as_float(h.z), __uint_as_float(v.w) - __uint_as_float(h.w));
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) bar_arrive_cluster_release(conv0 + (unsigned)(s * sizeof(unsigned long long)));
                if (++s == kSt) { s = 0; ph ^= 1; }
            }
        }
        __syncwarp();
      }
    } else {                                   // ---- epilogue: warps 2..9 own TMEM lanes 32 * (warp % 4) ..
        const int q = warp & 3, grp = (warp - 2) >> 2;
        const int c_lo = (grp * kCellsPerTile + kEpiGroups - 1) / kEpiGroups, c_hi = ((grp + 1) * kCellsPerTile + kEpiGroups - 1) / kEpiGroups;
        double sum_coord = 0.0, sum_obj = 0.0, sum_noobj = 0.0, sum_cls = 0.0;
        int as = 0, aph = 0;
        for (int t = unit; t < total; t += units) {
            const int m_unit = t / kNTiles, n_blk = t - m_unit * kNTiles;
            const int img0 = (m_unit * CTAS + (int)rank) * kBM + q * 32;    // the warp's first row
            const int img = img0 + lane;
            const bool live = img < a.M;
            const size_t cell0 = (size_t)img * kCells + n_blk * kCellsPerTile;
            // the two object flags (t[4] > 0, t[9] > 0) of the row's cells, in flight while the accumulator is still being made
            unsigned has0 = 0u, has1 = 0u;
            if (live && a.tgt != nullptr) {
#pragma unroll
                for (int c = 0; c < kCellsPerTile; ++c) {
                    if (c < c_lo || c >= c_hi) continue;
                    const float *tc = reinterpret_cast<const float *>(a.tgt + (cell0 + c) * (kD / 2));
                    has0 |= (unsigned)(__ldg(tc + 4) > 0.f) << c;
                    has1 |= (unsigned)(__ldg(tc + 9) > 0.f) << c;
                }
            }
            bar_wait(&sm.acc_full[as], aph);
            fence_after();
            const unsigned taddr = tmem + ((unsigned)(q * 32) << 16) + as * kAccStride;
            float2 *stage = sm.out[warp - 2];
            LossSums acc = {0.f, 0.f, 0.f, 0.f};
            const int rows_live = a.M - img0 < 32 ? (a.M - img0 < 0 ? 0 : a.M - img0) : 32;
#pragma unroll 1
            for (int c = c_lo; c < c_hi; ++c) {
                float v[32];
                __syncwarp();                  // tcgen05.ld is warp-wide: lanes that left the previous cell early wait here
                tmem_ld32(taddr + c * kD, v);
                if constexpr (PASSES == 3) {   // + the cross terms' accumulator
                    float u[32];
                    tmem_ld32(taddr + kAccStride + c * kD, u);
#pragma unroll
                    for (int j = 0; j < kD; ++j) v[j] += u[j];
                }
                const float *b = sm.bias + n_blk * kBN + c * kD;
#pragma unroll
                for (int j = 0; j < kD; ++j) {
                    v[j] += b[j];
                    if (a.round_pred) v[j] = Elem<T>::round(v[j]);
                }
                // A cell of the warp's 32 rows leaves through shared memory: a lane puts its row's 30 values down (row
                // stride 15 x 8 B, odd: conflict-free), then the warp writes row after row with neighbouring lanes on
                // neighbouring words -- 60 / 120 contiguous bytes per row instead of 32 scattered words per instruction.
                const size_t out0 = ((size_t)img0 * kCells + n_blk * kCellsPerTile + c) * (kD / 2);   // word index, row 0
                if (a.pred_out != nullptr) {
#pragma unroll
                    for (int j = 0; j < kD / 2; ++j) stage[lane * (kD / 2) + j] = make_float2(v[2 * j], v[2 * j + 1]);
                    __syncwarp();
                    float2 *o = reinterpret_cast<float2 *>(a.pred_out);
#pragma unroll
                    for (int k = 0; k < kD / 2; ++k) {
                        const int w = k * 32 + lane, r = w / (kD / 2), j = w - r * (kD / 2);
                        if (r < rows_live) o[out0 + (size_t)r * (kCells * kD / 2) + j] = stage[w];
                    }
                    __syncwarp();
                }
                if (a.tgt != nullptr) {
                    float g[30];
                    if (live) cell_loss(v, a.tgt + (cell0 + c) * (kD / 2), (has0 >> c) & 1u, (has1 >> c) & 1u, a.inv_n, a.lambda_coord, a.lambda_noobj, g, acc);
                    if (a.grad != nullptr) {
                        __syncwarp();
                        if (sizeof(GT) == 4) {
#pragma unroll
                            for (int j = 0; j < kD / 2; ++j)
                                stage[lane * (kD / 2) + j] = make_float2(g[2 * j] * a.grad_scale, g[2 * j + 1] * a.grad_scale);
                            __syncwarp();
                            float2 *o = reinterpret_cast<float2 *>(a.grad);
#pragma unroll
                            for (int k = 0; k < kD / 2; ++k) {
                                const int w = k * 32 + lane, r = w / (kD / 2), j = w - r * (kD / 2);
                                if (r < rows_live) o[out0 + (size_t)r * (kCells * kD / 2) + j] = stage[w];
                            }
                        } else {
                            unsigned *st32 = reinterpret_cast<unsigned *>(stage);
#pragma unroll
                            for (int j = 0; j < kD / 2; ++j)
                                st32[lane * (kD / 2) + j] = Elem<T>::pack(g[2 * j] * a.grad_scale, g[2 * j + 1] * a.grad_scale);
                            __syncwarp();
                            unsigned *o = reinterpret_cast<unsigned *>(a.grad);
#pragma unroll
                            for (int k = 0; k < kD / 2; ++k) {
                                const int w = k * 32 + lane, r = w / (kD / 2), j = w - r * (kD / 2);
                                if (r < rows_live) o[out0 + (size_t)r * (kCells * kD / 2) + j] = st32[w];
                            }
                        }
                        __syncwarp();
                    }
                }
            }
            sum_coord += (double)acc.coord;
            sum_obj += (double)acc.obj;
            sum_noobj += (double)acc.noobj;
            sum_cls += (double)acc.cls;
            fence_before();
            __syncwarp();
            if (lane == 0) {
                if (CTAS == 2) bar_arrive_cluster(map_to_rank(s32(&sm.acc_empty[as]), 0));
                else bar_arrive(&sm.acc_empty[as]);
            }
            if (PASSES == 3) aph ^= 1;
            else { as ^= 1; if (as == 0) aph ^= 1; }
        }
        // the warp's four sums (fp64 from here on)
        double s4[4] = {sum_coord, sum_obj, sum_noobj, sum_cls};
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) s4[k] += __shfl_down_sync(0xffffffffu, s4[k], off);
        if (lane == 0)
#pragma unroll
            for (int k = 0; k < 4; ++k) sm.red[warp - 2][k] = s4[k];
    }
    fence_before();
    if (CTAS == 2) cluster_sync(); else __syncthreads();          // nobody leaves while its partner may still signal it
    if (warp == 1) {
        fence_after();
        if (CTAS == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
    }
    if (a.tgt == nullptr) return;

    // per-CTA partials in a fixed place, folded in a fixed order by the last CTA to finish: deterministic for a grid size
    if (threadIdx.x < 4) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kEpiWarps; ++w) s += sm.red[w][threadIdx.x];
        a.partials[(size_t)blockIdx.x * 4 + threadIdx.x] = s;
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) sm.last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!sm.last) return;
    __threadfence();
    if (threadIdx.x < 4) {
        double s = 0.0;
        for (unsigned b = 0; b < gridDim.x; ++b) s += __ldcg(a.partials + (size_t)b * 4 + threadIdx.x);
        sm.red[0][threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const double invn = (double)a.inv_n;
        const double coord = (double)a.lambda_coord * sm.red[0][0], cobj = sm.red[0][1];
        const double noobj = (double)a.lambda_noobj * sm.red[0][2], cls = sm.red[0][3];
        a.parts[0] = (float)((coord + cobj + noobj + cls) * invn);
        a.parts[1] = (float)(coord * invn);
        a.parts[2] = (float)(cobj * invn);
        a.parts[3] = (float)(noobj * invn);
        a.parts[4] = (float)(cls * invn);
        *a.ticket = 0u;
    }
}

// lo = w - (w with its low 13 bits cleared): the part of an fp32 weight that TF32 drops (exact in fp32)
__global__ void __launch_bounds__(256) split_lo_kernel(const uint4 *__restrict__ w, float4 *__restrict__ lo, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
        const uint4 v = __ldg(w + i);
        lo[i] = make_float4(__uint_as_float(v.x) - __uint_as_float(v.x & 0xFFFFE000u), __uint_as_float(v.y) - __uint_as_float(v.y & 0xFFFFE000u),
                            __uint_as_float(v.z) - __uint_as_float(v.z & 0xFFFFE000u), __uint_as_float(v.w) - __uint_as_float(v.w & 0xFFFFE000u));
    }
}

}  // namespace hl
}  // namespace yh

using namespace yh;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
    static std::atomic<EncodeTiledFn> cached{nullptr};
    EncodeTiledFn fn = cached.load();
    if (fn != nullptr) return fn;
    void *sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess || sym == nullptr ||
        q != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = (EncodeTiledFn)sym;
    cached.store(fn);
    return fn;
}

// rows x 4096 K-major matrix (fp16 / bf16 / fp32), boxes of one 128-byte K block x box_rows rows, 128-byte swizzle, zero fill outside
bool make_map(CUtensorMap *map, const void *base, int dtype, long long rows, int box_rows) {
    EncodeTiledFn fn = encode_tiled();
    if (fn == nullptr) return false;
    const int esize = (dtype == YH_F16 || dtype == YH_BF16) ? 2 : 4;
    const cuuint64_t dims[2] = {(cuuint64_t)hl::kK, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)hl::kK * esize};
    const cuuint32_t box[2] = {(cuuint32_t)(hl::kRowBytes / esize), (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = dtype == YH_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                   : dtype == YH_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    return fn(map, dt, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

constexpr int kMaxCtas = 256;   // partials are sized for this many CTAs (one per SM)

// CTAs that share a tile: 2 (an SM pair, tcgen05 cta_group::2) unless YH_HL_CTAS=1 asks for the single-CTA form
int ctas_per_tile() {
    static int cached = 0;
    if (cached == 0) {
        const char *e = getenv("YH_HL_CTAS");
        cached = (e != nullptr && e[0] == '1') ? 1 : 2;
    }
    return cached;
}

template <typename T, typename GT, int CTAS, int PASSES = 1>
int launch(const CUtensorMap &ma, const CUtensorMap &mb, const CUtensorMap &mblo, const hl::Args &a, int ctas, cudaStream_t st) {
    static std::atomic<bool> attr_done[kMaxDevices];
    const int smem = (int)sizeof(hl::SmemT<CTAS, PASSES>);
    auto kern = hl::head_linear_loss_kernel<T, GT, CTAS, PASSES>;
    if (first_use_on_device(attr_done)) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ctas);
    cfg.blockDim = dim3(hl::Cfg<CTAS, PASSES>::kThreads1);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CTAS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ma, mb, mblo, a);
    if (e != cudaSuccess) {
        set_error("yh_head_linear_loss: launch failed: %s", cudaGetErrorString(e));
        return (int)e;
    }
    return check_launch("yh_head_linear_loss");
}

template <int CTAS>
int dispatch(const CUtensorMap &ma, const CUtensorMap &mb, const CUtensorMap &ml, const hl::Args &a, int ctas, int dtype, bool g32,
             cudaStream_t st) {
    if (dtype == YH_F32) return launch<float, float, CTAS, 3>(ma, mb, ml, a, ctas, st);     // fp32 operands, 3xTF32: fp32's accuracy
    if (dtype == YH_TF32) return launch<float, float, CTAS, 1>(ma, mb, ml, a, ctas, st);    // fp32 operands, one TF32 pass
    if (dtype == YH_F16)
        return g32 ? launch<__half, float, CTAS>(ma, mb, ml, a, ctas, st) : launch<__half, __half, CTAS>(ma, mb, ml, a, ctas, st);
    return g32 ? launch<__nv_bfloat16, float, CTAS>(ma, mb, ml, a, ctas, st)
               : launch<__nv_bfloat16, __nv_bfloat16, CTAS>(ma, mb, ml, a, ctas, st);
}

}  // namespace

// ticket + per-CTA partial sums + (3xTF32) the lo part of the weight, split off once per call
static const size_t kWeightLoOffset = 256 + ((sizeof(double) * 4 * kMaxCtas + 255) & ~(size_t)255);
extern "C" size_t yh_head_linear_loss_workspace_bytes(void) {
    return kWeightLoOffset + sizeof(float) * (size_t)hl::kNOut * hl::kK;
}

extern "C" int yh_head_linear_loss(const void *x, const void *weight, const float *bias, int dtype, const float *tgt, int M,
                                   int S, int B, int C, int in_features, float lambda_coord, float lambda_noobj,
                                   float grad_scale, int round_pred, float *parts, void *grad, int grad_dtype,
                                   float *pred_out, void *workspace, size_t workspace_bytes, void *stream) {
    if (!shape_ok(S, B, C) || in_features != hl::kK) {
        set_error("yh_head_linear_loss: built for Linear(4096 -> 7*7*30), got in_features=%d S=%d B=%d C=%d", in_features, S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (dtype != YH_F16 && dtype != YH_BF16 && dtype != YH_F32 && dtype != YH_TF32) {
        set_error("yh_head_linear_loss: operands must be fp16, bf16 or fp32 (dtype %d)", dtype);
        return YH_ERR_UNSUPPORTED;
    }
    const bool narrow = dtype == YH_F16 || dtype == YH_BF16;
    if (grad != nullptr && grad_dtype != YH_F32 && !(narrow && grad_dtype == dtype)) {
        set_error("yh_head_linear_loss: the gradient is written in the operand type or in fp32 (grad_dtype %d)", grad_dtype);
        return YH_ERR_BAD_ARG;
    }
    if (M < 0 || !x || !weight || (tgt != nullptr && (!parts || !workspace)) || (tgt == nullptr && pred_out == nullptr)) {
        set_error("yh_head_linear_loss: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if ((tgt != nullptr || dtype == YH_F32) && (workspace == nullptr || workspace_bytes < yh_head_linear_loss_workspace_bytes())) {
        set_error("yh_head_linear_loss: workspace too small (or missing)");
        return YH_ERR_WORKSPACE;
    }
    if (((uintptr_t)x | (uintptr_t)weight) & 15u) {
        set_error("yh_head_linear_loss: activations and weight must be 16-byte aligned (TMA)");
        return YH_ERR_ALIGNMENT;
    }
    if ((((uintptr_t)tgt | (uintptr_t)pred_out) & 7u) || ((uintptr_t)grad & 7u)) {
        set_error("yh_head_linear_loss: targets, predictions and gradient must be 8-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    if (M == 0) return 0;
    const int per = ctas_per_tile();
    CUtensorMap ma, mb;
    if (!make_map(&ma, x, dtype, M, hl::kBM) || !make_map(&mb, weight, dtype, hl::kNOut, hl::kBN / per)) {
        set_error("yh_head_linear_loss: cuTensorMapEncodeTiled failed");
        return YH_ERR_UNSUPPORTED;
    }
    int dev = current_device(), sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int units = ((M + hl::kBM * per - 1) / (hl::kBM * per)) * hl::kNTiles;   // tiles of one CTA / one SM pair
    int ctas = units * per < sms ? units * per : sms - sms % per;
    if (ctas > kMaxCtas) ctas = kMaxCtas;
    hl::Args a;
    a.bias = bias;
    a.tgt = (const float2 *)tgt;
    a.M = M;
    a.round_pred = round_pred;
    a.lambda_coord = lambda_coord;
    a.lambda_noobj = lambda_noobj;
    a.inv_n = 1.0f / (float)M;
    a.grad_scale = grad_scale;
    a.grad = grad;
    a.pred_out = pred_out;
    a.ticket = (unsigned *)workspace;
    a.partials = workspace ? (double *)((char *)workspace + 256) : nullptr;
    a.parts = parts;
    cudaStream_t st = (cudaStream_t)stream;
    const bool g32 = grad == nullptr || grad_dtype == YH_F32;
    CUtensorMap ml = mb;
    if (dtype == YH_F32) {
        float *wlo = (float *)((char *)workspace + kWeightLoOffset);
        const size_t n4 = (size_t)hl::kNOut * hl::kK / 4;
        hl::split_lo_kernel<<<148 * 8, 256, 0, st>>>((const uint4 *)weight, (float4 *)wlo, n4);
        const int rc = check_launch("yh_head_linear_loss (weight split)");
        if (rc != 0) return rc;
        if (!make_map(&ml, wlo, dtype, hl::kNOut, hl::kBN / per)) {
            set_error("yh_head_linear_loss: cuTensorMapEncodeTiled failed");
            return YH_ERR_UNSUPPORTED;
        }
    }
    return per == 2 ? dispatch<2>(ma, mb, ml, a, ctas, dtype, g32, st) : dispatch<1>(ma, mb, ml, a, ctas, dtype, g32, st);
}
