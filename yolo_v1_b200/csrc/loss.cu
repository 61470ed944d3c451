// loss.cu -- fused YOLOv1 loss forward + backward (sm_100a).
//
// Replaces YOLOLoss.forward / compute_iou and their autograd backward
// (reference src/yolo/loss.py:55-172, 174-212).  One pass over the batch reads
// pred and target once and writes d(total)/d(pred) once: 17,640 B per image.
//
// Mapping: a CTA owns a tile of kTile consecutive cells (a cell = 30 floats).
// The tile is staged through shared memory with coalesced 8-byte accesses
// (a cell is 120 B, so 8 B is the widest alignment every cell start has); each
// thread then owns one cell, reads it with conflict-free LDS.64 (cell stride =
// 15 x 8 B, odd), and writes its gradient back in place for the coalesced store.
// Loss terms are reduced warp -> CTA -> (fp64) global partials; the last CTA to
// finish folds the partials in a fixed order, so results are deterministic.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace yh {

constexpr int kTile = 128;              // cells per CTA == threads per CTA
constexpr int kPairs = kD / 2;          // 15 two-element pairs per cell
constexpr int kTilePairs = kTile * kPairs;

// ---- element-pair global I/O for the three prediction dtypes ----------------
template <typename T> struct Pair;
// The tile sits in shared memory in the tensor's own dtype (so a bulk copy can put it there);
// a thread widens its cell when it reads it and narrows the gradient when it writes it back.
template <> struct Pair<float> {
    using G = float2;
    static __device__ __forceinline__ G load(const G *p) { return __ldg(p); }
    static __device__ __forceinline__ float2 widen(G v) { return v; }
    static __device__ __forceinline__ G narrow(float2 v) { return v; }
};
template <> struct Pair<__half> {
    using G = __half2;
    static __device__ __forceinline__ G load(const G *p) { return *p; }
    static __device__ __forceinline__ float2 widen(G v) { return __half22float2(v); }
    static __device__ __forceinline__ G narrow(float2 v) { return __float22half2_rn(v); }
};
template <> struct Pair<__nv_bfloat16> {
    using G = __nv_bfloat162;
    static __device__ __forceinline__ G load(const G *p) { return *p; }
    static __device__ __forceinline__ float2 widen(G v) { return __bfloat1622float2(v); }
    static __device__ __forceinline__ G narrow(float2 v) { return __float22bfloat162_rn(v); }
};

// ---- bulk asynchronous copies (TMA engine, cp.async.bulk) + the mbarrier they complete on ----
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gmem_src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void *gmem_dst, const void *smem_src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// torch.maximum / torch.minimum / clamp(min=0): NaN propagates.
__device__ __forceinline__ float t_max(float a, float b) { return (a > b || a != a) ? a : b; }
__device__ __forceinline__ float t_min(float a, float b) { return (a < b || a != a) ? a : b; }
__device__ __forceinline__ float t_relu(float v) { return (v < 0.f) ? 0.f : v; }

struct IouCtx {  // forward values the backward needs
    float iou, inter, U, iw, ih, rawx, rawy;
    float px1, px2, py1, py2;
};

// loss.py:191-212 in fp32, one rounding per operator (-fmad=false).
__device__ __forceinline__ IouCtx iou_fwd(float x, float y, float w, float h, float tx1, float tx2,
                                          float ty1, float ty2, float tarea) {
    IouCtx c;
    float hw = w * 0.5f, hh = h * 0.5f;
    c.px1 = x - hw; c.px2 = x + hw;
    c.py1 = y - hh; c.py2 = y + hh;
    c.rawx = t_min(c.px2, tx2) - t_max(c.px1, tx1);
    c.rawy = t_min(c.py2, ty2) - t_max(c.py1, ty1);
    c.iw = t_relu(c.rawx);
    c.ih = t_relu(c.rawy);
    c.inter = c.iw * c.ih;
    float uni = (w * h + tarea) - c.inter;
    c.U = uni + 1e-6f;
    c.iou = c.inter / c.U;
    return c;
}

// gradient of torch.maximum(self, other) w.r.t. self: 1, 1/2 at equality, 0.
__device__ __forceinline__ float max_mask(float self, float other) {
    return (self < other) ? 0.f : ((self == other) ? 0.5f : 1.f);
}
__device__ __forceinline__ float min_mask(float self, float other) {
    return (self > other) ? 0.f : ((self == other) ? 0.5f : 1.f);
}

#ifndef YH_LOSS_MINBLOCKS_BULK
#define YH_LOSS_MINBLOCKS_BULK 7
#endif
#ifndef YH_LOSS_MINBLOCKS
#define YH_LOSS_MINBLOCKS 5
#endif
// BULK (build with -DYH_LOSS_BULK): the tile is brought in by two bulk asynchronous copies (one thread arms an mbarrier and
// issues them, nobody's registers hold data in flight) and the gradient tile leaves by one bulk
// store; needs 16-byte aligned tensors and a full tile (a full tile is 15,360 / 7,680 B).  The
// other form stages with 8-byte loads and serves unaligned tensors and the ragged last tile.
template <typename T, bool BULK>
__global__ void __launch_bounds__(kTile, BULK ? YH_LOSS_MINBLOCKS_BULK : YH_LOSS_MINBLOCKS)
loss_kernel(const typename Pair<T>::G *__restrict__ pred, const float2 *__restrict__ tgt,
            typename Pair<T>::G *__restrict__ grad, long long n_cells, long long first_cell, float lambda_coord,
            float lambda_noobj, float inv_n, double *__restrict__ partials, unsigned part_base,
            unsigned total_ctas, unsigned int *__restrict__ ticket, float *__restrict__ parts) {
    using G = typename Pair<T>::G;
    __shared__ alignas(128) G s_pred[kTilePairs];
    __shared__ alignas(128) float2 s_tgt[kTilePairs];
    __shared__ double s_red[kTile / 32][4];
    __shared__ alignas(8) unsigned long long s_bar;
    __shared__ bool s_last;

    const int tid = threadIdx.x;
    const long long cell0 = first_cell + (long long)blockIdx.x * kTile;
    const int cells_here = (int)min((long long)kTile, n_cells - cell0);
    const int pairs_here = cells_here * kPairs;
    const long long pair0 = cell0 * kPairs;

    if (BULK) {
        if (tid == 0) {
            mbar_init(&s_bar, 1);
            fence_async_smem();  // the copy engine must see the initialised barrier
            mbar_expect_tx(&s_bar, (unsigned)(kTilePairs * (sizeof(G) + sizeof(float2))));
            bulk_load(s_pred, pred + pair0, (unsigned)(kTilePairs * sizeof(G)), &s_bar);
            bulk_load(s_tgt, tgt + pair0, (unsigned)(kTilePairs * sizeof(float2)), &s_bar);
        }
        __syncthreads();         // the barrier is initialised before anyone polls it
        mbar_wait(&s_bar, 0);
    } else {
        // stage the tile: 15 independent 8-byte loads per thread per tensor
#pragma unroll
        for (int i = 0; i < kPairs; ++i) {
            int p = tid + i * kTile;
            if (p < pairs_here) {
                s_pred[p] = Pair<T>::load(pred + pair0 + p);
                s_tgt[p] = __ldg(tgt + pair0 + p);
            }
        }
        __syncthreads();
    }

    float acc_coord = 0.f, acc_obj = 0.f, acc_noobj = 0.f, acc_cls = 0.f;
    if (tid < cells_here) {
        // the two boxes (5 pairs) live in registers; the 20 class values are streamed through
        // shared memory pair by pair further down, which keeps the kernel at ~half the registers
        constexpr int kBoxPairs = kB * 5 / 2;
        float p[kB * 5], t[kB * 5];
#pragma unroll
        for (int i = 0; i < kBoxPairs; ++i) {
            float2 a = Pair<T>::widen(s_pred[tid * kPairs + i]), b = s_tgt[tid * kPairs + i];
            p[2 * i] = a.x; p[2 * i + 1] = a.y;
            t[2 * i] = b.x; t[2 * i + 1] = b.y;
        }
        float g[kB * 5];
#pragma unroll
        for (int i = 0; i < kB * 5; ++i) g[i] = 0.f;

        const bool has0 = t[4] > 0.f, has1 = t[9] > 0.f;
        const bool obj = has0 || has1;                    // loss.py:98-99
        const int slot = (!has0 && has1) ? 1 : 0;         // loss.py:102 (first filled slot)
        const float gn = inv_n * lambda_noobj;
        if (!obj) {
            // no-object cell: both confidences are "noobj" terms (loss.py:150-153)
            acc_noobj = p[4] * p[4] + p[9] * p[9];
            g[4] = gn * (2.f * p[4]);
            g[9] = gn * (2.f * p[9]);
        } else {
            const float tx = slot ? t[5] : t[0], ty = slot ? t[6] : t[1];
            const float tw = slot ? t[7] : t[2], th = slot ? t[8] : t[3];
            const float thw = tw * 0.5f, thh = th * 0.5f;
            const float tx1 = tx - thw, tx2 = tx + thw, ty1 = ty - thh, ty2 = ty + thh;
            const float tarea = tw * th;
            IouCtx c0 = iou_fwd(p[0], p[1], p[2], p[3], tx1, tx2, ty1, ty2, tarea);
            IouCtx c1 = iou_fwd(p[5], p[6], p[7], p[8], tx1, tx2, ty1, ty2, tarea);
            // argmax over B=2 (loss.py:110): first maximum, NaN counts as maximal
            const bool pick1 = (c1.iou > c0.iou) || (c1.iou != c1.iou && c0.iou == c0.iou);
            const IouCtx c = pick1 ? c1 : c0;
            // responsible box (x, y, w, h, conf) and the other box's confidence, by select
            // (no runtime-indexed register arrays)
            const float x = pick1 ? p[5] : p[0], y = pick1 ? p[6] : p[1];
            const float w = pick1 ? p[7] : p[2], h = pick1 ? p[8] : p[3];
            const float cf = pick1 ? p[9] : p[4], co = pick1 ? p[4] : p[9];

            // --- coordinate term (loss.py:126-137)
            const float dx = x - tx, dy = y - ty;
            const float wc = t_max(w, 1e-6f), hc = t_max(h, 1e-6f);
            const float sw = sqrtf(wc), sh = sqrtf(hc);
            const float dw = sw - sqrtf(t_max(tw, 1e-6f)), dh = sh - sqrtf(t_max(th, 1e-6f));
            acc_coord = (dx * dx + dy * dy) + (dw * dw + dh * dh);
            // --- confidence terms (loss.py:142-153); IoU target is not detached
            const float dc = cf - c.iou;
            acc_obj = dc * dc;
            acc_noobj = co * co;

            // --- backward
            const float gc = inv_n * lambda_coord;
            const float g_other = gn * (2.f * co);
            const float g_resp = inv_n * (2.f * dc);
            float gx = gc * (2.f * dx), gy = gc * (2.f * dy);
            // sqrt'(v) = 1/(2 sqrt v); clamp(min) passes the gradient where w >= 1e-6
            float gw = (w >= 1e-6f) ? (gc * (2.f * dw)) / (2.f * sw) : 0.f;
            float gh = (h >= 1e-6f) ? (gc * (2.f * dh)) / (2.f * sh) : 0.f;
            // through the IoU (d conf_obj / d iou = -2 dc)
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
            g[0] = pick1 ? 0.f : gx; g[1] = pick1 ? 0.f : gy;
            g[2] = pick1 ? 0.f : gw; g[3] = pick1 ? 0.f : gh;
            g[4] = pick1 ? g_other : g_resp;
            g[5] = pick1 ? gx : 0.f; g[6] = pick1 ? gy : 0.f;
            g[7] = pick1 ? gw : 0.f; g[8] = pick1 ? gh : 0.f;
            g[9] = pick1 ? g_resp : g_other;
        }
        // --- class term (loss.py:156-157), summed in class order, and its gradient in place
        float cs = 0.f;
#pragma unroll
        for (int i = kBoxPairs; i < kPairs; ++i) {
            float2 gc2 = make_float2(0.f, 0.f);
            if (obj) {
                const float2 a = Pair<T>::widen(s_pred[tid * kPairs + i]), b = s_tgt[tid * kPairs + i];
                const float d0 = a.x - b.x, d1 = a.y - b.y;
                cs += d0 * d0;
                cs += d1 * d1;
                gc2 = make_float2(inv_n * (2.f * d0), inv_n * (2.f * d1));
            }
            if (grad != nullptr) s_pred[tid * kPairs + i] = Pair<T>::narrow(gc2);
        }
        acc_cls = cs;
        if (grad != nullptr) {
#pragma unroll
            for (int i = 0; i < kBoxPairs; ++i)
                s_pred[tid * kPairs + i] = Pair<T>::narrow(make_float2(g[2 * i], g[2 * i + 1]));
            if (BULK) fence_async_smem();  // my writes, visible to the copy engine
        }
    }

    // CTA reduction of the four terms
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        acc_coord += __shfl_down_sync(0xffffffffu, acc_coord, off);
        acc_obj += __shfl_down_sync(0xffffffffu, acc_obj, off);
        acc_noobj += __shfl_down_sync(0xffffffffu, acc_noobj, off);
        acc_cls += __shfl_down_sync(0xffffffffu, acc_cls, off);
    }
    if ((tid & 31) == 0) {
        s_red[tid >> 5][0] = acc_coord; s_red[tid >> 5][1] = acc_obj;
        s_red[tid >> 5][2] = acc_noobj; s_red[tid >> 5][3] = acc_cls;
    }
    __syncthreads();

    // The partials are published (four stores, fenced by the four threads that made them) BEFORE
    // the gradient tile goes out: a fence after 15 KB of stores would hold every thread until
    // all of them are acknowledged.
    if (tid < 4) {
        double s = 0.0;
#pragma unroll
        for (int wdx = 0; wdx < kTile / 32; ++wdx) s += s_red[wdx][tid];
        partials[((size_t)part_base + blockIdx.x) * 4 + tid] = s;
        __threadfence();
    }
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(ticket, 1u) == total_ctas - 1);

    if (grad != nullptr) {
        if (BULK) {
            if (tid == 0) {
                bulk_store(grad + pair0, s_pred, (unsigned)(kTilePairs * sizeof(G)));
                bulk_store_wait_read();  // shared memory may go once the engine has read it
            }
        } else {
#pragma unroll
            for (int i = 0; i < kPairs; ++i) {
                int q = tid + i * kTile;
                if (q < pairs_here) grad[pair0 + q] = s_pred[q];
            }
        }
    }
    __syncthreads();
    if (!s_last) return;

    // last CTA: fold all partials in a fixed order (deterministic)
    __threadfence();
    double a[4] = {0.0, 0.0, 0.0, 0.0};
    for (unsigned b = tid; b < total_ctas; b += kTile) {
#pragma unroll
        for (int k = 0; k < 4; ++k) a[k] += __ldcg(partials + (size_t)b * 4 + k);
    }
    double(*s_fin)[4] = reinterpret_cast<double(*)[4]>(s_tgt);  // the target tile is dead by now (4 KB of its 15)
#pragma unroll
    for (int k = 0; k < 4; ++k) s_fin[tid][k] = a[k];
    __syncthreads();
    for (int st = kTile / 2; st > 0; st >>= 1) {
        if (tid < st) {
#pragma unroll
            for (int k = 0; k < 4; ++k) s_fin[tid][k] += s_fin[tid + st][k];
        }
        __syncthreads();
    }
    if (tid == 0) {
        double coord = (double)lambda_coord * s_fin[0][0];
        double cobj = s_fin[0][1];
        double noobj = (double)lambda_noobj * s_fin[0][2];
        double cls = s_fin[0][3];
        double invn = (double)inv_n;
        parts[0] = (float)((coord + cobj + noobj + cls) * invn);
        parts[1] = (float)(coord * invn);
        parts[2] = (float)(cobj * invn);
        parts[3] = (float)(noobj * invn);
        parts[4] = (float)(cls * invn);
        *ticket = 0u;  // leave the workspace ready for the next call
    }
}

template <typename T>
__global__ void scale_kernel(T *__restrict__ g, size_t n, const float *__restrict__ scale) {
    const float s = __ldg(scale);
    if (s == 1.0f) return;  // plain fp32 training: upstream gradient is exactly 1
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (size_t)gridDim.x * blockDim.x)
        g[i] = (T)((float)g[i] * s);
}

}  // namespace yh

using namespace yh;

namespace yh {
struct LossLaunch {
    const float2 *tgt;
    long long cells;
    float lambda_coord, lambda_noobj, inv_n;
    double *partials;
    unsigned int *ticket;
    float *parts;
    unsigned blocks, full;  // CTAs in all, CTAs the bulk kernel takes
    cudaStream_t st;
};
template <typename T>
static void launch_loss(const LossLaunch &L, const void *pred, void *grad) {
    using G = typename Pair<T>::G;
    if (L.full > 0)
        loss_kernel<T, true><<<L.full, kTile, 0, L.st>>>((const G *)pred, L.tgt, (G *)grad, L.cells, 0, L.lambda_coord,
                                                         L.lambda_noobj, L.inv_n, L.partials, 0, L.blocks, L.ticket,
                                                         L.parts);
    if (L.blocks > L.full)
        loss_kernel<T, false><<<L.blocks - L.full, kTile, 0, L.st>>>(
            (const G *)pred, L.tgt, (G *)grad, L.cells, (long long)L.full * kTile, L.lambda_coord, L.lambda_noobj,
            L.inv_n, L.partials, L.full, L.blocks, L.ticket, L.parts);
}
}  // namespace yh

extern "C" size_t yh_loss_workspace_bytes(int N, int S, int B, int C) {
    if (N < 0) return 0;
    if (!shape_ok(S, B, C)) return (S >= 1 && B >= 1 && B <= 8 && C >= 1) ? generic_loss_workspace_bytes(N, S) : 0;
    size_t cells = (size_t)N * kCells;
    size_t blocks = (cells + kTile - 1) / kTile;
    return 256 + blocks * 4 * sizeof(double);
}

extern "C" int yh_loss_fwd_bwd(const void *pred, int pred_dtype, const float *tgt, int N, int S,
                               int B, int C, float lambda_coord, float lambda_noobj, float *parts,
                               void *grad_pred, void *workspace, size_t workspace_bytes,
                               void *stream) {
    const bool fast = shape_ok(S, B, C);
    if (!fast && !(S >= 1 && B >= 1 && B <= 8 && C >= 1)) {
        set_error("yh_loss_fwd_bwd: shape S=%d,B=%d,C=%d outside this build (B <= 8)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (N <= 0 || !pred || !tgt || !parts || !workspace) {
        set_error("yh_loss_fwd_bwd: bad argument (N=%d)", N);
        return YH_ERR_BAD_ARG;
    }
    if (workspace_bytes < yh_loss_workspace_bytes(N, S, B, C)) {
        set_error("yh_loss_fwd_bwd: workspace too small");
        return YH_ERR_WORKSPACE;
    }
    if (!fast) {   // any other head shape: one thread per cell (generic.cu)
        if (pred_dtype != YH_F32 && pred_dtype != YH_F16 && pred_dtype != YH_BF16) {
            set_error("yh_loss_fwd_bwd: unknown pred_dtype %d", pred_dtype);
            return YH_ERR_BAD_ARG;
        }
        return generic_loss(pred, pred_dtype, tgt, N, S, B, C, lambda_coord, lambda_noobj, parts, grad_pred, workspace,
                            (cudaStream_t)stream);
    }
    const size_t esz = (pred_dtype == YH_F32) ? 4 : 2;
    if (((uintptr_t)pred % (2 * esz)) || ((uintptr_t)tgt % 8) || ((uintptr_t)grad_pred % (2 * esz)) ||
        ((uintptr_t)workspace % 256)) {
        set_error("yh_loss_fwd_bwd: pointers must be aligned to one element pair (workspace: 256 B)");
        return YH_ERR_ALIGNMENT;
    }
    cudaStream_t st = (cudaStream_t)stream;
    long long cells = (long long)N * kCells;
    unsigned blocks = (unsigned)((cells + kTile - 1) / kTile);
    unsigned int *ticket = (unsigned int *)workspace;
    double *partials = (double *)((char *)workspace + 256);
    cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) {
        set_error("yh_loss_fwd_bwd: memset: %s", cudaGetErrorString(e));
        return (int)e;
    }
    const float inv_n = 1.0f / (float)N;
    // full tiles through the bulk-copy kernel when every tensor is 16-byte aligned; the ragged
    // last tile (and unaligned tensors) through the 8-byte staging kernel.  Both share the ticket.
    const bool aligned16 = !((uintptr_t)pred % 16) && !((uintptr_t)tgt % 16) && !((uintptr_t)grad_pred % 16);
    LossLaunch L;
    L.tgt = (const float2 *)tgt; L.cells = cells; L.lambda_coord = lambda_coord; L.lambda_noobj = lambda_noobj;
    L.inv_n = inv_n; L.partials = partials; L.ticket = ticket; L.parts = parts; L.blocks = blocks;
    L.full = aligned16 ? (unsigned)(cells / kTile) : 0u; L.st = st;
#ifndef YH_LOSS_BULK
    // Measured on B200 (65,536 images, fp32): bulk copies 0.2395 ms, 8-byte staging 0.2333 ms.  The
    // staging kernel already keeps ~200 KB of loads in flight per SM, so the copy engine buys
    // nothing here and costs a barrier round trip per tile; it stays selectable for experiments.
    L.full = 0;
#endif
    switch (pred_dtype) {
        case YH_F32: launch_loss<float>(L, pred, grad_pred); break;
        case YH_F16: launch_loss<__half>(L, pred, grad_pred); break;
        case YH_BF16: launch_loss<__nv_bfloat16>(L, pred, grad_pred); break;
        default:
            set_error("yh_loss_fwd_bwd: unknown pred_dtype %d", pred_dtype);
            return YH_ERR_BAD_ARG;
    }
    return check_launch("yh_loss_fwd_bwd", (L.full > 0 ? 1 : 0) + (L.blocks > L.full ? 1 : 0));
}

extern "C" int yh_scale_inplace(void *grad, int dtype, size_t n_elems, const float *scale,
                                void *stream) {
    if (!grad || !scale) {
        set_error("yh_scale_inplace: null pointer");
        return YH_ERR_BAD_ARG;
    }
    if (n_elems == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned blocks = (unsigned)min((size_t)148 * 16, (n_elems + 255) / 256);
    switch (dtype) {
        case YH_F32: scale_kernel<float><<<blocks, 256, 0, st>>>((float *)grad, n_elems, scale); break;
        case YH_F16: scale_kernel<__half><<<blocks, 256, 0, st>>>((__half *)grad, n_elems, scale); break;
        case YH_BF16:
            scale_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((__nv_bfloat16 *)grad, n_elems, scale);
            break;
        default:
            set_error("yh_scale_inplace: unknown dtype %d", dtype);
            return YH_ERR_BAD_ARG;
    }
    return check_launch("yh_scale_inplace");
}

// ------------------------------------------------------------------ yh_loss_fwd_bwd_static
// The same loss forward + backward for callers that cannot allocate (code under CUDA-graph capture, tight serving
// loops, harnesses that forbid cudaMalloc): the scratch yh_loss_fwd_bwd takes from its caller -- the ticket word
// and the per-CTA partial sums -- lives in a file-scope device array instead.  No counterpart in the
// reference.
//
//   * kStaticScratchBytes = 128 KiB per GPU that loads the library, held until the process exits.  A call whose
//     yh_loss_workspace_bytes(N, S, B, C) exceeds it is rejected (S=7, B=2, C=20: N <= 10,678 images).
//   * In HOST code `g_loss_scratch` is a host-side stand-in, not a device address (a kernel handed it faults, or
//     silently uses the wrong memory where the GPU can reach host memory).  The device address comes from
//     cudaGetSymbolAddress, looked up once per device with that device current, return code checked.
//   * Every call on a device shares the array, and it keeps the previous call's contents: the ticket word is cleared
//     on the stream at the start of every call and every partial that is read has been written by the same call,
//     so calls that follow each other on one stream (or are otherwise ordered) are independent.  Calls that
//     OVERLAP on the GPU -- two streams, or a graph replay next to an eager call -- race on it and give wrong sums.
//   * Calls no cudaMalloc / cudaFree / allocating library routine; enqueues on the caller's stream only.
namespace yh {
constexpr size_t kStaticScratchBytes = 128 * 1024;
static __device__ __align__(256) unsigned char g_loss_scratch[kStaticScratchBytes];
static std::atomic<void *> g_loss_scratch_ptr[kMaxDevices];   // device address of the array, per device
}  // namespace yh

extern "C" size_t yh_loss_static_scratch_bytes(void) { return kStaticScratchBytes; }

extern "C" int yh_loss_fwd_bwd_static(const void *pred, int pred_dtype, const float *tgt, int N, int S, int B, int C,
                                      float lambda_coord, float lambda_noobj, float *parts, void *grad_pred,
                                      void *stream) {
    if (N <= 0 || !pred || !tgt || !parts) {
        set_error("yh_loss_fwd_bwd_static: bad argument (N=%d)", N);
        return YH_ERR_BAD_ARG;
    }
    const size_t need = yh_loss_workspace_bytes(N, S, B, C);
    if (need == 0) {
        set_error("yh_loss_fwd_bwd_static: shape S=%d,B=%d,C=%d outside this build (B <= 8)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (need > kStaticScratchBytes) {
        set_error("yh_loss_fwd_bwd_static: N=%d images of %dx%d cells need %zu bytes of scratch, the built-in array holds "
                  "%zu (S=7: at most 10,678 images per call); split the batch or call yh_loss_fwd_bwd with a workspace",
                  N, S, S, need, kStaticScratchBytes);
        return YH_ERR_WORKSPACE;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        set_error("yh_loss_fwd_bwd_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        cudaGetLastError();
        return e != cudaSuccess ? (int)e : YH_ERR_UNSUPPORTED;
    }
    void *scratch = g_loss_scratch_ptr[dev].load(std::memory_order_acquire);
    if (scratch == nullptr) {   // first call on this device: this device's copy of the symbol
        e = cudaGetSymbolAddress(&scratch, g_loss_scratch);
        if (e != cudaSuccess || scratch == nullptr) {
            set_error("yh_loss_fwd_bwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            cudaGetLastError();
            return e != cudaSuccess ? (int)e : YH_ERR_UNSUPPORTED;
        }
        g_loss_scratch_ptr[dev].store(scratch, std::memory_order_release);
    }
    // the ticket word is cleared on the stream by yh_loss_fwd_bwd itself (cudaMemsetAsync, no allocation)
    return yh_loss_fwd_bwd(pred, pred_dtype, tgt, N, S, B, C, lambda_coord, lambda_noobj, parts, grad_pred, scratch,
                           kStaticScratchBytes, stream);
}
