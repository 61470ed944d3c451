// decode_nms.cu -- yh_decode_nms: batched decode + per-class NMS, one warp per image.
//
// Replaces mAPMetric._parse_predictions + _apply_nms (metrics.py:173-218, 258-296)
// and YOLOInference.parse_predictions + non_max_suppression (inference.py:141-210,
// 251-317).  Building blocks in head_image.cuh / head_nms.cuh.
#include "head_nms.cuh"

namespace yh {

constexpr int kWarpsPerCta = 4;
constexpr int kCtasPerSm = 7;

// Overlay of `stage` after NMS: the kept detections' sort keys, compact.
struct GlobalOrderOverlay {
    uint64_t key[kCand];
    uint8_t k[kPad];
};
static_assert(sizeof(GlobalOrderOverlay) <= sizeof(float2) * kHalfPairs, "overlay too large");

// one kept detection -> output slot `pos` of image `img`
__device__ __forceinline__ void emit_kept(const WarpWindow &w, int img, int pos, int k, uint8_t *__restrict__ keep_k,
                                          uint8_t *__restrict__ keep_cls, double *__restrict__ keep_conf,
                                          double *__restrict__ keep_box) {
    const size_t o = (size_t)img * kCand + pos;
    keep_k[o] = (uint8_t)k;
    if (keep_cls != nullptr) keep_cls[o] = w.cand.cls[k];
    if (keep_conf != nullptr) keep_conf[o] = w.cand.conf[k];
    if (keep_box != nullptr) {
        // (x, y, w, h) as the reference stores them (metrics.py:200-203, inference.py:184-193)
        const float2 wh = w.cand.wh[k];
        keep_box[4 * o + 0] = w.cand.xa[k];
        keep_box[4 * o + 1] = w.cand.ya[k];
        keep_box[4 * o + 2] = (double)wh.x;
        keep_box[4 * o + 3] = (double)wh.y;
    }
}

template <int MODE>
__global__ void __launch_bounds__(kWarpsPerCta * 32, kCtasPerSm)
decode_nms_kernel(const float *__restrict__ pred, int N, double conf_thr, double nms_thr, int order,
                  int32_t *__restrict__ n_keep, uint8_t *__restrict__ keep_k,
                  uint8_t *__restrict__ keep_cls, double *__restrict__ keep_conf,
                  double *__restrict__ keep_box, int32_t *__restrict__ n_cand,
                  uint8_t *__restrict__ bad_range, unsigned *__restrict__ ticket) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    WarpWindow *windows = reinterpret_cast<WarpWindow *>(smem_raw);
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    WarpWindow &w = windows[wib];
    const int warps_total = gridDim.x * kWarpsPerCta;

    // first image by position, further ones from the ticket counter (see map_update.cu)
    for (int img = blockIdx.x * kWarpsPerCta + wib; img < N;) {
        unsigned next_ticket = 0;
        if (lane == 0) next_ticket = atomicAdd(ticket, 1u);
        const float *pimg = pred + (size_t)img * kImgFloats;
        bool bad = false;
        const Mask128 valid = decode_image(w, pimg, lane, conf_thr, bad_range != nullptr ? &bad : nullptr);
        if (bad_range != nullptr) {
            const unsigned anybad = __ballot_sync(kFull, bad);
            if (lane == 0) bad_range[img] = anybad ? 1 : 0;
        }
        int n;
        const Mask128 keep = order_and_nms<MODE, true>(w, valid, nms_thr, lane, n);
        if (lane == 0) {
            n_keep[img] = keep.count();
            if (n_cand != nullptr) n_cand[img] = n;
        }
        if (order == YH_ORDER_CLASS_GROUPED) class_group_offsets(w, keep, lane);

        // global (confidence desc, candidate asc) order: rank each kept detection among the kept.
        // The kept are listed compactly (orderable key, candidate) so the rank loop is one 8-byte
        // shared load and an integer compare per rival.
        GlobalOrderOverlay &go = *reinterpret_cast<GlobalOrderOverlay *>(w.stage);
        int n_kept = 0;
        if (order == YH_ORDER_GLOBAL_CONF) {
            __syncwarp();
#pragma unroll
            for (int s = 0; s < kSlots; ++s) {
                const unsigned kw = keep.word(s);
                if ((kw >> lane) & 1u) {
                    const int k = w.order[lane + 32 * s];
                    const int q = n_kept + __popc(kw & ((1u << lane) - 1u));
                    // ascending composite = confidence descending, then candidate ascending
                    go.key[q] = ~orderable(w.cand.conf[k]);
                    go.k[q] = (uint8_t)k;
                }
                n_kept += __popc(kw);
            }
            __syncwarp();
        }

        if (order == YH_ORDER_GLOBAL_CONF) {
            // lanes take the compact kept list round-robin, so the rank loop runs ceil(kept / 32)
            // times however the kept boxes are spread over the sorted positions
#pragma unroll 1
            for (int q = lane; q < n_kept; q += 32) {
                const uint64_t mine = go.key[q];
                const int k = go.k[q];
                int pos = 0;
                for (int r = 0; r < n_kept; ++r) {
                    const uint64_t other = go.key[r];
                    pos += (other < mine) || (other == mine && go.k[r] < k);
                }
                emit_kept(w, img, pos, k, keep_k, keep_cls, keep_conf, keep_box);
            }
        } else {
#pragma unroll 1
            for (int s = 0; s < kSlots; ++s) {
                const int p = lane + 32 * s;
                if (!((keep.word(s) >> lane) & 1u)) continue;
                const int k = w.order[p];
                const int pos = w.cls_off[w.cand.cls[k]] + count_below(keep, p) - w.cls_kbase[w.cand.cls[k]];
                emit_kept(w, img, pos, k, keep_k, keep_cls, keep_conf, keep_box);
            }
        }
        __syncwarp();
        asm volatile("" : "+r"(next_ticket)::"memory");  // keep the (slow) ticket read down here
        const unsigned drawn = __shfl_sync(kFull, next_ticket, 0);
        img = drawn < (unsigned)N ? warps_total + (int)drawn : N;
    }
}

template <int MODE>
static int launch(const float *pred, int N, double conf_thr, double nms_thr, int order, int32_t *n_keep,
                  uint8_t *keep_k, uint8_t *keep_cls, double *keep_conf, double *keep_box,
                  int32_t *n_cand, uint8_t *bad_range, cudaStream_t st) {
    const size_t smem = sizeof(WarpWindow) * kWarpsPerCta;
    static std::atomic<bool> attr_done[kMaxDevices];
    if (first_use_on_device(attr_done)) {
        cudaFuncSetAttribute(decode_nms_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    int ctas = (N + kWarpsPerCta - 1) / kWarpsPerCta;
    const int cap = 148 * kCtasPerSm;
    if (ctas > cap) ctas = cap;
    unsigned *ticket = fresh_ticket(st);
    if (ticket == nullptr) return YH_ERR_UNSUPPORTED;
    decode_nms_kernel<MODE><<<ctas, kWarpsPerCta * 32, smem, st>>>(
        pred, N, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls, keep_conf, keep_box, n_cand, bad_range,
        ticket);
    return check_launch("yh_decode_nms");
}

}  // namespace yh

using namespace yh;

extern "C" int yh_decode_nms(const float *pred, int N, int S, int B, int C, double conf_thr,
                             double nms_thr, int iou_mode, int order, int32_t *n_keep,
                             uint8_t *keep_k, uint8_t *keep_cls, double *keep_conf,
                             double *keep_box, int32_t *n_cand, uint8_t *bad_range, void *stream) {
    if (!shape_ok(S, B, C)) {
        set_error("yh_decode_nms: only S=7,B=2,C=20 is built (got %d,%d,%d)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !pred || !n_keep || !keep_k) {
        set_error("yh_decode_nms: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if ((uintptr_t)pred % 8) {
        set_error("yh_decode_nms: pred must be 8-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    if (order != YH_ORDER_CLASS_GROUPED && order != YH_ORDER_GLOBAL_CONF) {
        set_error("yh_decode_nms: unknown order %d", order);
        return YH_ERR_BAD_ARG;
    }
    if (N == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (iou_mode == YH_IOU_METRICS)
        return launch<YH_IOU_METRICS>(pred, N, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls, keep_conf,
                                      keep_box, n_cand, bad_range, st);
    if (iou_mode == YH_IOU_INFER)
        return launch<YH_IOU_INFER>(pred, N, conf_thr, nms_thr, order, n_keep, keep_k, keep_cls, keep_conf,
                                    keep_box, n_cand, bad_range, st);
    set_error("yh_decode_nms: unknown iou_mode %d", iou_mode);
    return YH_ERR_BAD_ARG;
}
