// list_form.cu -- yh_list_form: turns the compact records of yh_map_update into the
// reference's list form, only when somebody asks for all_predictions /
// all_ground_truths (mAPMetric keeps its state as two Python lists,
// metrics.py:70-71, 99-100).  Off the hot path: one warp per image.
//
//   rec_pos[n*K+q]  position of record q inside the reference's per-image list:
//                   class groups in the order their best member appears in the
//                   confidence-sorted list (metrics.py:274-279), records of a group in
//                   record order (they are already confidence-descending)
//   rec_box / gt_box  (x_abs, y_abs, w, h) in fp64 exactly as metrics.py:200-203,
//                   247-248 build them: (col + x) / S, (row + y) / S, w, h
#include "head_image.cuh"

namespace yh {

__global__ void __launch_bounds__(128)
list_form_kernel(const int32_t *__restrict__ n_keep, const uint64_t *__restrict__ rec_key,
                 const uint64_t *__restrict__ rec_flags, const uint8_t *__restrict__ rec_k,
                 const uint64_t *__restrict__ rec_box3, const int32_t *__restrict__ n_gt,
                 const uint8_t *__restrict__ gt_cell, const float4 *__restrict__ gt_raw, int N,
                 uint8_t *__restrict__ rec_pos, double *__restrict__ rec_box, double *__restrict__ gt_box) {
    __shared__ int s_first[4][32], s_count[4][32], s_off[4][32];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int img = blockIdx.x * 4 + wib;
    if (img >= N) return;
    const int n = n_keep[img];
    const size_t base = (size_t)img * kCand;
    s_first[wib][lane] = 1 << 30;
    s_count[wib][lane] = 0;
    __syncwarp();
    for (int q = lane; q < n; q += 32) {
        const int c = (int)(rec_flags[base + q] >> 56) & 31;
        atomicMin(&s_first[wib][c], q);
        atomicAdd(&s_count[wib][c], 1);
    }
    __syncwarp();
    // lane c ranks class c among the classes present: by its first record's
    // (confidence desc, candidate asc); key order is confidence descending
    const int first = s_first[wib][lane], cnt = s_count[wib][lane];
    const bool present = cnt > 0;
    const uint64_t hkey = present ? rec_key[base + first] : 0;
    const int hk = present ? rec_k[base + first] : 0;
    int off = 0;
    for (unsigned m = __ballot_sync(kFull, present); m; m &= m - 1) {
        const int o = __ffs(m) - 1;
        const uint64_t okey = __shfl_sync(kFull, hkey, o);
        const int ok = __shfl_sync(kFull, hk, o);
        const int on = __shfl_sync(kFull, cnt, o);
        if (okey < hkey || (okey == hkey && ok < hk)) off += on;
    }
    s_off[wib][lane] = off;
    __syncwarp();
    for (int q = lane; q < n; q += 32) {
        const int c = (int)(rec_flags[base + q] >> 56) & 31;
        rec_pos[base + q] = (uint8_t)(s_off[wib][c] + q - s_first[wib][c]);
        const uint64_t *r = rec_box3 + 3 * (base + q);  // (x_abs, y_abs) fp64, (w, h) fp32
        const uint64_t wh = r[2];
        double *o = rec_box + 4 * (base + q);
        o[0] = __longlong_as_double((long long)r[0]);
        o[1] = __longlong_as_double((long long)r[1]);
        o[2] = (double)__uint_as_float((unsigned)wh);
        o[3] = (double)__uint_as_float((unsigned)(wh >> 32));
    }
    const int ng = n_gt[img];
    const size_t gb = (size_t)img * kCells;
    for (int g = lane; g < ng; g += 32) {
        const int cell = gt_cell[gb + g];
        const float4 r = gt_raw[gb + g];
        double *o = gt_box + 4 * (gb + g);
        o[0] = abs_coord(cell % kS, r.x);
        o[1] = abs_coord(cell / kS, r.y);
        o[2] = (double)r.z;
        o[3] = (double)r.w;
    }
}

}  // namespace yh

using namespace yh;

extern "C" int yh_list_form(const int32_t *n_keep, const uint64_t *rec_key, const uint64_t *rec_flags,
                            const uint8_t *rec_k, const void *rec_box3, const int32_t *n_gt,
                            const uint8_t *gt_cell, const float *gt_raw, int N, int S, int B, int C,
                            uint8_t *rec_pos, double *rec_box, double *gt_box, void *stream) {
    if (!shape_ok(S, B, C)) {
        set_error("yh_list_form: only S=7,B=2,C=20 is built (got %d,%d,%d)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !n_keep || !rec_key || !rec_flags || !rec_k || !rec_box3 || !n_gt || !gt_cell || !gt_raw ||
        !rec_pos || !rec_box || !gt_box) {
        set_error("yh_list_form: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if (N == 0) return 0;
    list_form_kernel<<<(N + 3) / 4, 128, 0, (cudaStream_t)stream>>>(
        n_keep, rec_key, rec_flags, rec_k, (const uint64_t *)rec_box3, n_gt, gt_cell, (const float4 *)gt_raw, N,
        rec_pos, rec_box, gt_box);
    return check_launch("yh_list_form");
}
