// encode.cu -- yh_encode_targets: boxes + class ids -> (S, S, B*5 + C) target tensors.
//
// Replaces VOCDetectionYOLO._encode_target (reference src/yolo/dataset.py:487-532) for a
// whole batch: the step that feeds the loss / metric kernels, done per sample in
// Python by the reference's DataLoader.  One warp per image:
//   - the image's 5,880 bytes are zero-filled with coalesced 8-byte stores;
//   - every box finds its cell in doubles exactly as the reference does
//     (i = int(S * y), j = int(S * x), clamped to S - 1; negative indices wrap like
//     Python's); the FIRST box of a cell in list order owns it (dataset.py:518:
//     `if target[i, j, 4] == 0`), decided with a shared-memory atomicMin over box index;
//   - owners write x_cell = S*x - j, y_cell = S*y - i (rounded to fp32 on store, as the
//     assignment into a float32 tensor does), w, h, confidence 1 and the class one-hot.
// A box the reference would raise on (NaN / infinite centre: int() raises; an index
// outside [-S, S); a class outside [-D + B*5, C)) sets bad[n] and is skipped.
#include "common.cuh"

namespace yh {

constexpr int kEncWarps = 4;

// min(int(v), S - 1) for a finite double v without overflowing an int: everything at or
// above S clamps to S - 1, everything below -S is out of range anyway (reported as -S - 1).
__device__ __forceinline__ int grid_index(double v) {
    if (v >= (double)kS) return kS - 1;
    if (v <= -(double)kS - 1.0) return -kS - 1;
    const int i = (int)v;  // truncation toward zero, like int()
    return i < kS - 1 ? i : kS - 1;
}
__device__ __forceinline__ bool is_finite(double v) { return fabs(v) <= 1.7976931348623157e308; }

__global__ void __launch_bounds__(kEncWarps * 32)
encode_targets_kernel(const double *__restrict__ boxes, const int32_t *__restrict__ cls,
                      const int64_t *__restrict__ off, int N, float *__restrict__ out,
                      uint8_t *__restrict__ bad) {
    __shared__ int s_owner[kEncWarps][kCells];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int img = blockIdx.x * kEncWarps + wib;
    if (img >= N) return;
    int *owner = s_owner[wib];
    for (int c = lane; c < kCells; c += 32) owner[c] = 0x7fffffff;
    float2 *dst = reinterpret_cast<float2 *>(out + (size_t)img * kImgFloats);
    for (int q = lane; q < kImgFloats / 2; q += 32) dst[q] = make_float2(0.f, 0.f);
    __syncwarp();
    const int64_t lo = off[img], hi = off[img + 1];
    bool any_bad = false;
    // pass 1: the first box (list order) of every cell
    for (int64_t b = lo + lane; b < hi; b += 32) {
        const double x = boxes[4 * b + 0], y = boxes[4 * b + 1];
        const double fy = __dmul_rn((double)kS, y), fx = __dmul_rn((double)kS, x);  // self.S * y_center
        // int() raises on NaN / infinity; indices outside [-S, S) raise IndexError
        const bool finite = is_finite(fx) && is_finite(fy);
        int i = finite ? grid_index(fy) : 0, j = finite ? grid_index(fx) : 0;
        const int c = cls[b];
        const int ci = kB * 5 + c;  // Python index into the last axis
        const bool ok = finite && i >= -kS && j >= -kS && ci >= -kD && ci < kD;
        if (!ok) {
            any_bad = true;
            continue;
        }
        if (i < 0) i += kS;
        if (j < 0) j += kS;
        atomicMin(&owner[i * kS + j], (int)(b - lo));
    }
    __syncwarp();
    // pass 2: owners write their cell
    for (int64_t b = lo + lane; b < hi; b += 32) {
        const double x = boxes[4 * b + 0], y = boxes[4 * b + 1];
        const double fy = __dmul_rn((double)kS, y), fx = __dmul_rn((double)kS, x);
        const bool finite = is_finite(fx) && is_finite(fy);
        int i = finite ? grid_index(fy) : 0, j = finite ? grid_index(fx) : 0;
        const int c = cls[b];
        int ci = kB * 5 + c;
        if (!(finite && i >= -kS && j >= -kS && ci >= -kD && ci < kD)) continue;
        // cell-relative coordinates use the index BEFORE Python's wrap-around (dataset.py:514-515)
        const double x_cell = __dsub_rn(fx, (double)j), y_cell = __dsub_rn(fy, (double)i);
        if (i < 0) i += kS;
        if (j < 0) j += kS;
        if (owner[i * kS + j] != (int)(b - lo)) continue;
        if (ci < 0) ci += kD;
        float *cell = out + (size_t)img * kImgFloats + (i * kS + j) * kD;
        cell[0] = (float)x_cell;
        cell[1] = (float)y_cell;
        cell[2] = (float)boxes[4 * b + 2];
        cell[3] = (float)boxes[4 * b + 3];
        cell[4] = 1.0f;
        cell[ci] = 1.0f;  // written last, as the reference does
    }
    if (bad != nullptr) {
        const unsigned m = __ballot_sync(0xffffffffu, any_bad);
        if (lane == 0) bad[img] = m ? 1 : 0;
    }
}

}  // namespace yh

using namespace yh;

extern "C" int yh_encode_targets(const double *boxes, const int32_t *cls, const int64_t *offsets, int N,
                                 int S, int B, int C, float *out, uint8_t *bad, void *stream) {
    if (!shape_ok(S, B, C)) {
        set_error("yh_encode_targets: only S=7,B=2,C=20 is built (got %d,%d,%d)", S, B, C);
        return YH_ERR_UNSUPPORTED;
    }
    if (N < 0 || !offsets || !out) {
        set_error("yh_encode_targets: bad argument");
        return YH_ERR_BAD_ARG;
    }
    if ((uintptr_t)out % 8) {
        set_error("yh_encode_targets: out must be 8-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    if (N == 0) return 0;
    encode_targets_kernel<<<(N + kEncWarps - 1) / kEncWarps, kEncWarps * 32, 0, (cudaStream_t)stream>>>(
        boxes, cls, offsets, N, out, bad);
    return check_launch("yh_encode_targets");
}
