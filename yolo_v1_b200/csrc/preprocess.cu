// preprocess.cu -- yh_preprocess_rgb8: the image preprocessing of YOLOInference on the GPU.
//
// Replaces `self.transform(image)` of the reference (src/yolo/inference.py:57-66, 88-101): torchvision
// Resize((448, 448)) on a PIL image, ToTensor(), Normalize(mean, std).  The arithmetic is Pillow's
// (src/libImaging/Resample.c: precompute_coeffs in doubles, normalize_coeffs_8bpc to 22-bit fixed point,
// a horizontal then a vertical 8-bit pass with an 8-bit image in between) and torchvision's
// (uint8 -> float32, / 255, - mean, / std, each rounded to float32); results are bit-identical
// (oracle/preprocess_oracle.py, tests/test_preprocess.py).  Byte work bounded by HBM: SURVEY.md 8f row 2.
#include "common.cuh"

namespace yh {

constexpr int kPrecisionBits = 32 - 8 - 2;

struct AxisPlan {
    int in_size, out_size, ksize;
};
static AxisPlan axis_plan(int in_size, int out_size) {
    AxisPlan p;
    p.in_size = in_size;
    p.out_size = out_size;
    double filterscale = (double)in_size / (double)out_size;
    if (filterscale < 1.0) filterscale = 1.0;
    const double support = 1.0 * filterscale;  // bilinear: support 1.0
    // taps per output position (Pillow's ksize), padded to a multiple of four: rows of the tap table are
    // read as int4, and the taps past the last real one are zero
    p.ksize = ((int)ceil(support) * 2 + 1 + 3) & ~3;
    return p;
}

__device__ __forceinline__ double bilinear_filter(double x) {
    if (x < 0.0) x = -x;
    return x < 1.0 ? __dsub_rn(1.0, x) : 0.0;
}

// precompute_coeffs + normalize_coeffs_8bpc for one axis: thread xx owns output position xx.
// bounds[2*xx] = first input index, bounds[2*xx + 1] = number of taps; kk[xx * ksize + x] = tap x.
struct AxisArgs {
    int in_size, out_size, ksize;
    int *bounds, *kk;
};
// blockIdx.y picks the axis (both tables in one launch; an axis with out_size 0 is skipped)
__global__ void __launch_bounds__(128)
pp_coeffs(const AxisArgs ax0, const AxisArgs ax1) {
    const AxisArgs &ax = blockIdx.y ? ax1 : ax0;
    const int in_size = ax.in_size, out_size = ax.out_size, ksize = ax.ksize;
    int *__restrict__ bounds = ax.bounds, *__restrict__ kk = ax.kk;
    const int xx = blockIdx.x * 128 + threadIdx.x;
    if (xx >= out_size) return;
    const double scale = __ddiv_rn((double)in_size, (double)out_size);
    const double filterscale = scale < 1.0 ? 1.0 : scale;
    const double support = filterscale;
    const double ss = __ddiv_rn(1.0, filterscale);
    const double center = __dmul_rn(__dadd_rn((double)xx, 0.5), scale);
    int xmin = (int)__dadd_rn(__dsub_rn(center, support), 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)__dadd_rn(__dadd_rn(center, support), 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x)
        ww = __dadd_rn(ww, bilinear_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ss)));
    for (int x = 0; x < ksize; ++x) {
        double v = 0.0;
        if (x < xmax) {
            v = bilinear_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ss));
            if (ww != 0.0) v = __ddiv_rn(v, ww);
        }
        const double scaled = __dmul_rn(v, (double)(1 << kPrecisionBits));
        kk[(size_t)xx * ksize + x] = v < 0.0 ? (int)__dadd_rn(-0.5, scaled) : (int)__dadd_rn(0.5, scaled);
    }
    bounds[2 * xx] = xmin;
    bounds[2 * xx + 1] = xmax;
}

__device__ __forceinline__ unsigned clip8(int v) {
    v >>= kPrecisionBits;
    return (unsigned)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// Horizontal pass: (H, W, 3) -> (H, OW, 3), one thread per output pixel; the taps of neighbouring
// output pixels overlap, so a warp's loads fall into a few consecutive lines.
__global__ void __launch_bounds__(256)
pp_horizontal(const uint8_t *__restrict__ in, int H, int W, int OW, int ksize, const int *__restrict__ bounds,
              const int *__restrict__ kk, uint8_t *__restrict__ out) {
    in += (size_t)blockIdx.z * H * W * 3;
    out += (size_t)blockIdx.z * H * OW * 3;
    const int xx = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    if (xx >= OW) return;
    const int x0 = bounds[2 * xx], n = bounds[2 * xx + 1];
    const int *k = kk + (size_t)xx * ksize;
    const uint8_t *row = in + ((size_t)y * W + x0) * 3;
    int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
    for (int x = 0; x < n; ++x) {
        const int c = __ldg(k + x);
        a0 += (int)__ldg(row + 3 * x) * c;
        a1 += (int)__ldg(row + 3 * x + 1) * c;
        a2 += (int)__ldg(row + 3 * x + 2) * c;
    }
    uint8_t *o = out + ((size_t)y * OW + xx) * 3;
    o[0] = (uint8_t)clip8(a0);
    o[1] = (uint8_t)clip8(a1);
    o[2] = (uint8_t)clip8(a2);
}

// The same pass with the input staged through shared memory: a CTA owns kHSeg consecutive output pixels of
// one row, copies the span of input bytes they tap with aligned 4-byte loads (a row starts at an arbitrary
// byte), and every thread then reads its taps from shared memory.  Used whenever the span fits 48 KB.
constexpr int kHSeg = 128;
__global__ void __launch_bounds__(kHSeg)
pp_horizontal_staged(const uint8_t *__restrict__ in, int H, int W, int OW, int ksize, const int *__restrict__ bounds,
                     const int *__restrict__ kk, uint8_t *__restrict__ out) {
    extern __shared__ uint32_t s_words[];
    in += (size_t)blockIdx.z * H * W * 3;    // image blockIdx.z of the batch
    out += (size_t)blockIdx.z * H * OW * 3;
    const int xx0 = blockIdx.x * kHSeg, y = blockIdx.y, tid = threadIdx.x;
    const int xlast = min(xx0 + kHSeg, OW) - 1;
    const int p0 = bounds[2 * xx0], p1 = bounds[2 * xlast] + bounds[2 * xlast + 1];  // input pixels [p0, p1)
    const uintptr_t first = (uintptr_t)in + ((size_t)y * W + p0) * 3, last = (uintptr_t)in + ((size_t)y * W + p1) * 3;
    const uintptr_t a0 = first & ~(uintptr_t)3;
    const uintptr_t buf_end = (uintptr_t)in + (size_t)H * W * 3;
    const int nwords = (int)((last - a0 + 3) >> 2);
    if (a0 + 4 * (uintptr_t)nwords <= buf_end) {
        // four independent loads per thread and round: the copy is latency-bound otherwise
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a0);
        int i = tid;
        for (; i + 3 * kHSeg < nwords; i += 4 * kHSeg) {
            const uint32_t v0 = __ldg(src + i), v1 = __ldg(src + i + kHSeg), v2 = __ldg(src + i + 2 * kHSeg),
                           v3 = __ldg(src + i + 3 * kHSeg);
            s_words[i] = v0;
            s_words[i + kHSeg] = v1;
            s_words[i + 2 * kHSeg] = v2;
            s_words[i + 3 * kHSeg] = v3;
        }
        for (; i < nwords; i += kHSeg) s_words[i] = __ldg(src + i);
    } else {  // the last words of the image: the one that would run past its end is read bytewise
        for (int i = tid; i < nwords; i += kHSeg) {
            const uintptr_t a = a0 + 4 * (uintptr_t)i;
            uint32_t v;
            if (a + 4 <= buf_end) {
                v = __ldg(reinterpret_cast<const uint32_t *>(a));
            } else {
                v = 0;
                for (int b = 0; b < 4 && a + b < buf_end; ++b)
                    v |= (uint32_t)__ldg(reinterpret_cast<const uint8_t *>(a + b)) << (8 * b);
            }
            s_words[i] = v;
        }
    }
    __syncthreads();
    const int xx = xx0 + tid;
    if (xx >= OW) return;
    const int x0 = bounds[2 * xx], n = bounds[2 * xx + 1];
    const int *k = kk + (size_t)xx * ksize;
    // The taps of this pixel are 3 * n consecutive bytes at an arbitrary byte offset of the staged span.  They are
    // read as words and realigned with a funnel shift: three realigned words hold exactly four RGB pixels, and the
    // four taps that go with them are one int4 of the (zero-padded) tap row.  A quarter of the shared-memory
    // and tap loads of the byte-wise loop (ncu had it at 13.7 warps per issue in mio_throttle).
    const unsigned boff = (unsigned)(first - a0) + (unsigned)(x0 - p0) * 3u;
    const uint32_t *wp = s_words + (boff >> 2);
    const unsigned sh = (boff & 3u) * 8u;
    const int4 *k4 = reinterpret_cast<const int4 *>(k);
    int a0c = 1 << (kPrecisionBits - 1), a1c = a0c, a2c = a0c;
    uint32_t w0 = wp[0];
    for (int g = 0; 4 * g < n; ++g) {
        const uint32_t w1 = wp[3 * g + 1], w2 = wp[3 * g + 2], w3 = wp[3 * g + 3];
        const uint32_t r0 = __funnelshift_r(w0, w1, sh), r1 = __funnelshift_r(w1, w2, sh), r2 = __funnelshift_r(w2, w3, sh);
        w0 = w3;
        const int4 c = __ldg(k4 + g);  // taps past n are zero, so whatever bytes follow the span do not count
        a0c += (int)(r0 & 0xffu) * c.x + (int)(r0 >> 24) * c.y + (int)((r1 >> 16) & 0xffu) * c.z + (int)((r2 >> 8) & 0xffu) * c.w;
        a1c += (int)((r0 >> 8) & 0xffu) * c.x + (int)(r1 & 0xffu) * c.y + (int)(r1 >> 24) * c.z + (int)((r2 >> 16) & 0xffu) * c.w;
        a2c += (int)((r0 >> 16) & 0xffu) * c.x + (int)((r1 >> 8) & 0xffu) * c.y + (int)(r2 & 0xffu) * c.z + (int)(r2 >> 24) * c.w;
    }
    uint8_t *o = out + ((size_t)y * OW + xx) * 3;
    o[0] = (uint8_t)clip8(a0c);
    o[1] = (uint8_t)clip8(a1c);
    o[2] = (uint8_t)clip8(a2c);
}

struct Norm3 {
    float mean[3], std[3];
};

// Vertical pass fused with ToTensor + Normalize: (H, OW, 3) uint8 -> (3, OH, OW) float32.
// Threads of a warp take consecutive columns: the 3-byte pixels they read and the floats they
// write per channel are contiguous.
__global__ void __launch_bounds__(256)
pp_vertical_normalize(const uint8_t *__restrict__ in, int H, int OW, int OH, int ksize,
                      const int *__restrict__ bounds, const int *__restrict__ kk, const __grid_constant__ Norm3 nm,
                      float *__restrict__ out) {
    in += (size_t)blockIdx.z * H * OW * 3;
    out += (size_t)blockIdx.z * 3 * OH * OW;
    const int xx = blockIdx.x * 256 + threadIdx.x, yy = blockIdx.y;
    if (xx >= OW) return;
    const int y0 = bounds[2 * yy], n = bounds[2 * yy + 1];
    const int *k = kk + (size_t)yy * ksize;
    int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
    for (int y = 0; y < n; ++y) {
        const int c = __ldg(k + y);
        const uint8_t *p = in + ((size_t)(y0 + y) * OW + xx) * 3;
        a0 += (int)__ldg(p) * c;
        a1 += (int)__ldg(p + 1) * c;
        a2 += (int)__ldg(p + 2) * c;
    }
    const int acc[3] = {a0, a1, a2};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float v = __fdiv_rn((float)clip8(acc[c]), 255.0f);                     // to_tensor: .div(255)
        out[((size_t)c * OH + yy) * OW + xx] = __fdiv_rn(__fsub_rn(v, nm.mean[c]), nm.std[c]);  // .sub_(mean).div_(std)
    }
}

struct PpLayout {
    size_t bounds_h, kk_h, bounds_v, kk_v, tmp, total;
    AxisPlan ph, pv;
};
static PpLayout pp_layout(int N, int H, int W, int OH, int OW) {
    PpLayout L;
    L.ph = axis_plan(W, OW);
    L.pv = axis_plan(H, OH);
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t o = 0;
    L.bounds_h = o; o = up(o + sizeof(int) * 2 * (size_t)OW);
    L.kk_h = o; o = up(o + sizeof(int) * (size_t)OW * L.ph.ksize);
    L.bounds_v = o; o = up(o + sizeof(int) * 2 * (size_t)OH);
    L.kk_v = o; o = up(o + sizeof(int) * (size_t)OH * L.pv.ksize);
    L.tmp = o; o = up(o + (size_t)N * H * OW * 3);
    L.total = o;
    return L;
}

}  // namespace yh

using namespace yh;

static bool pp_shape_ok(int N, int H, int W, int OH, int OW) {
    return N >= 1 && N <= 65535 && H >= 1 && W >= 1 && OH >= 1 && OW >= 1 && H <= 65535 && OH <= 65535 && W <= (1 << 20) && OW <= (1 << 20);
}

extern "C" size_t yh_preprocess_workspace_bytes(int N, int H, int W, int out_h, int out_w) {
    if (!pp_shape_ok(N, H, W, out_h, out_w)) return 0;
    return pp_layout(N, H, W, out_h, out_w).total;
}

extern "C" int yh_preprocess_rgb8(const uint8_t *rgb, int N, int H, int W, int out_h, int out_w, const float *mean3,
                                  const float *std3, float *out, void *workspace, size_t workspace_bytes,
                                  void *stream) {
    if (!rgb || !out || !workspace || !mean3 || !std3 || !pp_shape_ok(N, H, W, out_h, out_w)) {
        set_error("yh_preprocess_rgb8: bad argument (N=%d H=%d W=%d -> %dx%d)", N, H, W, out_h, out_w);
        return YH_ERR_BAD_ARG;
    }
    const PpLayout L = pp_layout(N, H, W, out_h, out_w);
    if (workspace_bytes < L.total) {
        set_error("yh_preprocess_rgb8: workspace too small (%zu < %zu)", workspace_bytes, L.total);
        return YH_ERR_WORKSPACE;
    }
    if ((uintptr_t)workspace % 256 || (uintptr_t)rgb % 4) {
        set_error("yh_preprocess_rgb8: workspace must be 256-byte aligned, rgb 4-byte aligned");
        return YH_ERR_ALIGNMENT;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char *ws = (char *)workspace;
    int *bounds_h = (int *)(ws + L.bounds_h), *kk_h = (int *)(ws + L.kk_h);
    int *bounds_v = (int *)(ws + L.bounds_v), *kk_v = (int *)(ws + L.kk_v);
    uint8_t *tmp = (uint8_t *)(ws + L.tmp);
    int launched = 0;
    const uint8_t *vin = rgb;
    // Pillow skips the horizontal pass when the width already matches.  The vertical pass with equal heights is
    // the identity (taps 1, 0), so it always runs: it carries the normalisation.
    const bool horiz = W != out_w;
    const AxisArgs axh = {W, horiz ? out_w : 0, L.ph.ksize, bounds_h, kk_h};
    const AxisArgs axv = {H, out_h, L.pv.ksize, bounds_v, kk_v};
    const int widest = (horiz && out_w > out_h) ? out_w : out_h;
    pp_coeffs<<<dim3((widest + 127) / 128, 2), 128, 0, st>>>(axh, axv);
    ++launched;
    if (horiz) {
        // input pixels a segment of kHSeg outputs can tap: its centres span kHSeg * scale, plus the support on both sides
        const double scale = (double)W / (double)out_w;
        const size_t span_px = (size_t)ceil(kHSeg * scale) + (size_t)L.ph.ksize + 2;
        const size_t span_bytes = ((span_px * 3 + 8 + 3) & ~(size_t)3) + 32;  // + the words a last group reads past the span
        if (span_bytes <= 48 * 1024)
            pp_horizontal_staged<<<dim3((out_w + kHSeg - 1) / kHSeg, H, N), kHSeg, span_bytes, st>>>(
                rgb, H, W, out_w, L.ph.ksize, bounds_h, kk_h, tmp);
        else
            pp_horizontal<<<dim3((out_w + 255) / 256, H, N), 256, 0, st>>>(rgb, H, W, out_w, L.ph.ksize, bounds_h, kk_h, tmp);
        vin = tmp;
        ++launched;
    }
    Norm3 nm;
    for (int c = 0; c < 3; ++c) {
        nm.mean[c] = mean3[c];
        nm.std[c] = std3[c];
    }
    pp_vertical_normalize<<<dim3((out_w + 255) / 256, out_h, N), 256, 0, st>>>(vin, H, out_w, out_h, L.pv.ksize, bounds_v,
                                                                           kk_v, nm, out);
    ++launched;
    return check_launch("yh_preprocess_rgb8", launched);
}
