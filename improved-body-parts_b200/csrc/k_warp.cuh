// The rotation_search branch of predict / predict_refactor: utils/parse_skeletons.py:215-218
// (forward cv2.warpAffine of the padded float32 input image), :265-267 (backward warpAffine of the x4 up-sampled 20- and
// 30-channel maps), :98-100 (predict_refactor: backward warp of the fused feature-resolution maps); knob utils/config:15.
//
// warp_affine_kernel restates cv2.warpAffine(src, M, (0, 0)) for float32 data with the defaults the reference uses
// (INTER_LINEAR, BORDER_CONSTANT 0, dsize = source size) — OpenCV's fixed-point remap path (imgwarp.cpp hal::warpAffine +
// remapBilinear), bit-exact against cv2 4.13 for 1, 3, 4, 20 and 30 channels (tests/test_rotation.py):
//   * the host inverts M in double exactly as cv::warpAffine does and hands the inverse in;
//   * per destination pixel: adelta = rint(M0 * x * 1024), bdelta = rint(M3 * x * 1024),
//     X0 = rint((M1 * y + M2) * 1024) + 16, Y0 likewise (round_delta = AB_SCALE / INTER_TAB_SIZE / 2);
//     X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5: source pixel (X >> 5, Y >> 5), fractions (X & 31, Y & 31) in 1/32;
//   * weights from the 32 x 32 table: wy * wx with w0 = 1.f - i * (1.f / 32), w1 = i * (1.f / 32) (float);
//   * value = ((v00 * w00 + v01 * w01) + v10 * w10) + v11 * w11, separate multiplies and adds, pixels outside the image = 0.
// The weights are the same for every channel: one thread per destination pixel, channels in a loop.  Generic element strides
// serve both layouts on the path: planar (B, C, h, w) maps and the HWC (B, h, w, 3) input image.
//
// crop_resize_accumulate_kernel is predict's tail for ONE (scale, rotation) item when the fused multi-scale kernel cannot be
// used (a warp sits between the two resizes): crop the x4 maps to the scaled image (l.272-273), cv2.resize to the image size
// (INTER_CUBIC, l.276-277), divide by the number of items in float32 and add to the float64 averages (l.280-281).
#pragma once
#include "ibp_device.cuh"
#include "k_multiscale.cuh"

namespace ibp {

struct WarpArgs {
    const float *src;
    float *dst;
    int batch, channels, h, w;
    long long bs, cs, ys, xs;  // element strides of batch, channel, row, pixel (same for src and dst)
    double m[6];               // INVERSE map (destination -> source), inverted on the host like cv::warpAffine does
};

__global__ void __launch_bounds__(128) warp_affine_kernel(WarpArgs a) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= a.w) return;
    const double dx = (double)x, dy = (double)y;
    const int adelta = __double2int_rn(__dmul_rn(__dmul_rn(a.m[0], dx), 1024.0));
    const int bdelta = __double2int_rn(__dmul_rn(__dmul_rn(a.m[3], dx), 1024.0));
    const int X0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m[1], dy), a.m[2]), 1024.0)) + 16;
    const int Y0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m[4], dy), a.m[5]), 1024.0)) + 16;
    const int X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
    const int sx = clampi(X >> 5, -32768, 32767), sy = clampi(Y >> 5, -32768, 32767);  // saturate_cast<short>
    const float tx = __fmul_rn(__int2float_rn(X & 31), 1.0f / 32.0f), ty = __fmul_rn(__int2float_rn(Y & 31), 1.0f / 32.0f);
    const float wx0 = __fsub_rn(1.0f, tx), wx1 = tx, wy0 = __fsub_rn(1.0f, ty), wy1 = ty;
    const float w00 = __fmul_rn(wy0, wx0), w01 = __fmul_rn(wy0, wx1), w10 = __fmul_rn(wy1, wx0), w11 = __fmul_rn(wy1, wx1);
    const bool in_x0 = sx >= 0 && sx < a.w, in_x1 = sx + 1 >= 0 && sx + 1 < a.w;
    const bool in_y0 = sy >= 0 && sy < a.h, in_y1 = sy + 1 >= 0 && sy + 1 < a.h;
    const float *s = a.src + (long long)b * a.bs;
    float *d = a.dst + (long long)b * a.bs + (long long)y * a.ys + (long long)x * a.xs;
    const long long o00 = (long long)sy * a.ys + (long long)sx * a.xs, o01 = o00 + a.xs, o10 = o00 + a.ys, o11 = o10 + a.xs;
    for (int c = 0; c < a.channels; c++) {
        const float *p = s + (long long)c * a.cs;
        const float v00 = (in_y0 && in_x0) ? __ldg(p + o00) : 0.0f, v01 = (in_y0 && in_x1) ? __ldg(p + o01) : 0.0f;
        const float v10 = (in_y1 && in_x0) ? __ldg(p + o10) : 0.0f, v11 = (in_y1 && in_x1) ? __ldg(p + o11) : 0.0f;
        d[(long long)c * a.cs] =
            __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(v00, w00), __fmul_rn(v01, w01)), __fmul_rn(v10, w10)), __fmul_rn(v11, w11));
    }
}

struct CropResizeArgs {
    const float *src;  // (batch, 50, ph, pw) planar x4 maps at the padded input size
    int batch, ph, pw;
    int sh, sw;        // rows / columns kept by the crop (the scaled image before padding)
    int H, W;          // image size
    double scale_x, scale_y;  // 1. / ((double)W / sw), 1. / ((double)H / sh): what cv2 derives from an explicit dsize
    float divisor;     // len(multiplier) * len(rotate_angle)
    double *out;       // (batch, 50, H, W) planar float64 averages, accumulated into
};

__global__ void __launch_bounds__(128) crop_resize_accumulate_kernel(CropResizeArgs a) {
    const int X = blockIdx.x * blockDim.x + threadIdx.x, Y = blockIdx.y, b = blockIdx.z;
    if (X >= a.W) return;
    float wx[4], wy[4];
    const int sx = ms_tap(X, a.scale_x, wx), sy = ms_tap(Y, a.scale_y, wy);
    int xi[4], yo[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        xi[k] = clampi(sx - 1 + k, 0, a.sw - 1);            // replicate border at the edge of the CROPPED map
        yo[k] = clampi(sy - 1 + k, 0, a.sh - 1) * a.pw;
    }
    const size_t plane = (size_t)a.ph * a.pw, oplane = (size_t)a.H * a.W;
    const float *src = a.src + (size_t)b * kLayers * plane;
    double *out = a.out + (size_t)b * kLayers * oplane + (size_t)Y * a.W + X;
    for (int c = 0; c < kLayers; c++) {
        // the reference resizes the 30 limb channels and the 20 keypoint channels as two interleaved images: the last
        // (W * cn) % 4 elements of a destination row take the scalar tail of OpenCV's vertical pass (left-associated sum)
        const int cn = (c < kLimbs) ? kLimbs : (kLayers - kLimbs), cl = (c < kLimbs) ? c : c - kLimbs;
        const bool vec = (X * cn + cl) < ((a.W * cn) & ~3);
        const float *p = src + (size_t)c * plane;
        float r[4];
#pragma unroll
        for (int k = 0; k < 4; k++) r[k] = hsum4(__ldg(p + yo[k] + xi[0]), __ldg(p + yo[k] + xi[1]), __ldg(p + yo[k] + xi[2]), __ldg(p + yo[k] + xi[3]), wx);
        const float v = vec ? vsum4(r[0], r[1], r[2], r[3], wy) : vsum4_tail(r[0], r[1], r[2], r[3], wy);
        out[(size_t)c * oplane] = __dadd_rn(out[(size_t)c * oplane], (double)__fdiv_rn(v, a.divisor));
    }
}

}  // namespace ibp
