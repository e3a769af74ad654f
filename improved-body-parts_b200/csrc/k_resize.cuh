// Image resize of the scale search: cv2.resize(image, (0, 0), fx=scale, fy=scale, INTER_CUBIC) on a
// uint8 HWC image — utils/parse_skeletons.py:204 (predict) and :52 (predict_refactor).
//
// OpenCV's own 8-bit cubic path (resize.cpp, resizeGeneric_<HResizeCubic<uchar,int,short>, VResizeCubic<uchar,int,short,
// FixedPtCast<int,uchar,22>, VResizeCubicVec_32s8u>>), restated and held bit-exact against cv2 (IPP off) by the tests:
//   * destination size: dsize = (rint(w * fx), rint(h * fy)) (saturate_cast<int>(double) rounds half to even);
//     with dsize = (0, 0) the source coordinate uses scale = 1. / fx — NOT src / dst;
//   * coordinates fx = (float)((d + 0.5) * scale - 0.5) (double scale), floor, fraction; cubic weights with A = -0.75 in
//     float (the same interpolateCubic as the float path), then FIXED POINT: short(rint(w * 2048)) per tap;
//   * horizontal pass in int: S[x0]*a0 + S[x1]*a1 + S[x2]*a2 + S[x3]*a3, tap indices clamped to the image;
//   * vertical pass: the SIMD body converts the four int rows to float and computes
//         s0*b0 + (s1*b1 + (s2*b2 + s3*b3))      b_k = (float)beta_k * 2^-22, separate multiplies and adds (the 128-bit
//     universal-intrinsics path has no FMA), rint, saturate to uint8; the scalar tail of a row — its last
//     (dw * 3) % lanes elements — stays in fixed point: (s0*β0 + s1*β1 + s2*β2 + s3*β3 + 2^21) >> 22, saturated.
//     The two agree except on exact .5 ties; `lanes` = 8 (v_int16 lanes of the 128-bit path that matched cv2 here).
// One thread per destination pixel (three channels): 48 byte loads that hit L1/L2 — the whole image is a few MB and the
// kernel runs once per scale in front of a 60 ms network forward; it is written for exactness, not for bandwidth.
#pragma once
#include "ibp_device.cuh"
#include "k_multiscale.cuh"

namespace ibp {

struct ResizeArgs {
    const uint8_t *img;  // (batch, h, w, 3) uint8 HWC
    uint8_t *out;        // (batch, dh, dw, 3) uint8 HWC
    int batch, h, w, dh, dw;
    double scale_x, scale_y;  // 1. / fx, 1. / fy
    int lanes;                // elements per step of OpenCV's SIMD body (the tail is what remains of dw * 3)
};

__device__ __forceinline__ int resize_fixed_coeff(float w) {
    // saturate_cast<short>(w * INTER_RESIZE_COEF_SCALE): float multiply, cvRound (half to even), saturate
    return clampi(__float2int_rn(__fmul_rn(w, 2048.0f)), -32768, 32767);
}

__global__ void __launch_bounds__(128) resize_u8_cubic_kernel(ResizeArgs a) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= a.dw) return;
    float wf[4];
    int xi[4], yi[4], ax[4], by[4];
    const int sx = ms_tap(x, a.scale_x, wf);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        ax[k] = resize_fixed_coeff(wf[k]);
        xi[k] = clampi(sx - 1 + k, 0, a.w - 1) * 3;
    }
    const int sy = ms_tap(y, a.scale_y, wf);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        by[k] = resize_fixed_coeff(wf[k]);
        yi[k] = clampi(sy - 1 + k, 0, a.h - 1);
    }
    const uint8_t *src = a.img + (size_t)b * a.h * a.w * 3;
    int H[4][3];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const uint8_t *row = src + (size_t)yi[r] * a.w * 3;
#pragma unroll
        for (int c = 0; c < 3; c++)
            H[r][c] = (int)row[xi[0] + c] * ax[0] + (int)row[xi[1] + c] * ax[1] + (int)row[xi[2] + c] * ax[2] + (int)row[xi[3] + c] * ax[3];
    }
    const float scale = 1.0f / (2048.0f * 2048.0f);
    float bf[4];
#pragma unroll
    for (int k = 0; k < 4; k++) bf[k] = __fmul_rn(__int2float_rn(by[k]), scale);
    const int n = a.dw * 3, n_vec = (n / a.lanes) * a.lanes;
    uint8_t *dst = a.out + ((size_t)(b * a.dh + y) * a.dw + x) * 3;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        int v;
        if (x * 3 + c < n_vec) {
            const float s0 = __int2float_rn(H[0][c]), s1 = __int2float_rn(H[1][c]), s2 = __int2float_rn(H[2][c]), s3 = __int2float_rn(H[3][c]);
            const float f = __fadd_rn(__fmul_rn(s0, bf[0]), __fadd_rn(__fmul_rn(s1, bf[1]), __fadd_rn(__fmul_rn(s2, bf[2]), __fmul_rn(s3, bf[3]))));
            v = __float2int_rn(f);
        } else {
            v = (H[0][c] * by[0] + H[1][c] * by[1] + H[2][c] * by[2] + H[3][c] * by[3] + (1 << 21)) >> 22;
        }
        dst[c] = (uint8_t)clampi(v, 0, 255);
    }
}

}  // namespace ibp
