// Second operator mode ("north-star operators"): the multi-scale + flip fusion of
// utils/parse_skeletons.py:180-283 with BILINEAR interpolation in both resizes (cv2.INTER_LINEAR instead of the
// reference's INTER_CUBIC), and the separable Gaussian smoothing of utils/util.py:103-174 (GaussianSmoothing, dead code
// in the reference).  These are NOT what the reference computes on its live path; they are the operators BASELINE.json's
// north_star names, offered beside the reference's own, each with its own CPU oracle (in the test tree:
// predict_fuse_linear, gaussian_smooth) and a tolerance instead of bit equality.
//
// Bilinear fusion.  Two bilinear resizes in a row are one linear operator with at most three taps per axis on the
// FEATURE map (the two taps of the second resize are neighbours in the x4 image, whose own taps overlap), so nothing
// intermediate is needed:
//     out[Y][X] = sum_s sum_i cy_s,i(Y) * ( sum_j cx_s,j(X) * F_s[fy_s(Y) + i][fx_s(X) + j] ),      i, j in 0..2
// with F_s the flip average of scale s.  The tap positions and weights follow cv2's rules (source coordinate
// (d + 0.5) * scale - 0.5, left/right clamp with the fraction forced to 0) and are composed on the host; 1 / n_scales
// is folded into cy.  A lane owns an output column and walks down the rows, keeping for every scale the horizontal
// sums of three consecutive feature rows in registers; a feature row enters the window every 4 / scale output rows.
// The kernel reads every network output once (through L1) and writes the float64 result: it is bound by HBM.
#pragma once
#include "ibp_device.cuh"

namespace ibp {

constexpr int kLinMaxScales = 8;
constexpr int kLinThreads = 128;

struct __align__(16) LinTap {  // one axis, one destination index, one scale
    int base;                  // first of three consecutive feature indices (base + 2 is inside the map)
    float c[3];
};
struct LinScale {
    const float *src;    // (batch, 2, 50, h, w)
    const LinTap *xtab;  // [W]
    const LinTap *ytab;  // [H]
    int h, w;
};
struct LinArgs {
    LinScale sc[kLinMaxScales];
    int n_scales, flip, H, W;
    int chain_rows;
    double *out;
    float *heat32;
};

// horizontal sums of feature row `fr` for the lane's column: sum_j cx_j * avg(F[fr][b + j], twin[fr][w - 1 - b - j])
__device__ __forceinline__ float lin_row(const float *upr, const float *twin, int w, int fr, const LinTap &t, bool flip) {
    const float *p = upr + (size_t)fr * w + t.base;
    float f0 = __ldg(p), f1 = __ldg(p + 1), f2 = __ldg(p + 2);
    if (flip) {
        const float *q = twin + (size_t)fr * w + (w - 1 - t.base);
        f0 = avg_like_numpy(f0, __ldg(q));
        f1 = avg_like_numpy(f1, __ldg(q - 1));
        f2 = avg_like_numpy(f2, __ldg(q - 2));
    }
    return __fmaf_rn(t.c[2], f2, __fmaf_rn(t.c[1], f1, __fmul_rn(t.c[0], f0)));
}

template <int NS>
__global__ void __launch_bounds__(kLinThreads) fuse_linear_kernel(LinArgs a) {
    const int X = blockIdx.x * kLinThreads + threadIdx.x;
    const int Xl = min(X, a.W - 1);
    const int chan = blockIdx.z % kLayers, img = blockIdx.z / kLayers;
    const int Y0 = blockIdx.y * a.chain_rows, Y1 = min(Y0 + a.chain_rows, a.H);
    const bool flip = a.flip != 0;

    LinTap xt[NS];
    const float *upr[NS], *twin[NS];
    float w0[NS], w1[NS], w2[NS];  // horizontal sums of feature rows top, top + 1, top + 2
    int top[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const LinScale &S = a.sc[s];
        xt[s] = S.xtab[Xl];
        const size_t hw = (size_t)S.h * S.w;
        const float *base = S.src + (size_t)img * 2 * kLayers * hw;
        upr[s] = base + (size_t)chan * hw;
        twin[s] = base + (size_t)(kLayers + c_flip_src[chan]) * hw;
        top[s] = __ldg(&S.ytab[Y0].base);
        w0[s] = lin_row(upr[s], twin[s], S.w, top[s], xt[s], flip);
        w1[s] = lin_row(upr[s], twin[s], S.w, top[s] + 1, xt[s], flip);
        w2[s] = lin_row(upr[s], twin[s], S.w, top[s] + 2, xt[s], flip);
    }
    double *out = a.out + ((size_t)img * kLayers + chan) * (size_t)a.H * a.W;
    float *h32 = (a.heat32 && chan >= kLimbs && chan < kLimbs + kParts) ? a.heat32 + ((size_t)img * kParts + (chan - kLimbs)) * (size_t)a.H * a.W : nullptr;
    for (int Y = Y0; Y < Y1; Y++) {
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < NS; s++) {
            const LinScale &S = a.sc[s];
            const int4 e = __ldg(reinterpret_cast<const int4 *>(&S.ytab[Y]));  // warp-uniform: base, c0, c1, c2
            while (top[s] < e.x) {  // the window moves down one feature row (at most once per output row for scales <= 4)
                top[s]++;
                w0[s] = w1[s];
                w1[s] = w2[s];
                w2[s] = lin_row(upr[s], twin[s], S.w, top[s] + 2, xt[s], flip);
            }
            const float v = __fmaf_rn(__int_as_float(e.w), w2[s], __fmaf_rn(__int_as_float(e.z), w1[s], __fmul_rn(__int_as_float(e.y), w0[s])));
            acc = __dadd_rn(acc, (double)v);
        }
        if (X < a.W) {
            const size_t o = (size_t)Y * a.W + X;
            out[o] = acc;
            if (h32) h32[o] = __double2float_rn(acc);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Separable Gaussian smoothing with a mirrored border: GaussianSmoothing(channels, ksize, sigma) applied to
// F.pad(x, (p, p, p, p), mode='reflect') — utils/util.py:103-174 and its docstring.  One CTA = a 128 x 32 tile of one
// plane: tile + halo to shared memory, horizontal pass into a second buffer, vertical pass to HBM.  Reads and writes
// every pixel once: bound by HBM.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kGsTileX = 128, kGsTileY = 32, kGsMaxK = 15;
struct GaussArgs {
    const float *src;
    float *dst;
    int H, W, ksize;
    float g[kGsMaxK];  // normalised 1-D taps
};
__device__ __forceinline__ int reflect_idx(int i, int n) {  // torch 'reflect': -1 -> 1, n -> n - 2
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    return i;
}
__global__ void __launch_bounds__(256) gauss_smooth_kernel(const __grid_constant__ GaussArgs a) {
    extern __shared__ __align__(16) float gs_smem[];
    const int p = a.ksize >> 1;
    const int X0 = blockIdx.x * kGsTileX, Y0 = blockIdx.y * kGsTileY;
    const float *src = a.src + (size_t)blockIdx.z * a.H * a.W;
    float *dst = a.dst + (size_t)blockIdx.z * a.H * a.W;
    const int in_w = kGsTileX + 2 * p, in_h = kGsTileY + 2 * p;
    float *tin = gs_smem;               // [in_h][in_w]
    float *tmid = gs_smem + in_h * in_w;  // [in_h][kGsTileX]
    for (int i = threadIdx.x; i < in_h * in_w; i += blockDim.x) {
        const int r = i / in_w, c = i - r * in_w;
        const int y = reflect_idx(min(Y0 + r - p, a.H - 1 + p), a.H), x = reflect_idx(min(X0 + c - p, a.W - 1 + p), a.W);
        tin[i] = __ldg(src + (size_t)y * a.W + x);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < in_h * kGsTileX; i += blockDim.x) {
        const int r = i / kGsTileX, c = i - r * kGsTileX;
        const float *q = tin + r * in_w + c;
        float v = 0.f;
        for (int k = 0; k < a.ksize; k++) v = __fmaf_rn(a.g[k], q[k], v);
        tmid[i] = v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kGsTileY * kGsTileX; i += blockDim.x) {
        const int r = i / kGsTileX, c = i - r * kGsTileX;
        if (Y0 + r < a.H && X0 + c < a.W) {
            const float *q = tmid + r * kGsTileX + c;
            float v = 0.f;
            for (int k = 0; k < a.ksize; k++) v = __fmaf_rn(a.g[k], q[k * kGsTileX], v);
            dst[(size_t)(Y0 + r) * a.W + X0 + c] = v;
        }
    }
}

}  // namespace ibp
