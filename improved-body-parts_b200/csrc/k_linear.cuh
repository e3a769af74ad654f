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
    int nc[kLinMaxScales];    // feature columns a CTA's 128 output columns can touch (capacity of a staged row)
    int maxr[kLinMaxScales];  // feature rows a segment can add (capacity of a staging buffer)
    int off[kLinMaxScales];   // float offset of the scale's rows inside a staging buffer: upright rows, then twin rows
    int buf_floats;           // floats per staging buffer (there are two)
    int n_scales, flip, H, W;
    int chain_rows;           // output rows one CTA walks (multiple of kLinSeg)
    double *out;
    float *heat32;
};
constexpr int kLinSeg = 16;  // output rows per staging step

__device__ __forceinline__ void lin_cp4(void *dst_smem, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void lin_cp16(void *dst_smem, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}

// A CTA owns 128 output columns of one channel and walks down `chain_rows` rows, kLinSeg at a time.  The feature rows a
// segment adds to the windows (and the row taps of its output rows) are copied to shared memory with cp.async one
// segment AHEAD, all requests independent, so HBM latency is paid once per segment and under the previous segment's
// arithmetic; the walk itself only reads shared memory.
template <int NS>
__global__ void __launch_bounds__(kLinThreads) fuse_linear_kernel(const __grid_constant__ LinArgs a) {
    extern __shared__ __align__(16) float lin_smem[];
    const int tid = threadIdx.x;
    const int X0 = blockIdx.x * kLinThreads, X = X0 + tid;
    const int Xl = min(X, a.W - 1);
    const int chan = blockIdx.z % kLayers, img = blockIdx.z / kLayers;
    const int Y0 = blockIdx.y * a.chain_rows, Y1 = min(Y0 + a.chain_rows, a.H);
    const bool flip = a.flip != 0;
    LinTap *YT = reinterpret_cast<LinTap *>(lin_smem + 2 * a.buf_floats);  // [2][NS][kLinSeg]
    __shared__ int s_flo[2][kLinMaxScales];                                  // first feature row held by each buffer

    float c0[NS], c1[NS], c2[NS];  // the lane's column taps
    int col[NS];                   // its first feature column, relative to the CTA's first
    int c_lo[NS];                  // the CTA's first feature column
    const float *upr[NS], *twin[NS];
    float w0[NS], w1[NS], w2[NS];  // horizontal sums of feature rows top, top + 1, top + 2
    int top[NS], loaded[NS];       // window position; last feature row requested so far
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const LinScale &S = a.sc[s];
        const LinTap t = S.xtab[Xl];
        c_lo[s] = __ldg(&S.xtab[X0].base);
        col[s] = t.base - c_lo[s];
        c0[s] = t.c[0], c1[s] = t.c[1], c2[s] = t.c[2];
        const size_t hw = (size_t)S.h * S.w;
        const float *base = S.src + (size_t)img * 2 * kLayers * hw;
        upr[s] = base + (size_t)chan * hw;
        twin[s] = base + (size_t)(kLayers + c_flip_src[chan]) * hw;
        top[s] = __ldg(&S.ytab[Y0].base) - 3;  // the first output row pulls three rows into the window
        loaded[s] = top[s] + 2;
        w0[s] = w1[s] = w2[s] = 0.f;
    }

    // requests of the segment that starts at row Ya into buffer b
    auto request = [&](int Ya, int b) {
        const int Yb = min(Ya + kLinSeg, Y1);
        float *buf = lin_smem + b * a.buf_floats;
#pragma unroll
        for (int s = 0; s < NS; s++) {
            const LinScale &S = a.sc[s];
            const int f_hi = __ldg(&S.ytab[Yb - 1].base) + 2, f_lo = loaded[s] + 1;
            if (tid == 0) s_flo[b][s] = f_lo;
            for (int t = tid; t < a.nc[s]; t += kLinThreads) {
                const int c = min(c_lo[s] + t, S.w - 1);
                float *du = buf + a.off[s] + t, *dt = du + a.maxr[s] * a.nc[s];
                for (int fr = f_lo; fr <= f_hi; fr++) {
                    lin_cp4(du, upr[s] + (size_t)fr * S.w + c);
                    if (flip) lin_cp4(dt, twin[s] + (size_t)fr * S.w + (S.w - 1 - c));
                    du += a.nc[s], dt += a.nc[s];
                }
            }
            loaded[s] = max(loaded[s], f_hi);
            if (tid < Yb - Ya) lin_cp16(&YT[(b * NS + s) * kLinSeg + tid], &S.ytab[Ya + tid]);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    double *out = a.out + ((size_t)img * kLayers + chan) * (size_t)a.H * a.W;
    float *h32 = (a.heat32 && chan >= kLimbs && chan < kLimbs + kParts) ? a.heat32 + ((size_t)img * kParts + (chan - kLimbs)) * (size_t)a.H * a.W : nullptr;
    request(Y0, 0);
    int b = 0;
    for (int Ya = Y0; Ya < Y1; Ya += kLinSeg, b ^= 1) {
        if (Ya + kLinSeg < Y1) {
            request(Ya + kLinSeg, b ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const float *buf = lin_smem + b * a.buf_floats;
        const int Yb = min(Ya + kLinSeg, Y1);
        for (int Y = Ya; Y < Yb; Y++) {
            double acc = 0.0;
#pragma unroll
            for (int s = 0; s < NS; s++) {
                const float4 e = *reinterpret_cast<const float4 *>(&YT[(b * NS + s) * kLinSeg + (Y - Ya)]);  // base, c0, c1, c2 of the row
                const int ebase = __float_as_int(e.x);
                while (top[s] < ebase) {  // the window moves down one feature row
                    top[s]++;
                    const float *pu = buf + a.off[s] + (top[s] + 2 - s_flo[b][s]) * a.nc[s] + col[s];
                    float f0 = pu[0], f1 = pu[1], f2 = pu[2];
                    if (flip) {
                        const float *pt = pu + a.maxr[s] * a.nc[s];
                        f0 = avg_like_numpy(f0, pt[0]), f1 = avg_like_numpy(f1, pt[1]), f2 = avg_like_numpy(f2, pt[2]);
                    }
                    w0[s] = w1[s];
                    w1[s] = w2[s];
                    w2[s] = __fmaf_rn(c2[s], f2, __fmaf_rn(c1[s], f1, __fmul_rn(c0[s], f0)));
                }
                const float v = __fmaf_rn(e.w, w2[s], __fmaf_rn(e.z, w1[s], __fmul_rn(e.y, w0[s])));
                acc = __dadd_rn(acc, (double)v);
            }
            if (X < a.W) {
                const size_t o = (size_t)Y * a.W + X;
                out[o] = acc;
                if (h32) h32[o] = __double2float_rn(acc);
            }
        }
        __syncthreads();  // buffer b is free for the segment after next
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
// K = kernel size as a compile-time constant (0: run-time size, any odd K <= kGsMaxK)
template <int K>
__global__ void __launch_bounds__(256) gauss_smooth_kernel(const __grid_constant__ GaussArgs a) {
    extern __shared__ __align__(16) float gs_smem[];
    const int ksize = K ? K : a.ksize;
    const int p = ksize >> 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int X0 = blockIdx.x * kGsTileX, Y0 = blockIdx.y * kGsTileY;
    const float *src = a.src + (size_t)blockIdx.z * a.H * a.W;
    float *dst = a.dst + (size_t)blockIdx.z * a.H * a.W;
    const int in_w = kGsTileX + 2 * p, in_h = kGsTileY + 2 * p;
    float *tin = gs_smem;                 // [in_h][in_w]
    float *tmid = gs_smem + in_h * in_w;  // [in_h][kGsTileX]
    float g[K ? K : kGsMaxK];
#pragma unroll
    for (int k = 0; k < (K ? K : kGsMaxK); k++) g[k] = k < ksize ? a.g[k] : 0.f;
    // tile + mirrored halo: a warp per row, lanes along the row
    for (int r = warp; r < in_h; r += 8) {
        const int y = reflect_idx(min(Y0 + r - p, a.H - 1 + p), a.H);
        const float *row = src + (size_t)y * a.W;
        for (int c = lane; c < in_w; c += 32) tin[r * in_w + c] = __ldg(row + reflect_idx(min(X0 + c - p, a.W - 1 + p), a.W));
    }
    __syncthreads();
    // horizontal pass: a warp per row, a lane four columns 32 apart
    for (int r = warp; r < in_h; r += 8) {
        const float *q = tin + r * in_w + lane;
        float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll
        for (int k = 0; k < (K ? K : kGsMaxK); k++) {
            if (!K && k >= ksize) break;
            v0 = __fmaf_rn(g[k], q[k], v0);
            v1 = __fmaf_rn(g[k], q[k + 32], v1);
            v2 = __fmaf_rn(g[k], q[k + 64], v2);
            v3 = __fmaf_rn(g[k], q[k + 96], v3);
        }
        float *o = tmid + r * kGsTileX + lane;
        o[0] = v0, o[32] = v1, o[64] = v2, o[96] = v3;
    }
    __syncthreads();
    // vertical pass: a warp per output row
    for (int r = warp; r < kGsTileY; r += 8) {
        if (Y0 + r >= a.H) break;
        const float *q = tmid + r * kGsTileX + lane;
        float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll
        for (int k = 0; k < (K ? K : kGsMaxK); k++) {
            if (!K && k >= ksize) break;
            v0 = __fmaf_rn(g[k], q[k * kGsTileX], v0);
            v1 = __fmaf_rn(g[k], q[k * kGsTileX + 32], v1);
            v2 = __fmaf_rn(g[k], q[k * kGsTileX + 64], v2);
            v3 = __fmaf_rn(g[k], q[k * kGsTileX + 96], v3);
        }
        float *o = dst + (size_t)(Y0 + r) * a.W + X0 + lane;
        if (X0 + lane < a.W) o[0] = v0;
        if (X0 + lane + 32 < a.W) o[32] = v1;
        if (X0 + lane + 64 < a.W) o[64] = v2;
        if (X0 + lane + 96 < a.W) o[96] = v3;
    }
}

}  // namespace ibp
