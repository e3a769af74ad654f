// K1 (multi-scale) — flip / scale-search fusion at image resolution.
// Reference: utils/parse_skeletons.py:180-283 (predict).  Per scale s of the search (l.186, l.194):
//     fused_s   = (out_s[0] + mirror(out_s[1])[perm]) / 2                              l.232-248, feature resolution
//     up_s      = cv2.resize(fused_s, fx=stride, fy=stride, INTER_CUBIC)               l.252-263, padded input size
//     crop_s    = up_s[:sh, :sw]                                                        l.272-273, pad removed
//     img_s     = cv2.resize(crop_s, (img_w, img_h), INTER_CUBIC)                       l.276-277, image size
//     avg      += img_s / n_scales      (float32 division, float64 accumulation)        l.280-281
// The reference materialises up_s and img_s for every scale (two full-size cubic resizes of 50 channels each); here
// one CTA owns a tile of the OUTPUT of one channel and walks the scales: it stages the few feature pixels the tile
// depends on (flip average on the way in), rebuilds the part of up_s the tile taps in shared memory, resamples it to
// the tile and adds it to float64 accumulators held in registers.  HBM sees the network outputs (about once: tiles
// overlap by the cubic halo) and one write of the float64 result; nothing intermediate exists in memory.
//
// Both resizes follow OpenCV's own float32 code path exactly (the restatement in the test tree is checked against
// cv2): coordinates fx = (float)((d+0.5)*scale - 0.5) with a double `scale`, A = -0.75 weights, replicate border at
// the edge of the image being resized (the padded map for the first resize, the CROPPED map for the second),
// horizontal pass ((s0*a0 + s1*a1) + s2*a2) + s3*a3, vertical pass s0*b0 + (s1*b1 + (s2*b2 + s3*b3)) in the 4-lane
// vector body of a row and ((s0*b0 + s1*b1) + s2*b2) + s3*b3 in its scalar tail (the last (W*C) % 4 elements of a
// row of the interleaved C-channel image; the first resize has no tail because padded widths are multiples of 64).
#pragma once
#include "ibp_device.cuh"

namespace ibp {

constexpr int kMsMaxScales = 8;
constexpr int kMsTileX = 64, kMsTileY = 16;
constexpr int kMsThreads = 256;
constexpr int kMsPerThread = kMsTileX * kMsTileY / kMsThreads;

struct MsScale {
    const float *src;   // (batch, 2, 50, h, w) network output of this scale
    int h, w;           // feature dims; padded input size = h*up, w*up
    int sh, sw;         // size of the scaled image before padding (rows/cols kept by the crop)
    double scale_y, scale_x;  // second resize: 1. / ((double)img / s), what cv2 derives from an explicit dsize
};

struct MsArgs {
    MsScale sc[kMsMaxScales];
    int n_scales;
    int up;             // model stride (first resize factor); the kernel supports 4
    int flip;           // average with the flipped twin
    int H, W;           // image size
    double *out;        // (batch, 50, H, W) planar float64: [0,30) limb maps, [30,50) keypoint maps + background
    float *heat32;      // optional (batch, 18, H, W) float32 copy of the keypoint planes (find_peaks casts, l.291)
    // shared-memory carve-up (floats), sized on the host for the largest scale
    int cap_f, cap_t1, cap_i, cap_t2;
};

// left-associated four-tap sum: the scalar tail of OpenCV's vertical pass
__device__ __forceinline__ float vsum4_tail(float s0, float s1, float s2, float s3, const float *b) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(s0, b[0]), __fmul_rn(s1, b[1])), __fmul_rn(s2, b[2])), __fmul_rn(s3, b[3]));
}

// one axis of the second resize: first tap (unclamped) and weights of destination index d
__device__ __forceinline__ int ms_tap(int d, double scale, float *w) {
    float f = __double2float_rn(__dsub_rn(__dmul_rn(__dadd_rn((double)d, 0.5), scale), 0.5));
    const int s = __float2int_rd(f);
    f = __fsub_rn(f, __int2float_rn(s));
    cubic_weights(f, w);
    return s;
}

__global__ void __launch_bounds__(kMsThreads) fuse_multiscale_kernel(MsArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int chan = blockIdx.z % kLayers, img = blockIdx.z / kLayers;
    const int X0 = blockIdx.x * kMsTileX, Y0 = blockIdx.y * kMsTileY;
    const int nX = min(kMsTileX, a.W - X0), nY = min(kMsTileY, a.H - Y0);
    // the reference resizes the 30 limb channels and the 20 keypoint channels as two interleaved images
    const int cn = (chan < kLimbs) ? kLimbs : (kLayers - kLimbs);
    const int c_local = (chan < kLimbs) ? chan : chan - kLimbs;
    const int n_vec2 = (a.W * cn) & ~3;  // elements of a destination row handled by the 4-lane vector loop

    float *F = reinterpret_cast<float *>(smem_raw);  // fused feature pixels       [n_fy][n_fx]
    float *T1 = F + a.cap_f;                         // first resize, horizontal   [n_fy][n_rx]
    float *I = T1 + a.cap_t1;                        // first resize, vertical     [n_ry][n_rx] = up_s on the tile's taps
    float *T2 = I + a.cap_i;                         // second resize, horizontal  [n_ry][nX]
    int4 *row_q = reinterpret_cast<int4 *>(T2 + a.cap_t2);   // [<= cap_rows] stage C row-group table (cap_t2 = cap_rows * tile width)
    __shared__ __align__(16) int x_idx[kMsTileX][4], y_idx[kMsTileY][4];  // read back as one 128-bit load per row
    __shared__ __align__(16) float x_w[kMsTileX][4], y_w[kMsTileY][4];

    // the four weight sets of the x4 resize (phases 1/8, 3/8, 5/8, 7/8), same evaluation as the table
    float tw[4][4];
#pragma unroll
    for (int ph = 0; ph < 4; ph++) cubic_weights(__fadd_rn(__int2float_rn(ph) * 0.25f, 0.125f), tw[ph]);

    double acc[kMsPerThread];
#pragma unroll
    for (int k = 0; k < kMsPerThread; k++) acc[k] = 0.0;
    const float n_scales_f = __int2float_rn(a.n_scales);

    for (int s = 0; s < a.n_scales; s++) {
        const MsScale S = a.sc[s];
        __syncthreads();  // previous scale's tables and tiles are no longer read
        // ---- taps of the second resize for this tile's columns and rows (clamped to the CROPPED map) ----
        if (tid < nX) {
            float w4[4];
            const int sx = ms_tap(X0 + tid, S.scale_x, w4);
            for (int k = 0; k < 4; k++) {
                x_idx[tid][k] = clampi(sx - 1 + k, 0, S.sw - 1);
                x_w[tid][k] = w4[k];
            }
        } else if (tid >= 128 && tid - 128 < nY) {
            const int j = tid - 128;
            float w4[4];
            const int sy = ms_tap(Y0 + j, S.scale_y, w4);
            for (int k = 0; k < 4; k++) {
                y_idx[j][k] = clampi(sy - 1 + k, 0, S.sh - 1);
                y_w[j][k] = w4[k];
            }
        }
        __syncthreads();
        // taps are monotonic in the destination index: the first and last entries bound the range
        const int rx_lo = x_idx[0][0], rx_hi = x_idx[nX - 1][3];
        const int ry_lo = y_idx[0][0], ry_hi = y_idx[nY - 1][3];
        const int n_rx = rx_hi - rx_lo + 1, n_ry = ry_hi - ry_lo + 1;
        // feature pixels tapped by the first resize (x`up`, up == 4: tap (r-2)>>2, phase (r-2)&3), clamped to the map
        const int fx_lo = clampi(((rx_lo - 2) >> 2) - 1, 0, S.w - 1), fx_hi = clampi(((rx_hi - 2) >> 2) + 2, 0, S.w - 1);
        const int fy_lo = clampi(((ry_lo - 2) >> 2) - 1, 0, S.h - 1), fy_hi = clampi(((ry_hi - 2) >> 2) + 2, 0, S.h - 1);
        const int n_fx = fx_hi - fx_lo + 1, n_fy = fy_hi - fy_lo + 1;

        // ---- A: feature tile, flip average on the way in (l.232-248) ----
        {
            const size_t hw = (size_t)S.h * S.w;
            const float *base = S.src + (size_t)img * 2 * kLayers * hw;
            const float *upr = base + (size_t)chan * hw;
            const float *twin = base + (size_t)(kLayers + c_flip_src[chan]) * hw;
            for (int i = tid; i < n_fy * n_fx; i += kMsThreads) {
                const int r = i / n_fx, c = i - r * n_fx;
                const int y = fy_lo + r, x = fx_lo + c;
                const float v = __ldg(upr + (size_t)y * S.w + x);
                F[i] = a.flip ? avg_like_numpy(v, __ldg(twin + (size_t)y * S.w + (S.w - 1 - x))) : v;
            }
        }
        // The x4 resize has four phases: outputs 4k+2 .. 4k+5 read the SAME four source pixels k-1 .. k+2 (clamped) and
        // differ only in the weights.  Stages B and C therefore work on groups of four outputs: four shared-memory
        // loads feed four results (the weights of the four phases sit in registers), instead of four loads each.
        const int kx_lo = (rx_lo - 2) >> 2, n_kx = ((rx_hi - 2) >> 2) - kx_lo + 1;
        const int ky_lo = (ry_lo - 2) >> 2, n_ky = ((ry_hi - 2) >> 2) - ky_lo + 1;
        // per group of intermediate rows: offsets of its four source rows in T1 (stage C)
        for (int k = tid; k < n_ky; k += kMsThreads) {
            const int sy = ky_lo + k;
            row_q[k] = make_int4((clampi(sy - 1, 0, S.h - 1) - fy_lo) * n_rx, (clampi(sy, 0, S.h - 1) - fy_lo) * n_rx,
                                 (clampi(sy + 1, 0, S.h - 1) - fy_lo) * n_rx, (clampi(sy + 2, 0, S.h - 1) - fy_lo) * n_rx);
        }
        __syncthreads();
        // ---- B: first resize, horizontal pass: feature rows x intermediate columns, four columns per item ----
        for (int i = tid; i < n_fy * n_kx; i += kMsThreads) {
            const int r = i / n_kx, k = i - r * n_kx;
            const int sx = kx_lo + k;
            const float *row = F + r * n_fx - fx_lo;
            const float f0 = row[clampi(sx - 1, 0, S.w - 1)], f1 = row[clampi(sx, 0, S.w - 1)];
            const float f2 = row[clampi(sx + 1, 0, S.w - 1)], f3 = row[clampi(sx + 2, 0, S.w - 1)];
            float *dst = T1 + r * n_rx - rx_lo;
            const int rx = 4 * sx + 2;
#pragma unroll
            for (int ph = 0; ph < 4; ph++)
                if (rx + ph >= rx_lo && rx + ph <= rx_hi) dst[rx + ph] = hsum4(f0, f1, f2, f3, tw[ph]);
        }
        __syncthreads();
        // ---- C: first resize, vertical pass (always the vector formula: padded widths are multiples of 64) ----
        // the largest stage: a flat walk over (row group, column), (k, c) advanced without a division
        {
            const int dk = kMsThreads / n_rx, dc = kMsThreads - dk * n_rx;
            int k = tid / n_rx, c = tid - k * n_rx;
            while (k < n_ky) {
                const int4 q = row_q[k];
                const float t0 = T1[q.x + c], t1 = T1[q.y + c], t2 = T1[q.z + c], t3 = T1[q.w + c];
                const int ry = 4 * (ky_lo + k) + 2;
                float *dst = I + c - ry_lo * n_rx;
#pragma unroll
                for (int ph = 0; ph < 4; ph++)
                    if (ry + ph >= ry_lo && ry + ph <= ry_hi) dst[(ry + ph) * n_rx] = vsum4(t0, t1, t2, t3, tw[ph]);
                c += dc;
                k += dk;
                if (c >= n_rx) {
                    c -= n_rx;
                    k++;
                }
            }
        }
        __syncthreads();
        // ---- D: second resize, horizontal pass: intermediate rows x output columns ----
        {
            constexpr int groups = kMsThreads / kMsTileX;
            const int g = tid / kMsTileX, c = tid % kMsTileX;
            if (c < nX) {
                const int4 xi = *reinterpret_cast<const int4 *>(x_idx[c]);
                const float4 xw4 = *reinterpret_cast<const float4 *>(x_w[c]);
                const int i0 = xi.x - rx_lo, i1 = xi.y - rx_lo, i2 = xi.z - rx_lo, i3 = xi.w - rx_lo;
                const float wv[4] = {xw4.x, xw4.y, xw4.z, xw4.w};
                for (int r = g; r < n_ry; r += groups) {
                    const float *row = I + r * n_rx;
                    T2[r * kMsTileX + c] = hsum4(row[i0], row[i1], row[i2], row[i3], wv);
                }
            }
        }
        __syncthreads();
        // ---- E: second resize, vertical pass, /n_scales in float32, accumulate in float64 (l.280-281) ----
#pragma unroll
        for (int k = 0; k < kMsPerThread; k++) {
            const int i = tid + k * kMsThreads;
            const int r = i / kMsTileX, c = i - r * kMsTileX;
            if (r < nY && c < nX) {
                const float *col = T2 + c - ry_lo * kMsTileX;
                const int4 yi = *reinterpret_cast<const int4 *>(y_idx[r]);
                const float4 yw4 = *reinterpret_cast<const float4 *>(y_w[r]);
                const float yw[4] = {yw4.x, yw4.y, yw4.z, yw4.w};
                const float s0 = col[yi.x * kMsTileX], s1 = col[yi.y * kMsTileX];
                const float s2 = col[yi.z * kMsTileX], s3 = col[yi.w * kMsTileX];
                const bool vec = ((X0 + c) * cn + c_local) < n_vec2;
                const float v = vec ? vsum4(s0, s1, s2, s3, yw) : vsum4_tail(s0, s1, s2, s3, yw);
                acc[k] = __dadd_rn(acc[k], (double)__fdiv_rn(v, n_scales_f));
            }
        }
    }

    double *out = a.out + ((size_t)img * kLayers + chan) * (size_t)a.H * a.W;
    float *h32 = (a.heat32 && chan >= kLimbs && chan < kLimbs + kParts)
                     ? a.heat32 + ((size_t)img * kParts + (chan - kLimbs)) * (size_t)a.H * a.W
                     : nullptr;
#pragma unroll
    for (int k = 0; k < kMsPerThread; k++) {
        const int i = tid + k * kMsThreads;
        const int r = i / kMsTileX, c = i - r * kMsTileX;
        if (r < nY && c < nX) {
            const size_t o = (size_t)(Y0 + r) * a.W + (X0 + c);
            out[o] = acc[k];
            if (h32) h32[o] = __double2float_rn(acc[k]);
        }
    }
}

}  // namespace ibp
