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
    int4 *row_q = reinterpret_cast<int4 *>(T2 + a.cap_t2);   // [cap_rows] stage C row table (cap_t2 = cap_rows * tile width)
    int *row_ph = reinterpret_cast<int *>(row_q + a.cap_t2 / kMsTileX);
    __shared__ int x_idx[kMsTileX][4], y_idx[kMsTileY][4];
    __shared__ float x_w[kMsTileX][4], y_w[kMsTileY][4];
    __shared__ float tab4[4][4];
    cubic_tab4_init(tab4);

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
        // per intermediate row: offsets of its four source rows in T1 and its phase (stage C)
        for (int r = tid; r < n_ry; r += kMsThreads) {
            const int ry = ry_lo + r;
            const int sy = (ry - 2) >> 2;
            row_q[r] = make_int4((clampi(sy - 1, 0, S.h - 1) - fy_lo) * n_rx, (clampi(sy, 0, S.h - 1) - fy_lo) * n_rx,
                                 (clampi(sy + 1, 0, S.h - 1) - fy_lo) * n_rx, (clampi(sy + 2, 0, S.h - 1) - fy_lo) * n_rx);
            row_ph[r] = (ry - 2) & 3;
        }
        __syncthreads();
        // Stages B-D give every thread ONE column and let it walk rows: the four taps and weights of the column live in
        // registers, there is no index division in the loops, and a warp reads consecutive columns of a row.
        // ---- B: first resize, horizontal pass: feature rows x intermediate columns ----
        if (n_rx <= kMsThreads) {
            const int groups = kMsThreads / n_rx, g = tid / n_rx, c = tid - g * n_rx;
            if (g < groups) {
                const int rx = rx_lo + c;
                const int sx = (rx - 2) >> 2;
                const float w0 = tab4[(rx - 2) & 3][0], w1 = tab4[(rx - 2) & 3][1], w2 = tab4[(rx - 2) & 3][2], w3 = tab4[(rx - 2) & 3][3];
                const float wv[4] = {w0, w1, w2, w3};
                const int i0 = clampi(sx - 1, 0, S.w - 1) - fx_lo, i1 = clampi(sx, 0, S.w - 1) - fx_lo;
                const int i2 = clampi(sx + 1, 0, S.w - 1) - fx_lo, i3 = clampi(sx + 2, 0, S.w - 1) - fx_lo;
                for (int r = g; r < n_fy; r += groups) {
                    const float *row = F + r * n_fx;
                    T1[r * n_rx + c] = hsum4(row[i0], row[i1], row[i2], row[i3], wv);
                }
            }
        } else {
            for (int i = tid; i < n_fy * n_rx; i += kMsThreads) {
                const int r = i / n_rx, c = i - r * n_rx;
                const int rx = rx_lo + c;
                const int sx = (rx - 2) >> 2;
                const float *row = F + r * n_fx - fx_lo;
                T1[i] = hsum4(row[clampi(sx - 1, 0, S.w - 1)], row[clampi(sx, 0, S.w - 1)], row[clampi(sx + 1, 0, S.w - 1)],
                              row[clampi(sx + 2, 0, S.w - 1)], tab4[(rx - 2) & 3]);
            }
        }
        __syncthreads();
        // ---- C: first resize, vertical pass (always the vector formula: padded widths are multiples of 64) ----
        // the largest stage: a flat walk over (row, column) that keeps all threads busy, (r, c) advanced without a
        // division, the four source rows and the phase of a row read from a small table filled before stage B's barrier
        {
            const int dr = kMsThreads / n_rx, dc = kMsThreads - dr * n_rx;
            int r = tid / n_rx, c = tid - r * n_rx;
            while (r < n_ry) {
                const int4 q = row_q[r];
                I[r * n_rx + c] = vsum4(T1[q.x + c], T1[q.y + c], T1[q.z + c], T1[q.w + c], tab4[row_ph[r]]);
                c += dc;
                r += dr;
                if (c >= n_rx) {
                    c -= n_rx;
                    r++;
                }
            }
        }
        __syncthreads();
        // ---- D: second resize, horizontal pass: intermediate rows x output columns ----
        {
            constexpr int groups = kMsThreads / kMsTileX;
            const int g = tid / kMsTileX, c = tid % kMsTileX;
            if (c < nX) {
                const int i0 = x_idx[c][0] - rx_lo, i1 = x_idx[c][1] - rx_lo, i2 = x_idx[c][2] - rx_lo, i3 = x_idx[c][3] - rx_lo;
                const float wv[4] = {x_w[c][0], x_w[c][1], x_w[c][2], x_w[c][3]};
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
                const float s0 = col[y_idx[r][0] * kMsTileX], s1 = col[y_idx[r][1] * kMsTileX];
                const float s2 = col[y_idx[r][2] * kMsTileX], s3 = col[y_idx[r][3] * kMsTileX];
                const bool vec = ((X0 + c) * cn + c_local) < n_vec2;
                const float v = vec ? vsum4(s0, s1, s2, s3, y_w[r]) : vsum4_tail(s0, s1, s2, s3, y_w[r]);
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
