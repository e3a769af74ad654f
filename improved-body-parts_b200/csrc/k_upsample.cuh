// A4 — limb-map up-sampling to the reference's full-resolution HWC layout.
// Reference: evaluate.py:77-80 / demo_image.py:94-97: cv2.resize(pafs (h,w,30), None, fx=4, fy=4, INTER_CUBIC).
// The batched pipeline never materialises this map (the limb kernel evaluates the same bicubic at its samples);
// this kernel exists for callers that want the reference's intermediate — the drop-in `process_paf(peaks, pafmap)`
// flow — and as the bit-exactness witness against cv2 at full size.  Write-bound: 30*H*W*4 bytes out per image
// (31.5 MB at 512x512) against 30*h*w*4 in.
//
// One CTA produces `up` consecutive output rows x kTileX output columns x 30 channels.  The source rows those rows
// tap (at most 5 for up=4: two phase groups share 4 rows each, offset by one) and the tile's source columns are
// staged in shared memory channel-major, then every thread owns one (X, c) column of the tile: 4..5 horizontal
// sums kept in registers, `up` vertical combinations, stores contiguous along (X, c) — the HWC inner dimensions.
#pragma once
#include "ibp_device.cuh"

namespace ibp {

constexpr int kUpTileX = 128;   // output columns per CTA
constexpr int kUpThreads = 256;
constexpr int kUpMaxRows = 6;   // source rows staged per CTA (up <= 4 needs 5)

struct UpArgs {
    const float *src;  // (batch, chan_stride, h, w) planar, limb channels first
    int chan_stride;
    int h, w, up;
    float *dst;        // (batch, h*up, w*up, 30)
};

__global__ void __launch_bounds__(kUpThreads) upsample_limbs_kernel(UpArgs a) {
    extern __shared__ float tile[];  // [30][n_rows][n_cols]
    const int img = blockIdx.z;
    const int Y0 = blockIdx.y * a.up;          // first output row of this CTA
    const int X0 = blockIdx.x * kUpTileX;      // first output column
    const int Wo = a.w * a.up, Ho = a.h * a.up;
    const int nY = min(a.up, Ho - Y0), nX = min(kUpTileX, Wo - X0);
    float wtmp[4];
    // source row / column ranges touched by this tile (unclamped first taps)
    const int sy_lo = cubic_tap(Y0, a.up, wtmp) - 1, sy_hi = cubic_tap(Y0 + nY - 1, a.up, wtmp) + 2;
    const int sx_lo = cubic_tap(X0, a.up, wtmp) - 1, sx_hi = cubic_tap(X0 + nX - 1, a.up, wtmp) + 2;
    const int n_rows = sy_hi - sy_lo + 1, n_cols = sx_hi - sx_lo + 1;
    const float *src = a.src + (size_t)img * a.chan_stride * a.h * a.w;
    for (int i = threadIdx.x; i < kLimbs * n_rows * n_cols; i += kUpThreads) {
        const int c = i / (n_rows * n_cols);
        const int rem = i - c * (n_rows * n_cols);
        const int r = rem / n_cols, q = rem - r * n_cols;
        const int sy = clampi(sy_lo + r, 0, a.h - 1), sx = clampi(sx_lo + q, 0, a.w - 1);  // replicate border
        tile[i] = __ldg(src + ((size_t)c * a.h + sy) * a.w + sx);
    }
    __syncthreads();
    float *dst = a.dst + (size_t)img * Ho * Wo * kLimbs;
    for (int i = threadIdx.x; i < nX * kLimbs; i += kUpThreads) {
        const int xl = i / kLimbs, c = i - xl * kLimbs;
        const int X = X0 + xl;
        float xw[4];
        const int sx = cubic_tap(X, a.up, xw);
        const int q0 = sx - 1 - sx_lo;
        float hrow[kUpMaxRows];
        const float *tc = tile + (size_t)c * n_rows * n_cols;
#pragma unroll
        for (int r = 0; r < kUpMaxRows; r++)
            if (r < n_rows) {
                const float *tr = tc + r * n_cols + q0;
                hrow[r] = hsum4(tr[0], tr[1], tr[2], tr[3], xw);
            }
        for (int yl = 0; yl < nY; yl++) {
            const int Y = Y0 + yl;
            float yw[4];
            const int r0 = cubic_tap(Y, a.up, yw) - 1 - sy_lo;
            float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll
            for (int r = 0; r < kUpMaxRows; r++) {
                if (r == r0) v0 = hrow[r];
                if (r == r0 + 1) v1 = hrow[r];
                if (r == r0 + 2) v2 = hrow[r];
                if (r == r0 + 3) v3 = hrow[r];
            }
            dst[((size_t)Y * Wo + X) * kLimbs + c] = vsum4(v0, v1, v2, v3, yw);
        }
    }
}

}  // namespace ibp
