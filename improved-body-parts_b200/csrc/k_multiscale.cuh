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
constexpr int kMsTileX = 64, kMsTileY = 24;
constexpr int kMsThreads = 256;
constexpr int kMsWarps = kMsThreads / 32;
constexpr int kMsRowsPerWarp = kMsTileY / kMsWarps;          // stage E: a warp owns this many rows of the tile

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
    float one;          // 1.0f, opaque to the compiler (see add2)
    int H, W;           // image size
    double *out;        // (batch, 50, H, W) planar float64: [0,30) limb maps, [30,50) keypoint maps + background
    float *heat32;      // optional (batch, 18, H, W) float32 copy of the keypoint planes (find_peaks casts, l.291)
    // shared-memory carve-up (floats), sized on the host for the largest scale (ms_footprint below)
    int cap_r1, cap_i;
};

// Upper bounds of one tile's footprints for a scale: the taps of T destination pixels span <= (T-1)*scale + 6 source
// pixels; they are produced in whole groups of four (the four phases of the x4 resize).
constexpr int kMsStride = 152;  // the compiled constant-stride variant: scale factors up to ~2.2 at 64-wide tiles
struct MsFootprint {
    size_t n_ry, n_kx, n_ky, r1, i;
};
inline MsFootprint ms_footprint(double scale_x, double scale_y, size_t stride /* 0 = 4 * n_kx */) {
    MsFootprint f;
    const size_t n_rx = (size_t)((kMsTileX - 1) * scale_x) + 6;
    f.n_ry = (size_t)((kMsTileY - 1) * scale_y) + 6;
    f.n_kx = n_rx / 4 + 2;
    f.n_ky = f.n_ry / 4 + 2;
    const size_t n_fx = f.n_kx + 3, n_fy = f.n_ky + 3;
    if (!stride) stride = 4 * f.n_kx;
    const size_t f_t1 = ((n_fy * n_fx + 3) & ~(size_t)3) + n_fy * stride;  // F + T1
    const size_t t2 = f.n_ry * kMsTileX;                                        // T2 reuses the same region
    f.r1 = (std::max(f_t1, t2) + 3) & ~(size_t)3;
    f.i = (4 * f.n_ky * stride + 3) & ~(size_t)3;
    return f;
}

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

// Packed float32x2 arithmetic of sm_100: FMUL2 / FFMA2 work on two floats per lane at the issue rate of the scalar
// instructions (measured, tools/ubench/f32x2.cu), so they halve the arithmetic instructions of an issue-bound kernel.
// Each half rounds like the scalar operation.  The ADD must not be written as add.rn.f32x2 or __fadd2_rn: ptxas 12.9
// contracts a packed multiply feeding a packed add into ONE FFMA2 (a single rounding) even with -fmad=false and
// explicit .rn, which breaks parity with OpenCV's separate multiply and add.  It is therefore issued as
// fma(x, one, y) with `one` = 1.0f passed in at run time: x*1 is exact, the sum is rounded once, and the compiler can
// neither simplify nor fuse it.
__device__ __forceinline__ float2 dup2(float v) { return make_float2(v, v); }
__device__ __forceinline__ float2 add2(float2 x, float2 y, float2 one) { return __ffma2_rn(x, one, y); }
// ((s0*a0 + s1*a1) + s2*a2) + s3*a3 on both halves
__device__ __forceinline__ float2 hsum4x2(float2 s0, float2 s1, float2 s2, float2 s3, float2 a0, float2 a1, float2 a2, float2 a3, float2 one) {
    return add2(add2(add2(__fmul2_rn(s0, a0), __fmul2_rn(s1, a1), one), __fmul2_rn(s2, a2), one), __fmul2_rn(s3, a3), one);
}
// s0*b0 + (s1*b1 + (s2*b2 + s3*b3)) on both halves
__device__ __forceinline__ float2 vsum4x2(float2 s0, float2 s1, float2 s2, float2 s3, float2 b0, float2 b1, float2 b2, float2 b3, float2 one) {
    return add2(__fmul2_rn(s0, b0), add2(__fmul2_rn(s1, b1), add2(__fmul2_rn(s2, b2), __fmul2_rn(s3, b3), one), one), one);
}

// i / n for 0 <= i < 2^16 and small n, with inv = 1.0f / n: (i + 0.5) * inv is at least 0.5 / n away from an integer,
// far more than the rounding error of the two float operations
__device__ __forceinline__ int ms_div(int i, float inv) { return __float2int_rz(__fmul_rn(__fadd_rn(__int2float_rn(i), 0.5f), inv)); }

// Stages per scale (all intermediates in shared memory; region R1 holds F and T1, later T2):
//   A  F  [n_fy][n_fx]    fused feature pixels of source rows ky_lo-1 .. and columns kx_lo-1 .., border replicated on load
//   B  T1 [n_fy][4 n_kx]  first resize, horizontal: item = (row, group of four columns), one 128-bit store
//   C  I  [4 n_ky][4 n_kx] first resize, vertical: a warp owns 32 columns and walks down the row groups with a sliding
//                          window of four T1 values (one load per four results)
//   D  T2 [n_ry][64]      second resize, horizontal: a warp owns 32 output columns (taps and weights in registers) and
//                          walks down the rows
//   E  acc                second resize, vertical, /n_scales, float64 accumulate in registers
// The instruction count per tile is what bounds this kernel (it is issue-bound), so every stage is written to need no
// division, no clamp and no predicate per element.
// Stage A in two halves: `ms_feature_loads` issues the global loads of up to kMsPrefetch feature pixels per thread
// (the upright value and the mirrored twin) into registers, `ms_feature_store` averages and stores them.  Called back
// to back for the first scale; for the others the loads are issued one scale ahead, before stage C, so that their
// latency is covered by C, D and E.  (Tiles wider than kMsPrefetch * 256 pixels fetch the rest at store time.)
constexpr int kMsPrefetch = 3;
struct MsFeat {
    int kx_lo, ky_lo, n_fx, n_fy;
};
__device__ __forceinline__ MsFeat ms_feat(const int *b) {
    MsFeat f;
    f.kx_lo = (b[0] - 2) >> 2;
    f.ky_lo = (b[2] - 2) >> 2;
    f.n_fx = ((b[1] - 2) >> 2) - f.kx_lo + 4;
    f.n_fy = ((b[3] - 2) >> 2) - f.ky_lo + 4;
    return f;
}
__device__ __forceinline__ void ms_planes(const MsScale &S, int img, int chan, const float *&upr, const float *&twin) {
    const size_t hw = (size_t)S.h * S.w;
    const float *base = S.src + (size_t)img * 2 * kLayers * hw;
    upr = base + (size_t)chan * hw;
    twin = base + (size_t)(kLayers + c_flip_src[chan]) * hw;
}
__device__ __forceinline__ void ms_feature_loads(const MsScale &S, const MsFeat &f, int img, int chan, bool flip, float *pu, float *pt) {
    const float *upr, *twin;
    ms_planes(S, img, chan, upr, twin);
    const float inv = __frcp_rn(__int2float_rn(f.n_fx));
    const int n = f.n_fy * f.n_fx;
#pragma unroll
    for (int j = 0; j < kMsPrefetch; j++) {
        const int i = threadIdx.x + j * kMsThreads;
        if (i < n) {
            const int r = ms_div(i, inv), c = i - r * f.n_fx;
            const int y = clampi(f.ky_lo - 1 + r, 0, S.h - 1), x = clampi(f.kx_lo - 1 + c, 0, S.w - 1);
            pu[j] = __ldg(upr + y * S.w + x);
            if (flip) pt[j] = __ldg(twin + y * S.w + (S.w - 1 - x));
        }
    }
}
__device__ __forceinline__ void ms_feature_store(const MsScale &S, const MsFeat &f, int img, int chan, bool flip, const float *pu, const float *pt,
                                                 float *F) {
    const int n = f.n_fy * f.n_fx;
#pragma unroll
    for (int j = 0; j < kMsPrefetch; j++) {
        const int i = threadIdx.x + j * kMsThreads;
        if (i < n) F[i] = flip ? avg_like_numpy(pu[j], pt[j]) : pu[j];
    }
    if (n > kMsPrefetch * kMsThreads) {
        const float *upr, *twin;
        ms_planes(S, img, chan, upr, twin);
        const float inv = __frcp_rn(__int2float_rn(f.n_fx));
        for (int i = threadIdx.x + kMsPrefetch * kMsThreads; i < n; i += kMsThreads) {
            const int r = ms_div(i, inv), c = i - r * f.n_fx;
            const int y = clampi(f.ky_lo - 1 + r, 0, S.h - 1), x = clampi(f.kx_lo - 1 + c, 0, S.w - 1);
            const float v = __ldg(upr + y * S.w + x);
            F[i] = flip ? avg_like_numpy(v, __ldg(twin + y * S.w + (S.w - 1 - x))) : v;
        }
    }
}

// STRIDE = row stride of T1 and I as a compile-time constant (row offsets become immediates), 0 = 4 * n_kx at run time.
// POW2: the number of scales is a power of two, v / n is then v * (1 / n) bit for bit.
template <int STRIDE, bool POW2>
__global__ void __launch_bounds__(kMsThreads, 4) fuse_multiscale_kernel(MsArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chan = blockIdx.z % kLayers, img = blockIdx.z / kLayers;
    const int X0 = blockIdx.x * kMsTileX, Y0 = blockIdx.y * kMsTileY;
    const int nX = min(kMsTileX, a.W - X0), nY = min(kMsTileY, a.H - Y0);
    // the reference resizes the 30 limb channels and the 20 keypoint channels as two interleaved images
    const int cn = (chan < kLimbs) ? kLimbs : (kLayers - kLimbs);
    const int c_local = (chan < kLimbs) ? chan : chan - kLimbs;
    const int n_vec2 = (a.W * cn) & ~3;  // elements of a destination row handled by the 4-lane vector loop

    float *R1 = reinterpret_cast<float *>(smem_raw);
    float *I = R1 + a.cap_r1;
    // taps of the second resize for the scale at hand: clamped source index (x: absolute; y: times the T2 row stride)
    // and weights, one 128-bit entry each
    __shared__ int4 x_idx[kMsTileX], y_idx[kMsTileY];  // y: byte offset of the T2 row
    __shared__ float4 x_w[kMsTileX], y_w[kMsTileY][2];  // y: each weight twice (packed pairs of columns)
    __shared__ int s_bounds[kMsMaxScales][4];  // per scale: first / last source column and row the tile taps

    if (tid < 4 * a.n_scales) {
        const int s = tid >> 2, which = tid & 3;
        float w4[4];
        int v;
        if (which == 0) v = clampi(ms_tap(X0, a.sc[s].scale_x, w4) - 1, 0, a.sc[s].sw - 1);
        else if (which == 1) v = clampi(ms_tap(X0 + nX - 1, a.sc[s].scale_x, w4) + 2, 0, a.sc[s].sw - 1);
        else if (which == 2) v = clampi(ms_tap(Y0, a.sc[s].scale_y, w4) - 1, 0, a.sc[s].sh - 1);
        else v = clampi(ms_tap(Y0 + nY - 1, a.sc[s].scale_y, w4) + 2, 0, a.sc[s].sh - 1);
        s_bounds[s][which] = v;
    }

    // the four weight sets of the x4 resize (phases 1/8, 3/8, 5/8, 7/8), same evaluation as the table
    float tw[4][4];
#pragma unroll
    for (int ph = 0; ph < 4; ph++) cubic_weights(__fadd_rn(__int2float_rn(ph) * 0.25f, 0.125f), tw[ph]);

    // stage E geometry: a warp owns kMsRowsPerWarp full rows of the tile, a lane two adjacent columns
    double acc[kMsRowsPerWarp][2];
#pragma unroll
    for (int k = 0; k < kMsRowsPerWarp; k++) acc[k][0] = acc[k][1] = 0.0;
    const float n_scales_f = __int2float_rn(a.n_scales);
    const float2 inv_scales2 = dup2(__fdiv_rn(1.0f, n_scales_f));
    const float2 one2 = dup2(a.one);
    const int e_col = 2 * lane, e_row0 = warp * kMsRowsPerWarp;
    const bool e_vec0 = ((X0 + e_col) * cn + c_local) < n_vec2, e_vec1 = ((X0 + e_col + 1) * cn + c_local) < n_vec2;
    const bool e_all_vec = __all_sync(0xffffffffu, e_vec0 && e_vec1);

    float pu[kMsPrefetch], pt[kMsPrefetch];  // feature pixels in flight (stage A)
    for (int s = 0; s < a.n_scales; s++) {
        const MsScale S = a.sc[s];
        __syncthreads();  // bounds written (first pass); previous scale's T2 and tap tables no longer read
        const int rx_lo = s_bounds[s][0], rx_hi = s_bounds[s][1], ry_lo = s_bounds[s][2], ry_hi = s_bounds[s][3];
        const int n_ry = ry_hi - ry_lo + 1;
        // x4 resize: destination r reads source (r-2)>>2 - 1 .. +2 with the weights of phase (r-2)&3, so destinations
        // 4k+2 .. 4k+5 ("group k") share their four sources.  Whole groups are produced.
        const int kx_lo = (rx_lo - 2) >> 2, n_kx = ((rx_hi - 2) >> 2) - kx_lo + 1;
        const int ky_lo = (ry_lo - 2) >> 2, n_ky = ((ry_hi - 2) >> 2) - ky_lo + 1;
        const int cx0 = 4 * kx_lo + 2, cy0 = 4 * ky_lo + 2;  // first column / row of I
        const int nrx = STRIDE ? STRIDE : 4 * n_kx;          // row stride of T1 and I
        const int n_fx = n_kx + 3, n_fy = n_ky + 3;
        float *F = R1;
        float *T1 = R1 + ((n_fy * n_fx + 3) & ~3);
        float *T2 = R1;

        // tap tables of this scale (read by D and E, two barriers from here)
        if (tid < kMsTileX + kMsTileY) {
            float w4[4];
            if (tid < kMsTileX) {
                const int sx = ms_tap(min(X0 + tid, a.W - 1), S.scale_x, w4);
                x_idx[tid] = make_int4(clampi(sx - 1, 0, S.sw - 1), clampi(sx, 0, S.sw - 1), clampi(sx + 1, 0, S.sw - 1), clampi(sx + 2, 0, S.sw - 1));
                x_w[tid] = make_float4(w4[0], w4[1], w4[2], w4[3]);
            } else {
                const int r = tid - kMsTileX;
                const int sy = ms_tap(min(Y0 + r, a.H - 1), S.scale_y, w4);
                constexpr int rb = kMsTileX * (int)sizeof(float);
                y_idx[r] = make_int4(clampi(sy - 1, 0, S.sh - 1) * rb, clampi(sy, 0, S.sh - 1) * rb, clampi(sy + 1, 0, S.sh - 1) * rb,
                                     clampi(sy + 2, 0, S.sh - 1) * rb);
                y_w[r][0] = make_float4(w4[0], w4[0], w4[1], w4[1]);
                y_w[r][1] = make_float4(w4[2], w4[2], w4[3], w4[3]);
            }
        }
        // ---- A: feature tile, flip average on the way in (l.232-248); the replicate border of the first resize is
        //         materialised here so that B and C index without clamps.  The loads were issued a scale ago. ----
        {
            const MsFeat f = ms_feat(s_bounds[s]);
            if (s == 0) ms_feature_loads(S, f, img, chan, a.flip != 0, pu, pt);
            ms_feature_store(S, f, img, chan, a.flip != 0, pu, pt, F);
        }
        __syncthreads();
        // ---- B: first resize, horizontal pass; the pairs are phases (0,1) and (2,3) of one group ----
        {
            const float inv = __frcp_rn(__int2float_rn(n_kx));
            for (int i = tid; i < n_fy * n_kx; i += kMsThreads) {
                const int r = ms_div(i, inv), k = i - r * n_kx;
                const float *row = F + r * n_fx + k;
                const float2 f0 = dup2(row[0]), f1 = dup2(row[1]), f2 = dup2(row[2]), f3 = dup2(row[3]);
                const float2 lo = hsum4x2(f0, f1, f2, f3, make_float2(tw[0][0], tw[1][0]), make_float2(tw[0][1], tw[1][1]),
                                          make_float2(tw[0][2], tw[1][2]), make_float2(tw[0][3], tw[1][3]), one2);
                const float2 hi = hsum4x2(f0, f1, f2, f3, make_float2(tw[2][0], tw[3][0]), make_float2(tw[2][1], tw[3][1]),
                                          make_float2(tw[2][2], tw[3][2]), make_float2(tw[2][3], tw[3][3]), one2);
                *reinterpret_cast<float4 *>(T1 + r * nrx + 4 * k) = make_float4(lo.x, lo.y, hi.x, hi.y);
            }
        }
        __syncthreads();
        if (s + 1 < a.n_scales) ms_feature_loads(a.sc[s + 1], ms_feat(s_bounds[s + 1]), img, chan, a.flip != 0, pu, pt);
        // ---- C: first resize, vertical pass (always the vector formula: a row of the x4 image has 4*w*cn elements);
        //         the pairs are two adjacent columns ----
        {
            const int n_cols = 4 * n_kx;
            // tasks = (64-column chunk, run of row groups): as many runs per chunk as gives every warp about one task
            // (8, 4, 2 or 1: a shift, the run length needs no division)
            const int n_chunks = (n_cols + 63) >> 6;
            const int run_shift = n_chunks == 1 ? 3 : (n_chunks == 2 ? 2 : (n_chunks <= 4 ? 1 : 0));
            const int run_len = (n_ky + (1 << run_shift) - 1) >> run_shift;
            const float inv = __frcp_rn(__int2float_rn(n_chunks));
            for (int t = warp; t < (n_chunks << run_shift); t += kMsWarps) {
                const int seg = ms_div(t, inv), c = (t - seg * n_chunks) * 64 + 2 * lane;
                const int k0 = seg * run_len, k1 = min(k0 + run_len, n_ky);
                if (c < n_cols && k0 < n_ky) {
                    const float *p = T1 + k0 * nrx + c;
                    float *q = I + 4 * k0 * nrx + c;
                    float2 t0 = *reinterpret_cast<const float2 *>(p), t1 = *reinterpret_cast<const float2 *>(p + nrx);
                    float2 t2 = *reinterpret_cast<const float2 *>(p + 2 * nrx);
                    p += 3 * nrx;
                    for (int k = k0; k < k1; k++) {
                        const float2 t3 = *reinterpret_cast<const float2 *>(p);
#pragma unroll
                        for (int ph = 0; ph < 4; ph++)
                            *reinterpret_cast<float2 *>(q + ph * nrx) =
                                vsum4x2(t0, t1, t2, t3, dup2(tw[ph][0]), dup2(tw[ph][1]), dup2(tw[ph][2]), dup2(tw[ph][3]), one2);
                        t0 = t1, t1 = t2, t2 = t3;
                        p += nrx;
                        q += 4 * nrx;
                    }
                }
            }
        }
        __syncthreads();
        // ---- D: second resize, horizontal pass: rows ry_lo .. ry_hi of I -> T2[n_ry][64]; the pairs are two rows ----
        {
            const int c = (warp & 1) * 32 + lane, blk = warp >> 1;
            const int rb = (n_ry + kMsWarps / 2 - 1) / (kMsWarps / 2);
            const int r0 = blk * rb, r1 = min(r0 + rb, n_ry);
            if (c < nX) {
                const int4 xi = x_idx[c];
                const float4 xw4 = x_w[c];
                const float2 w0 = dup2(xw4.x), w1 = dup2(xw4.y), w2 = dup2(xw4.z), w3 = dup2(xw4.w);
                const float *row = I + (ry_lo - cy0 + r0) * nrx - cx0;
                const float *p0 = row + xi.x, *p1 = row + xi.y, *p2 = row + xi.z, *p3 = row + xi.w;
                float *dst = T2 + r0 * kMsTileX + c;
                int r = r0;
#pragma unroll 2
                for (; r + 1 < r1; r += 2) {
                    const float2 v = hsum4x2(make_float2(p0[0], p0[nrx]), make_float2(p1[0], p1[nrx]), make_float2(p2[0], p2[nrx]),
                                             make_float2(p3[0], p3[nrx]), w0, w1, w2, w3, one2);
                    dst[0] = v.x;
                    dst[kMsTileX] = v.y;
                    p0 += 2 * nrx, p1 += 2 * nrx, p2 += 2 * nrx, p3 += 2 * nrx;
                    dst += 2 * kMsTileX;
                }
                if (r < r1) {
                    const float wv[4] = {xw4.x, xw4.y, xw4.z, xw4.w};
                    *dst = hsum4(*p0, *p1, *p2, *p3, wv);
                }
            }
        }
        __syncthreads();
        // ---- E: second resize, vertical pass, /n_scales in float32, accumulate in float64 (l.280-281);
        //         the pairs are two adjacent columns ----
        if (e_col < nX) {
            const char *col = reinterpret_cast<const char *>(T2 + e_col) - ry_lo * (kMsTileX * (int)sizeof(float));
#pragma unroll
            for (int k = 0; k < kMsRowsPerWarp; k++) {
                const int r = e_row0 + k;
                if (r < nY) {
                    const int4 yi = y_idx[r];
                    const float4 wa = y_w[r][0], wb = y_w[r][1];
                    const float2 s0 = *reinterpret_cast<const float2 *>(col + yi.x), s1 = *reinterpret_cast<const float2 *>(col + yi.y);
                    const float2 s2 = *reinterpret_cast<const float2 *>(col + yi.z), s3 = *reinterpret_cast<const float2 *>(col + yi.w);
                    float2 v = vsum4x2(s0, s1, s2, s3, make_float2(wa.x, wa.y), make_float2(wa.z, wa.w), make_float2(wb.x, wb.y),
                                       make_float2(wb.z, wb.w), one2);
                    if (!e_all_vec) {  // the scalar tail of a destination row sums left to right
                        const float yw[4] = {wa.x, wa.z, wb.x, wb.z};
                        if (!e_vec0) v.x = vsum4_tail(s0.x, s1.x, s2.x, s3.x, yw);
                        if (!e_vec1) v.y = vsum4_tail(s0.y, s1.y, s2.y, s3.y, yw);
                    }
                    if (POW2) v = __fmul2_rn(v, inv_scales2);
                    else v = make_float2(__fdiv_rn(v.x, n_scales_f), __fdiv_rn(v.y, n_scales_f));
                    acc[k][0] = __dadd_rn(acc[k][0], (double)v.x);
                    acc[k][1] = __dadd_rn(acc[k][1], (double)v.y);
                }
            }
        }
    }

    if (e_col < nX) {
        double *out = a.out + ((size_t)img * kLayers + chan) * (size_t)a.H * a.W;
        float *h32 = (a.heat32 && chan >= kLimbs && chan < kLimbs + kParts)
                         ? a.heat32 + ((size_t)img * kParts + (chan - kLimbs)) * (size_t)a.H * a.W
                         : nullptr;
        const bool two = e_col + 1 < nX, wide = two && ((a.W & 1) == 0);  // X0 and e_col are even: 16-byte aligned pairs
#pragma unroll
        for (int k = 0; k < kMsRowsPerWarp; k++) {
            const int r = e_row0 + k;
            if (r < nY) {
                const size_t o = (size_t)(Y0 + r) * a.W + (X0 + e_col);
                if (wide) {
                    *reinterpret_cast<double2 *>(out + o) = make_double2(acc[k][0], acc[k][1]);
                    if (h32) *reinterpret_cast<float2 *>(h32 + o) = make_float2(__double2float_rn(acc[k][0]), __double2float_rn(acc[k][1]));
                } else {
                    out[o] = acc[k][0];
                    if (two) out[o + 1] = acc[k][1];
                    if (h32) {
                        h32[o] = __double2float_rn(acc[k][0]);
                        if (two) h32[o + 1] = __double2float_rn(acc[k][1]);
                    }
                }
            }
        }
    }
}

}  // namespace ibp
