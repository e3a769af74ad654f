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
constexpr int kLinThreads = 128;   // threads of a CTA (the wide form; the narrow form has 64), a thread owns two adjacent output columns
constexpr int kLinSeg = 8;        // output rows per staging step
constexpr int kLinMaxSegs = 64;   // staging steps of one CTA (chain_rows <= 512)

struct __align__(16) LinTap {  // one axis, one destination index, one scale
    int base;                  // first of three consecutive feature indices (base + 2 is inside the map)
    float c[3];
};
struct LinScale {
    const float *src;    // (batch, 2, 50, h, w)
    const LinTap *xtab;  // [W]
    const LinTap *ytab;  // [H]
    const int *seg_hi;   // [ceil(H / kLinSeg)]: last feature row the output rows of a segment touch
    int h, w;
};
struct LinArgs {
    LinScale sc[kLinMaxScales];
    int nc[kLinMaxScales];    // feature columns a CTA's 256 output columns can touch (capacity of a staged row)
    int maxr[kLinMaxScales];  // feature rows a segment can add (capacity of the raw staging buffer)
    int off[kLinMaxScales];   // float offset of the scale's raw rows in the staging buffer: upright rows, then twin rows
    int hoff[kLinMaxScales];  // float offset of the scale's horizontal sums: (maxr + 4) rows of the CTA's columns
    int buf_floats;           // floats of the raw staging buffer
    int hr_floats;            // floats of the horizontal sums of all scales
    int step1;                // the taps of two adjacent output columns start at most one feature column apart (every scale <= 4)
    int ystep1;               // the same for two adjacent output rows
    int n_scales, flip, H, W;
    int chain_rows;           // output rows one CTA walks (multiple of kLinSeg)
    double *out;
    float *heat32;
};

__device__ __forceinline__ void lin_cp16(void *dst_smem, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}

// A CTA owns 256 (or 128) output columns of one channel and walks down `chain_rows` rows, kLinSeg at a time.  Per segment:
//   1. the feature rows the segment adds (and the row taps of its output rows) have been copied to shared memory with
//      16-byte cp.async while the PREVIOUS segment was computed: HBM latency is paid once per segment, under arithmetic;
//   2. horizontal pass: every thread turns the new raw rows into the three-tap sums of its two columns (flip average on
//      the way) and appends them to its own column of the HR table — thread-private shared memory, no barrier;
//   3. the raw buffer is free: the next segment's rows are requested;
//   4. vertical pass, straight-line: per output row and scale one uniform 128-bit load of the row's taps, three loads
//      from HR at immediate offsets, three packed multiply-adds.  The scales of a pixel are summed in float32 and
//      widened once (this mode is held to a tolerance, not to the bits of a float64 running sum).
// THREADS = 128 or 64: 256 or 128 output columns per CTA, whichever tiles the image width with less waste.
template <int NS, int THREADS>
__global__ void __launch_bounds__(THREADS, 384 / THREADS) fuse_linear_kernel(const __grid_constant__ LinArgs a) {
    constexpr int kLinCols = 2 * THREADS;
    extern __shared__ __align__(16) float lin_smem[];
    const int tid = threadIdx.x;
    const int X0 = blockIdx.x * kLinCols, X = X0 + 2 * tid;
    const int chan = blockIdx.z % kLayers, img = blockIdx.z / kLayers;
    const int Y0 = blockIdx.y * a.chain_rows, Y1 = min(Y0 + a.chain_rows, a.H);
    const bool flip = a.flip != 0;
    const bool warp_active = X0 + 2 * (tid & ~31) < a.W;  // a warp entirely right of the image only helps with the copies
    float *RAW = lin_smem;
    float *HR = lin_smem + a.buf_floats;
    LinTap *YT = reinterpret_cast<LinTap *>(HR + a.hr_floats);  // [2][NS][kLinSeg]
    __shared__ int s_hi[kLinMaxScales][kLinMaxSegs];            // last feature row of each segment of this CTA
    const int n_segs = (Y1 - Y0 + kLinSeg - 1) / kLinSeg;

    float2 c0[NS], c1[NS], c2[NS];  // column taps of the thread's two columns
    int colA[NS], colB[NS];         // their first feature columns, relative to the first staged one
    const float *gsrc[NS];          // what this thread copies per feature row: 16 bytes of the upright plane (and of the twin) ...
    int sdst[NS];                   // ... to this float offset of a staged row (-1: nothing)
    int gtwin[NS];                  // float offset from the upright piece to the twin piece (the same for the whole CTA)
    int nc4[NS];                    // staged columns of this CTA (a multiple of four)
    int loaded[NS];                 // last feature row requested so far
    int hbase[NS], cnt[NS];         // HR row j of the thread's column = feature row hbase + j; rows 0 .. cnt + 2 are valid
#pragma unroll
    for (int s = 0; s < NS; s++) {
        const LinScale &S = a.sc[s];
        const LinTap ta = S.xtab[min(X, a.W - 1)], tb = S.xtab[min(X + 1, a.W - 1)];
        const int c_lo = __ldg(&S.xtab[X0].base) & ~3;  // staged columns start on a 16-byte boundary
        nc4[s] = min(a.nc[s], S.w - c_lo);
        colA[s] = ta.base - c_lo, colB[s] = tb.base - c_lo;
        c0[s] = make_float2(ta.c[0], tb.c[0]), c1[s] = make_float2(ta.c[1], tb.c[1]), c2[s] = make_float2(ta.c[2], tb.c[2]);
        const size_t hw = (size_t)S.h * S.w;
        const float *base = S.src + (size_t)img * 2 * kLayers * hw;
        // thread t < nc4 / 4 copies, per feature row, 16 bytes of the upright plane (columns c_lo + 4t ..) and 16 bytes of the
        // twin plane from the range that mirrors the staged one, [w - c_lo - nc4, w - c_lo), in its own order (read back mirrored)
        if (tid < (nc4[s] >> 2)) {
            gsrc[s] = base + (size_t)chan * hw + c_lo + 4 * tid;
            sdst[s] = a.off[s] + 4 * tid;
        } else {
            gsrc[s] = base;
            sdst[s] = -1;
        }
        gtwin[s] = (int)((size_t)(kLayers + c_flip_src[chan] - chan) * hw) + S.w - 2 * c_lo - nc4[s];
        for (int g = tid; g < n_segs; g += THREADS) s_hi[s][g] = __ldg(&S.seg_hi[Y0 / kLinSeg + g]);
        loaded[s] = __ldg(&S.ytab[Y0].base) - 1;  // the first segment brings in everything from the first row's window on
        hbase[s] = loaded[s] - 2;
        cnt[s] = 0;
    }
    __syncthreads();  // s_hi

    int req_lo[NS], req_n[NS];  // what the last request brought: feature rows req_lo .. req_lo + req_n - 1, raw rows 0 ..
    // requests of segment k (rows Y0 + k * kLinSeg ..): the feature rows it adds into RAW, its row taps into YT[k & 1]
    auto request = [&](int k) {
        const int Ya = Y0 + k * kLinSeg, Yb = min(Ya + kLinSeg, Y1);
#pragma unroll
        for (int s = 0; s < NS; s++) {
            const LinScale &S = a.sc[s];
            const int f_hi = s_hi[s][k], f_lo = loaded[s] + 1;
            req_lo[s] = f_lo;
            req_n[s] = max(f_hi - f_lo + 1, 0);
            if (sdst[s] >= 0) {
                const float *g = gsrc[s] + (size_t)f_lo * S.w;
                float *d = RAW + sdst[s];
                const int dtw = a.maxr[s] * a.nc[s];
                for (int fr = f_lo; fr <= f_hi; fr++, g += S.w, d += a.nc[s]) {
                    lin_cp16(d, g);
                    if (flip) lin_cp16(d + dtw, g + gtwin[s]);
                }
            }
            loaded[s] = max(loaded[s], f_hi);
            if (tid < Yb - Ya) lin_cp16(&YT[((k & 1) * NS + s) * kLinSeg + tid], &S.ytab[Ya + tid]);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    double *out = a.out + ((size_t)img * kLayers + chan) * (size_t)a.H * a.W;
    float *h32 = (a.heat32 && chan >= kLimbs && chan < kLimbs + kParts) ? a.heat32 + ((size_t)img * kParts + (chan - kLimbs)) * (size_t)a.H * a.W : nullptr;
    const bool two = X + 1 < a.W, wide = two && ((a.W & 1) == 0);  // X is even: aligned pairs when the row length is even
    constexpr int HP = kLinCols / 2;  // row pitch of HR in float2
    request(0);
    for (int k = 0; k < n_segs; k++) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        // ---- horizontal pass: the new raw rows -> three-tap sums of the thread's two columns, appended to its HR column ----
        if (warp_active) {
#pragma unroll
            for (int s = 0; s < NS; s++) {
                const int n = req_n[s];
                if (n > 0) {
                    float2 *hr = reinterpret_cast<float2 *>(HR + a.hoff[s]) + tid;
                    const int old = cnt[s];
#pragma unroll
                    for (int i = 0; i < 3; i++) hr[i * HP] = hr[(old + i) * HP];  // the last three rows stay: a window may start there
                    const int nc = a.nc[s], tstride = a.maxr[s] * nc;
                    const float *pa = RAW + a.off[s] + colA[s];
                    const float *qa = RAW + a.off[s] + tstride + (nc4[s] - 1 - colA[s]);
                    const int d = colB[s] - colA[s];  // the second column's taps start 0 or 1 to the right (a.step1), else anywhere
                    for (int i = 0; i < n; i++, pa += nc, qa += nc) {
                        float2 f0, f1, f2;
                        if (a.step1) {  // four loads cover both columns
                            float u0 = pa[0], u1 = pa[1], u2 = pa[2], u3 = pa[3];
                            if (flip) {  // the twin row is staged in its own order: feature column c sits at nc4 - 1 - c
                                u0 = avg_like_numpy(u0, qa[0]), u1 = avg_like_numpy(u1, qa[-1]), u2 = avg_like_numpy(u2, qa[-2]), u3 = avg_like_numpy(u3, qa[-3]);
                            }
                            f0 = make_float2(u0, d ? u1 : u0), f1 = make_float2(u1, d ? u2 : u1), f2 = make_float2(u2, d ? u3 : u2);
                        } else {
                            f0 = make_float2(pa[0], pa[d]), f1 = make_float2(pa[1], pa[d + 1]), f2 = make_float2(pa[2], pa[d + 2]);
                            if (flip) {
                                const float2 half2 = make_float2(0.5f, 0.5f);
                                f0 = __fmul2_rn(__fadd2_rn(f0, make_float2(qa[0], qa[-d])), half2);
                                f1 = __fmul2_rn(__fadd2_rn(f1, make_float2(qa[-1], qa[-d - 1])), half2);
                                f2 = __fmul2_rn(__fadd2_rn(f2, make_float2(qa[-2], qa[-d - 2])), half2);
                            }
                        }
                        hr[(3 + i) * HP] = __ffma2_rn(c2[s], f2, __ffma2_rn(c1[s], f1, __fmul2_rn(c0[s], f0)));
                    }
                    cnt[s] = n;
                    hbase[s] = req_lo[s] - 3;
                }
            }
        }
        __syncthreads();  // the raw rows are consumed
        if (k + 1 < n_segs) request(k + 1);
        // ---- vertical pass of the segment's rows: straight-line ----
        if (warp_active) {
            const int Ya = Y0 + k * kLinSeg, Yb = min(Ya + kLinSeg, Y1);
            const LinTap *yt = YT + (k & 1) * NS * kLinSeg;
            const float2 *hrs[NS];
#pragma unroll
            for (int s = 0; s < NS; s++) hrs[s] = reinterpret_cast<const float2 *>(HR + a.hoff[s]) + tid;
            auto store_row = [&](int Y, float2 acc) {
                const size_t o = (size_t)Y * a.W + X;
                if (wide) {
                    *reinterpret_cast<double2 *>(out + o) = make_double2((double)acc.x, (double)acc.y);
                    if (h32) *reinterpret_cast<float2 *>(h32 + o) = acc;
                } else {
                    out[o] = (double)acc.x;
                    if (two) out[o + 1] = (double)acc.y;
                    if (h32) {
                        h32[o] = acc.x;
                        if (two) h32[o + 1] = acc.y;
                    }
                }
            };
            if (a.ystep1) {
                // two output rows at a time: their windows start on the same feature row or one apart, so four loads from HR
                // serve both (the second row's three are picked with the warp-uniform step)
                for (int Y = Ya; Y < Yb; Y += 2, yt += 2) {
                    const bool has2 = Y + 1 < Yb;
                    float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
#pragma unroll
                    for (int s = 0; s < NS; s++) {
                        const float4 e0 = *reinterpret_cast<const float4 *>(yt + s * kLinSeg);
                        const float4 e1 = *reinterpret_cast<const float4 *>(yt + s * kLinSeg + (has2 ? 1 : 0));
                        const int b0 = __float_as_int(e0.x);
                        const bool step = __float_as_int(e1.x) != b0;
                        const float2 *p = hrs[s] + (b0 - hbase[s]) * HP;
                        const float2 h0 = p[0], h1 = p[HP], h2 = p[2 * HP], h3 = p[3 * HP];
                        acc0 = __fadd2_rn(acc0, __ffma2_rn(make_float2(e0.w, e0.w), h2, __ffma2_rn(make_float2(e0.z, e0.z), h1, __fmul2_rn(make_float2(e0.y, e0.y), h0))));
                        const float2 g0 = step ? h1 : h0, g1 = step ? h2 : h1, g2 = step ? h3 : h2;
                        acc1 = __fadd2_rn(acc1, __ffma2_rn(make_float2(e1.w, e1.w), g2, __ffma2_rn(make_float2(e1.z, e1.z), g1, __fmul2_rn(make_float2(e1.y, e1.y), g0))));
                    }
                    if (X < a.W) {
                        store_row(Y, acc0);
                        if (has2) store_row(Y + 1, acc1);
                    }
                }
            } else {
                for (int Y = Ya; Y < Yb; Y++, yt++) {
                    float2 acc = make_float2(0.f, 0.f);
#pragma unroll
                    for (int s = 0; s < NS; s++) {
                        const float4 e = *reinterpret_cast<const float4 *>(yt + s * kLinSeg);  // base, c0, c1, c2 of the row (uniform)
                        const float2 *p = hrs[s] + (__float_as_int(e.x) - hbase[s]) * HP;
                        const float2 v = __ffma2_rn(make_float2(e.w, e.w), p[2 * HP], __ffma2_rn(make_float2(e.z, e.z), p[HP], __fmul2_rn(make_float2(e.y, e.y), p[0])));
                        acc = __fadd2_rn(acc, v);
                    }
                    if (X < a.W) store_row(Y, acc);
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Separable Gaussian smoothing with a mirrored border: GaussianSmoothing(channels, ksize, sigma) applied to
// F.pad(x, (p, p, p, p), mode='reflect') — utils/util.py:103-174 and its docstring.  One CTA = a 128 x 64 tile of one
// plane: tile + halo to shared memory, horizontal pass into a second buffer, vertical pass to HBM.  Reads and writes
// every pixel once: bound by HBM.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kGsTileX = 128, kGsTileY = 64, kGsMaxK = 15;
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
// K = kernel size as a compile-time constant (0: run-time size, any odd K <= kGsMaxK; the generic, scalar variant).
// K = 3, 5, 7, 9 with a row length that is a multiple of four: everything in 128-bit pieces — the tile comes in as
// float4 (a 4-column halo on either side keeps the pieces aligned), a thread makes four adjacent outputs of the
// horizontal pass from three float4 loads, and four adjacent columns of the vertical pass with a sliding window down
// four rows.
template <int K>
__global__ void __launch_bounds__(256) gauss_smooth_kernel(const __grid_constant__ GaussArgs a) {
    extern __shared__ __align__(16) float gs_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int X0 = blockIdx.x * kGsTileX, Y0 = blockIdx.y * kGsTileY;
    const float *src = a.src + (size_t)blockIdx.z * a.H * a.W;
    float *dst = a.dst + (size_t)blockIdx.z * a.H * a.W;
    if constexpr (K != 0) {
        constexpr int P = K / 2;
        constexpr int IH = kGsTileY + 2 * P;
        float *tmid = gs_smem;  // [IH][kGsTileX]: the horizontal pass of the tile and its P halo rows
        float g[K];
#pragma unroll
        for (int k = 0; k < K; k++) g[k] = a.g[k];
        // horizontal pass straight from global memory (neighbouring threads share their pieces through L1): four adjacent
        // outputs from the three float4 around them; a piece that crosses the edge of the map is put together element by
        // element through the mirror
        auto piece = [&](const float *row, int x) {
            if (x >= 0 && x + 3 < a.W) return __ldg(reinterpret_cast<const float4 *>(row + x));
            float4 v;
            v.x = __ldg(row + reflect_idx(min(x, a.W - 1 + P), a.W));
            v.y = __ldg(row + reflect_idx(min(x + 1, a.W - 1 + P), a.W));
            v.z = __ldg(row + reflect_idx(min(x + 2, a.W - 1 + P), a.W));
            v.w = __ldg(row + reflect_idx(min(x + 3, a.W - 1 + P), a.W));
            return v;
        };
        for (int i = threadIdx.x; i < IH * (kGsTileX / 4); i += 256) {
            const int r = i >> 5, q = i & 31;
            const int y = reflect_idx(min(Y0 + r - P, a.H - 1 + P), a.H), x = X0 + 4 * q;
            const float *row = src + (size_t)y * a.W;
            const float4 l = piece(row, x - 4), m = piece(row, x), h = piece(row, x + 4);
            const float in[12] = {l.x, l.y, l.z, l.w, m.x, m.y, m.z, m.w, h.x, h.y, h.z, h.w};
            float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < K; k++) {
#pragma unroll
                for (int j = 0; j < 4; j++) o[j] = __fmaf_rn(g[k], in[4 - P + j + k], o[j]);
            }
            *reinterpret_cast<float4 *>(tmid + r * kGsTileX + 4 * q) = make_float4(o[0], o[1], o[2], o[3]);
        }
        __syncthreads();
        // vertical pass: a thread = four adjacent columns x four rows, window of K rows in registers
        for (int r0 = warp * 4; r0 < kGsTileY; r0 += 32) {  // 8 warps x 4 rows per round
            if (Y0 + r0 >= a.H) break;
            const int q = lane;
            float4 win[K + 3];
#pragma unroll
            for (int k = 0; k < K + 3; k++) win[k] = *reinterpret_cast<const float4 *>(tmid + (r0 + k) * kGsTileX + 4 * q);
            const bool cols_ok = X0 + 4 * q + 3 < a.W;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int k = 0; k < K; k++) {
                    o.x = __fmaf_rn(g[k], win[j + k].x, o.x);
                    o.y = __fmaf_rn(g[k], win[j + k].y, o.y);
                    o.z = __fmaf_rn(g[k], win[j + k].z, o.z);
                    o.w = __fmaf_rn(g[k], win[j + k].w, o.w);
                }
                const int y = Y0 + r0 + j;
                if (y < a.H) {
                    float *d = dst + (size_t)y * a.W + X0 + 4 * q;
                    if (cols_ok) *reinterpret_cast<float4 *>(d) = o;
                    else {
                        if (X0 + 4 * q < a.W) d[0] = o.x;
                        if (X0 + 4 * q + 1 < a.W) d[1] = o.y;
                        if (X0 + 4 * q + 2 < a.W) d[2] = o.z;
                    }
                }
            }
        }
    } else {
        const int ksize = a.ksize, p = ksize >> 1;
        const int in_w = kGsTileX + 2 * p, in_h = kGsTileY + 2 * p;
        float *tin = gs_smem;                 // [in_h][in_w]
        float *tmid = gs_smem + in_h * in_w;  // [in_h][kGsTileX]
        for (int r = warp; r < in_h; r += 8) {
            const int y = reflect_idx(min(Y0 + r - p, a.H - 1 + p), a.H);
            const float *row = src + (size_t)y * a.W;
            for (int c = lane; c < in_w; c += 32) tin[r * in_w + c] = __ldg(row + reflect_idx(min(X0 + c - p, a.W - 1 + p), a.W));
        }
        __syncthreads();
        for (int r = warp; r < in_h; r += 8) {
            const float *q = tin + r * in_w + lane;
            float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
            for (int k = 0; k < ksize; k++) {
                const float gk = a.g[k];
                v0 = __fmaf_rn(gk, q[k], v0);
                v1 = __fmaf_rn(gk, q[k + 32], v1);
                v2 = __fmaf_rn(gk, q[k + 64], v2);
                v3 = __fmaf_rn(gk, q[k + 96], v3);
            }
            float *o = tmid + r * kGsTileX + lane;
            o[0] = v0, o[32] = v1, o[64] = v2, o[96] = v3;
        }
        __syncthreads();
        for (int r = warp; r < kGsTileY; r += 8) {
            if (Y0 + r >= a.H) break;
            const float *q = tmid + r * kGsTileX + lane;
            float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
            for (int k = 0; k < ksize; k++) {
                const float gk = a.g[k];
                v0 = __fmaf_rn(gk, q[k * kGsTileX], v0);
                v1 = __fmaf_rn(gk, q[k * kGsTileX + 32], v1);
                v2 = __fmaf_rn(gk, q[k * kGsTileX + 64], v2);
                v3 = __fmaf_rn(gk, q[k * kGsTileX + 96], v3);
            }
            float *o = dst + (size_t)(Y0 + r) * a.W + X0 + lane;
            if (X0 + lane < a.W) o[0] = v0;
            if (X0 + lane + 32 < a.W) o[32] = v1;
            if (X0 + lane + 64 < a.W) o[64] = v2;
            if (X0 + lane + 96 < a.W) o[96] = v3;
        }
    }
}

}  // namespace ibp
