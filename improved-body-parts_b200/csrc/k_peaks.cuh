// K2 — keypoint peaks: (flip average) -> non-maximum suppression -> ordered candidate lists -> sub-pixel refine.
//
// Flavour IBP_PEAKS_CROSS_REFINE restates heatmap_nms (utils/parse_skeletons.py:126-176):
//   * find_peaks_refactor (l.106-119): maximum_filter with the 4-connected cross footprint == value, AND value > 0.1
//     (strict).  scipy's default 'reflect' border repeats the edge pixel, so neighbours outside the map never win.
//     np.nonzero order = row-major (y, then x)  ->  peak ids.
//   * per peak (l.141-173): patch [y-2..y+2] x [x-2..x+2] clipped to the map, cv2.resize(patch, fx=4, fy=4, CUBIC)
//     with the border replicated AT THE PATCH EDGE, first arg-max in row-major order; x = 4*x_min + col,
//     y = 4*y_min + row (integer valued), score = the up-sampled maximum.
// Flavour IBP_PEAKS_BOX_CENTROID restates find_peaks (utils/parse_skeletons.py:286-321) with
// util.keypoint_heatmap_nms (utils/util.py:177-185: 3x3 max-pool on a reflect-padded map, value >= thre1, and the
// `heat * keep` product non-zero) and util.refine_centroid (utils/util.py:188-213, float64 moments with numpy's
// summation order); see centroid_refine below.
//
// Grid (bands, 18 parts, images).  A CTA owns a band of whole rows of one keypoint plane (+`halo` rows either side),
// stages it in shared memory — averaging the upright plane with the mirrored, channel-permuted plane of the flipped
// image on the way in when the source is the raw network output (utils/parse_skeletons.py:91-93), so the fused map
// is never written to HBM — and runs the stencil and the 5x5 patches from shared memory.  HBM sees every input value
// once: 18*h*w*sizeof(T) per image, twice that with the flip twin.
// Peaks are rare (a few per plane): threads that find one append its pixel index to a shared list, the list is
// rank-sorted (row-major == np.nonzero order), refined one warp per peak and written to a band-local list.
// The last CTA of an image to finish (ticket counter) concatenates the band lists in (part, band) order, assigns the
// running ids across parts (parse_skeletons.py:172-173) and writes the part-major "line" array
// (pafprocess.cpp:43-48) that grouping and the results index by peak id.
#pragma once
#include <climits>

#include "ibp_device.cuh"

namespace ibp {

struct PeakArgs {
    const void *src;      // planes: element (img, chan, y, x) at src + img*image_stride + chan*h*w + y*w + x
    int dtype;            // IBP_F32 / IBP_F16
    size_t image_stride;  // elements between images
    int chan0;            // channel of keypoint 0 in the upright block
    int twin_base;        // channel offset of the flipped image's block (50 for the network output), or -1: no flip
    int h, w;
    int rows_per_band, nbands, halo;
    int band_bytes;       // shared bytes reserved for the band (>= rows*w*4 and >= (18*nbands+1)*4, multiple of 16)
    int ring_bytes;       // shared bytes after the lists for the twin band (bulk-copy flip pipeline), 0: none
    int up;               // up-sampling factor of the refine step
    int win;              // patch half-size / centroid radius
    float thr;
    int cap;              // max peaks per (image, part); also per band
    int flavour;
    int raw;              // cross flavour only: heatmap_nms(bool_refine_center=False), utils/parse_skeletons.py:163-166 — no patch
                          // refine: x, y stay the FEATURE-map pixel, score = the raw map value
    DevPeak *band_peaks;  // (batch, 18, nbands, cap) scratch
    double *band_subpix;  // (batch, 18, nbands, cap, 2) scratch, centroid flavour: float64 x, y like the reference
    int *band_count;      // (batch, 18, nbands) scratch
    unsigned *ticket;     // (batch) zero on entry, left zero on exit
    DevPeak *peaks;       // (batch, 18, cap)
    double *subpix;       // (batch, 18, cap, 2)
    int *part_count;      // (batch, 18)
    ibp_peak *line;       // (batch, line_stride)
    int line_stride;
    int *n_peaks;         // (batch)
    uint32_t *status;     // (batch)
    unsigned *batch_ticket;  // one counter, zero on entry, left zero on exit
    unsigned long long *timeline;  // experiments only: (batch, 18, nbands, 6) %globaltimer stamps, or null
    int *order;           // (batch) images sorted by descending number of peaks, written by the last CTA of the launch
    int batch;
    int *yext;               // (batch, 18, 4) or null, defer_pack only: running {max(-y), max(y), max(-x), max(x)} of each part's peaks;
                             // "nothing yet" is any value below -2^30 (the index kernel resets the entries it read)
    int defer_pack;          // streaming kernel: 1 = leave the band lists as they are (no tickets, no pack, no order): the
                             // limb kernel gathers them itself; only n_raw[img] += peaks of the band
    int *n_raw;              // (batch) zero on entry when defer_pack
    unsigned *work_counter;  // streaming kernel: items beyond the first gridDim.x are claimed here, ...
    unsigned work_base;      // ... counter value at launch (it is never reset: a launch adds exactly `items` to it)
};

constexpr int kPeakThreads = 256;

__device__ __forceinline__ int block_exclusive_scan(int v, int *warp_tot, int &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    __syncthreads();  // warp_tot may still be read from a previous call
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int base = 0, tot = 0;
    const int nw = (blockDim.x + 31) >> 5;
    for (int i = 0; i < nw; i++) {
        const int t = warp_tot[i];
        if (i < warp) base += t;
        tot += t;
    }
    total = tot;
    return base + incl - v;
}

// numpy's pairwise summation for a contiguous run of n < 128 float64 values (numpy/core/src/umath/loops_utils.h
// pairwise_sum): n < 8 -> plain left-to-right loop starting from 0.; otherwise 8 accumulators over blocks of 8,
// combined as ((r0+r1)+(r2+r3)) + ((r4+r5)+(r6+r7)), then the tail added one by one.
__device__ __forceinline__ double np_pairwise_sum_f64(const double *a, int n) {
    if (n < 8) {
        double res = 0.;
        for (int i = 0; i < n; i++) res = __dadd_rn(res, a[i]);
        return res;
    }
    double r[8];
    for (int k = 0; k < 8; k++) r[k] = a[k];
    int i;
    for (i = 8; i < n - (n % 8); i += 8)
        for (int k = 0; k < 8; k++) r[k] = __dadd_rn(r[k], a[i + k]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])), __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; i++) res = __dadd_rn(res, a[i]);
    return res;
}
// ---------------------------------------------------------------------------------------------------------------
// Views of a staged band.  PlainBand: the fused values are in shared memory.  FlipBand: the upright rows and the
// flipped image's rows are both in shared memory exactly as they lie in HBM, and the average
// (utils/parse_skeletons.py:91-93) is taken when a value is read — the scan reads every value once, everything
// else touches a few pixels around a peak, so the extra pass that would fold one into the other is not needed.
// ---------------------------------------------------------------------------------------------------------------
// four consecutive elements of a staged band as they lie in memory (x % 4 == 0, rows 16-byte multiples)
__device__ __forceinline__ float4 load4(const float *p) { return *reinterpret_cast<const float4 *>(p); }
struct Half4 {
    __half x, y, z, w;
};
__device__ __forceinline__ Half4 load4(const __half *p) {
    const uint2 raw = *reinterpret_cast<const uint2 *>(p);
    Half4 h;
    h.x = __ushort_as_half((unsigned short)(raw.x & 0xffffu));
    h.y = __ushort_as_half((unsigned short)(raw.x >> 16));
    h.z = __ushort_as_half((unsigned short)(raw.y & 0xffffu));
    h.w = __ushort_as_half((unsigned short)(raw.y >> 16));
    return h;
}
template <typename T>
struct PlainBandT {
    const T *p;
    int l0, w;
    __device__ __forceinline__ float at(int y, int x) const { return to_f32(p[(y - l0) * w + x]); }
    __device__ __forceinline__ float4 at4(int y, int x) const {
        const auto v = load4(p + (y - l0) * w + x);
        return make_float4(to_f32(v.x), to_f32(v.y), to_f32(v.z), to_f32(v.w));
    }
};
template <typename T>
struct FlipBandT {
    const T *up, *tw;
    int l0, w;
    __device__ __forceinline__ float at(int y, int x) const {
        const int r = (y - l0) * w;
        return avg_like_numpy(up[r + x], tw[r + (w - 1 - x)]);
    }
    __device__ __forceinline__ float4 at4(int y, int x) const {  // x % 4 == 0, w % 4 == 0
        const int r = (y - l0) * w;
        const auto a = load4(up + r + x);
        const auto m = load4(tw + r + (w - 4 - x));
        return make_float4(avg_like_numpy(a.x, m.w), avg_like_numpy(a.y, m.z), avg_like_numpy(a.z, m.y), avg_like_numpy(a.w, m.x));
    }
};
using PlainBand = PlainBandT<float>;
using FlipBand = FlipBandT<float>;

// ---- non-maximum suppression over the owned rows [r0, r1) ----
// Neighbour rows / columns are clamped to the map, which is what both border rules amount to: scipy's 'reflect'
// and torch's reflect-pad of 1 only ever add values that are already inside the 3x3 window of the pixel, so a
// clamped (repeated) neighbour changes no comparison.  All neighbour loads are unconditional and independent
// (one round of shared-memory latency per group of four pixels instead of one per comparison).
// Threads that find a peak append its pixel index to found[] (unordered; *s_found counts, may exceed cap).
template <int FLAVOUR, class Acc>
__device__ __forceinline__ void peaks_scan(const Acc &B, int r0, int r1, int h, int w, float thr, int cap, int *found, int *s_found,
                                           int tid, int nthr) {
    constexpr bool cross = (FLAVOUR == IBP_PEAKS_CROSS_REFINE);
    auto emit = [&](bool keep, int y, int x) {
        if (keep) {
            const int slot = atomicAdd(s_found, 1);
            if (slot < cap) found[slot] = y * w + x;
        }
    };
    // centre test: heatmap_nms keeps value > 0.1 (parse_skeletons.py:116); keypoint_heatmap_nms keeps value >= thre and
    // a non-zero product heat*keep (util.py:184-185)
    auto centre_ok = [&](float v) { return cross ? (v > thr) : ((v >= thr) && (v != 0.0f)); };
    const float gate = thr;  // both flavours need v >= thr at least
    if ((w & 3) == 0) {
        const int wv = w >> 2;
        // (row, column group) of this thread's items without a division per item
        int yy = tid / wv, c = tid - yy * wv;
        const int dy = nthr / wv, dc = nthr - dy * wv;
        for (; r0 + yy < r1;) {
            const int y = r0 + yy, x = c << 2;
            const float4 v = B.at4(y, x);
            // cheap reject: nothing at or above the threshold in these four pixels
            if (fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)) >= gate) {
                const int yu = max(y - 1, 0), yd = min(y + 1, h - 1), xl = max(x - 1, 0), xr = min(x + 4, w - 1);
                const float4 u = B.at4(yu, x);
                const float4 d = B.at4(yd, x);
                const float l = B.at(y, xl), r = B.at(y, xr);
                bool k0 = centre_ok(v.x) & (v.x >= u.x) & (v.x >= d.x) & (v.x >= l) & (v.x >= v.y);
                bool k1 = centre_ok(v.y) & (v.y >= u.y) & (v.y >= d.y) & (v.y >= v.x) & (v.y >= v.z);
                bool k2 = centre_ok(v.z) & (v.z >= u.z) & (v.z >= d.z) & (v.z >= v.y) & (v.z >= v.w);
                bool k3 = centre_ok(v.w) & (v.w >= u.w) & (v.w >= d.w) & (v.w >= v.z) & (v.w >= r);
                if (!cross) {  // 3x3 window: the four diagonals as well
                    const float ul = B.at(yu, xl), ur = B.at(yu, xr), dl = B.at(yd, xl), dr = B.at(yd, xr);
                    k0 = k0 & (v.x >= ul) & (v.x >= u.y) & (v.x >= dl) & (v.x >= d.y);
                    k1 = k1 & (v.y >= u.x) & (v.y >= u.z) & (v.y >= d.x) & (v.y >= d.z);
                    k2 = k2 & (v.z >= u.y) & (v.z >= u.w) & (v.z >= d.y) & (v.z >= d.w);
                    k3 = k3 & (v.w >= u.z) & (v.w >= ur) & (v.w >= d.z) & (v.w >= dr);
                }
                emit(k0, y, x);
                emit(k1, y, x + 1);
                emit(k2, y, x + 2);
                emit(k3, y, x + 3);
            }
            yy += dy;
            c += dc;
            if (c >= wv) {
                c -= wv;
                yy++;
            }
        }
    } else {
        const int npix = (r1 - r0) * w;
        for (int i = tid; i < npix; i += nthr) {
            const int yy = i / w, x = i - yy * w, y = r0 + yy;
            const float v = B.at(y, x);
            if (v >= gate) {
                const int yu = max(y - 1, 0), yd = min(y + 1, h - 1), xl = max(x - 1, 0), xr = min(x + 1, w - 1);
                bool k = centre_ok(v) & (v >= B.at(yu, x)) & (v >= B.at(yd, x)) & (v >= B.at(y, xl)) & (v >= B.at(y, xr));
                if (!cross) k = k & (v >= B.at(yu, xl)) & (v >= B.at(yu, xr)) & (v >= B.at(yd, xl)) & (v >= B.at(yd, xr));
                emit(k, y, x);
            }
        }
    }
}

// ---- refine, one warp per peak: sorted[0..n) holds the pixel indices in row-major order ----
template <int FLAVOUR, class Acc>
__device__ __forceinline__ void peaks_refine(const Acc &B, const PeakArgs &a, int n, const int *sorted, DevPeak *out, size_t slot_base,
                                             float *hbuf, const float (*tab4)[4], int warp, int nwarps, int *yext = nullptr) {
    const int lane = threadIdx.x & 31;
    const int w = a.w, h = a.h;
    for (int k = warp; k < n; k += nwarps) {
        const int pix = sorted[k];
        const int py = pix / w, px = pix - py * w;
        if (FLAVOUR == IBP_PEAKS_CROSS_REFINE && a.raw) {
            if (lane == 0) {
                DevPeak p;
                p.x = px;
                p.y = py;
                p.score = B.at(py, px);
                p.id = k;
                out[k] = p;
            }
            __syncwarp();
        } else if (FLAVOUR == IBP_PEAKS_CROSS_REFINE && a.up == 4 && a.win == 2 && px >= 2 && px + 2 < w && py >= 2 && py + 2 < h) {
            // The common case — a full 5x5 patch up-sampled x4 to 20x20 — entirely in registers: lane X < 20 owns
            // output column X; its five horizontal sums (one per patch row) stay in registers, and the twenty vertical
            // combinations use compile-time row indices and the 4x4 phase table.  Same operations in the same order
            // as the generic path below, so the same bits.
            const int x0 = px - 2, y0 = py - 2;
            float best = -INFINITY;
            int best_idx = 0x7fffffff;
            if (lane < 20) {
                const int sx = (lane - 2) >> 2;
                const float *xw = tab4[(lane - 2) & 3];
                const int c0 = x0 + clampi(sx - 1, 0, 4), c1 = x0 + clampi(sx, 0, 4);
                const int c2 = x0 + clampi(sx + 1, 0, 4), c3 = x0 + clampi(sx + 2, 0, 4);
                float hcol[5];
#pragma unroll
                for (int r = 0; r < 5; r++) hcol[r] = hsum4(B.at(y0 + r, c0), B.at(y0 + r, c1), B.at(y0 + r, c2), B.at(y0 + r, c3), xw);
#pragma unroll
                for (int Y = 0; Y < 20; Y++) {
                    const int sy = (Y - 2) >> 2;  // compile-time after unrolling
                    const int q0 = sy - 1 < 0 ? 0 : (sy - 1 > 4 ? 4 : sy - 1), q1 = sy < 0 ? 0 : (sy > 4 ? 4 : sy);
                    const int q2 = sy + 1 < 0 ? 0 : (sy + 1 > 4 ? 4 : sy + 1), q3 = sy + 2 < 0 ? 0 : (sy + 2 > 4 ? 4 : sy + 2);
                    const float v = vsum4(hcol[q0], hcol[q1], hcol[q2], hcol[q3], tab4[(Y - 2) & 3]);
                    if (v > best) {
                        best = v;
                        best_idx = Y * 20 + lane;
                    }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, best_idx, o);
                if (ov > best || (ov == best && oi < best_idx)) {
                    best = ov;
                    best_idx = oi;
                }
            }
            if (lane == 0) {
                if (best_idx == 0x7fffffff) best_idx = 0;
                DevPeak p;
                p.x = x0 * 4 + best_idx % 20;
                p.y = y0 * 4 + best_idx / 20;
                p.score = best;
                p.id = k;
                out[k] = p;
                if (yext) {  // extent of this part's peaks (up-sampled pixels): {max of -y, max of y, max of -x, max of x}
                    atomicMax(yext, -p.y);
                    atomicMax(yext + 1, p.y);
                    atomicMax(yext + 2, -p.x);
                    atomicMax(yext + 3, p.x);
                }
            }
            __syncwarp();
        } else if (FLAVOUR == IBP_PEAKS_CROSS_REFINE) {
            const int x0 = max(px - a.win, 0), x1 = min(px + a.win, w - 1);
            const int y0 = max(py - a.win, 0), y1 = min(py + a.win, h - 1);
            const int pw = x1 - x0 + 1, ph = y1 - y0 + 1;
            const int W4 = pw * a.up, H4 = ph * a.up;
            // horizontal pass: ph rows x W4 columns
            for (int idx = lane; idx < ph * W4; idx += 32) {
                const int r = idx / W4, X = idx - r * W4;
                float xw_[4];
                const float *xw;
                int sx;
                if (a.up == 4) {
                    sx = (X - 2) >> 2;
                    xw = tab4[(X - 2) & 3];
                } else {
                    sx = cubic_tap(X, a.up, xw_);
                    xw = xw_;
                }
                const int c0 = x0 + clampi(sx - 1, 0, pw - 1), c1 = x0 + clampi(sx, 0, pw - 1);
                const int c2 = x0 + clampi(sx + 1, 0, pw - 1), c3 = x0 + clampi(sx + 2, 0, pw - 1);
                const int y = y0 + r;
                hbuf[r * W4 + X] = hsum4(B.at(y, c0), B.at(y, c1), B.at(y, c2), B.at(y, c3), xw);
            }
            __syncwarp();
            float best = -INFINITY;
            int best_idx = 0x7fffffff;
            if (W4 <= 32 && a.up == 4) {
                // lane = column: the four taps and weights of a row are the same for every lane, no index division
                if (lane < W4) {
                    for (int Y = 0; Y < H4; Y++) {
                        const int sy = (Y - 2) >> 2;
                        const float *yw = tab4[(Y - 2) & 3];
                        const int q0 = clampi(sy - 1, 0, ph - 1), q1 = clampi(sy, 0, ph - 1);
                        const int q2 = clampi(sy + 1, 0, ph - 1), q3 = clampi(sy + 2, 0, ph - 1);
                        const float v = vsum4(hbuf[q0 * W4 + lane], hbuf[q1 * W4 + lane], hbuf[q2 * W4 + lane], hbuf[q3 * W4 + lane], yw);
                        if (v > best) {
                            best = v;
                            best_idx = Y * W4 + lane;
                        }
                    }
                }
            } else
            for (int idx = lane; idx < H4 * W4; idx += 32) {
                const int Y = idx / W4, X = idx - Y * W4;
                float yw_[4];
                const float *yw;
                int sy;
                if (a.up == 4) {
                    sy = (Y - 2) >> 2;
                    yw = tab4[(Y - 2) & 3];
                } else {
                    sy = cubic_tap(Y, a.up, yw_);
                    yw = yw_;
                }
                const int q0 = clampi(sy - 1, 0, ph - 1), q1 = clampi(sy, 0, ph - 1);
                const int q2 = clampi(sy + 1, 0, ph - 1), q3 = clampi(sy + 2, 0, ph - 1);
                const float v = vsum4(hbuf[q0 * W4 + X], hbuf[q1 * W4 + X], hbuf[q2 * W4 + X], hbuf[q3 * W4 + X], yw);
                if (v > best) {
                    best = v;
                    best_idx = idx;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, best_idx, o);
                if (ov > best || (ov == best && oi < best_idx)) {
                    best = ov;
                    best_idx = oi;
                }
            }
            if (lane == 0) {
                if (best_idx == 0x7fffffff) best_idx = 0;
                DevPeak p;
                p.x = x0 * a.up + best_idx % W4;
                p.y = y0 * a.up + best_idx / W4;
                p.score = best;
                p.id = k;
                out[k] = p;
                if (yext) {  // extent of this part's peaks (up-sampled pixels): {max of -y, max of y, max of -x, max of x}
                    atomicMax(yext, -p.y);
                    atomicMax(yext + 1, p.y);
                    atomicMax(yext + 2, -p.x);
                    atomicMax(yext + 3, p.x);
                }
            }
            __syncwarp();
        } else {
            // util.refine_centroid (util.py:188-213), radius = win
            if (lane == 0) {
                const int rad = a.win;
                double fx, fy;
                float sc;
                if (py + rad + 1 > h || py - rad < 0 || px + rad + 1 > w || px - rad < 0) {
                    fx = (double)px;
                    fy = (double)py;
                    sc = B.at(py, px);
                } else {
                    // box*grid promotes to float64 (grid is int64): 25 products summed by numpy's pairwise scheme in
                    // float64; box.sum() is the same pairwise scheme over the 25 float32 values in row-major order
                    // (measured against numpy: the strided 5x5 view is reduced as one flat run, not row by row);
                    // box.mean() = that float32 sum / 25 in float32.
                    double prod_r[25 * 1], prod_c[25 * 1];
                    float vals[25];
                    const int sidec = 2 * rad + 1;
                    int cnt = 0;
                    for (int q = -rad; q <= rad; q++) {
                        for (int p = -rad; p <= rad; p++) {
                            const float v = B.at(py + q, px + p);
                            if (cnt < 25) {
                                vals[cnt] = v;
                                prod_r[cnt] = __dmul_rn((double)v, (double)q);  // x_grid = ROW offset (l.207 quirk)
                                prod_c[cnt] = __dmul_rn((double)v, (double)p);  // y_grid = column offset
                            }
                            cnt++;
                        }
                    }
                    float racc[8];
                    for (int k = 0; k < 8; k++) racc[k] = vals[k];
                    for (int i = 8; i < 24; i += 8)
                        for (int k = 0; k < 8; k++) racc[k] = __fadd_rn(racc[k], vals[i + k]);
                    float total = __fadd_rn(__fadd_rn(__fadd_rn(racc[0], racc[1]), __fadd_rn(racc[2], racc[3])),
                                            __fadd_rn(__fadd_rn(racc[4], racc[5]), __fadd_rn(racc[6], racc[7])));
                    total = __fadd_rn(total, vals[24]);
                    const int cntp = min(cnt, 25);
                    const double denom = (double)total;
                    const double offx = __ddiv_rn(np_pairwise_sum_f64(prod_r, cntp), denom);
                    const double offy = __ddiv_rn(np_pairwise_sum_f64(prod_c, cntp), denom);
                    fx = __dadd_rn((double)px, offx);  // x_c + offset_x in float64 (util.py:210)
                    fy = __dadd_rn((double)py, offy);
                    sc = __fdiv_rn(total, (float)(sidec * sidec));
                }
                DevPeak p;
                p.x = (int)fx;  // truncated like the C++ route would (pafprocess.cpp:35-36)
                p.y = (int)fy;
                p.score = sc;
                p.id = k;
                out[k] = p;
                if (yext) {  // extent of this part's peaks (up-sampled pixels): {max of -y, max of y, max of -x, max of x}
                    atomicMax(yext, -p.y);
                    atomicMax(yext + 1, p.y);
                    atomicMax(yext + 2, -p.x);
                    atomicMax(yext + 3, p.x);
                }
                if (a.band_subpix) {
                    double *sp = a.band_subpix + (slot_base + k) * 2;
                    sp[0] = fx;
                    sp[1] = fy;
                }
            }
            __syncwarp();
        }
    }
}

// ---- the last CTA of an image to finish packs the image; called by warp 0 after the whole CTA's writes are fenced ----
// scratch: shared memory, >= max(18*nbands + 1, batch) ints.
__device__ __forceinline__ void peaks_tail(const PeakArgs &a, int img, int *scratch) {
    const int lane = threadIdx.x & 31;
    const unsigned ctas_per_image = (unsigned)a.nbands * kParts;
    int last = 0;
    if (lane == 0) {
        const unsigned t = atomicAdd(a.ticket + img, 1u);
        last = (t == ctas_per_image - 1);
        if (last) a.ticket[img] = 0;  // ready for the next launch
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    __threadfence();

    // exclusive scan of the band counts in (part, band) order; re-use the band buffer (>= 18*nbands+1 ints by launch)
    int *offs = scratch;
    const int items = kParts * a.nbands;
    const int *bc = a.band_count + (size_t)img * items;
    int total = 0;
    for (int base = 0; base < items; base += 32) {
        const int i = base + lane;
        const int c = (i < items) ? __ldcg(bc + i) : 0;
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (i < items) offs[i] = total + incl - c;
        total += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) offs[items] = total;
    __syncwarp();
    // per-part counts, clamped to the per-part capacity
    bool over = false;
    if (lane < kParts) {
        const int c = offs[(lane + 1) * a.nbands] - offs[lane * a.nbands];
        over = c > a.cap;
        a.part_count[img * kParts + lane] = min(c, a.cap);
    }
    if (__any_sync(0xffffffffu, over)) {
        // ids would no longer be a running counter: flag it; ibp_fetch turns the flag into IBP_ERR_CAPACITY
        if (lane == 0) atomicOr(a.status + img, IBP_ST_PEAK_OVERFLOW);
    }
    if (lane == 0) a.n_peaks[img] = min(total, a.line_stride);
    for (int j = lane; j < total; j += 32) {
        // item holding global position j: last i with offs[i] <= j
        int lo = 0, hi = items - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (offs[mid] <= j) lo = mid; else hi = mid - 1;
        }
        const int p = lo / a.nbands;
        const int k_in_band = j - offs[lo];
        const int k_in_part = j - offs[p * a.nbands];
        const size_t srci = ((size_t)img * items + lo) * a.cap + k_in_band;
        const int4 raw = __ldcg(reinterpret_cast<const int4 *>(a.band_peaks + srci));
        DevPeak pk;
        pk.x = raw.x;
        pk.y = raw.y;
        pk.score = __int_as_float(raw.z);
        pk.id = j;
        if (k_in_part < a.cap) {
            const size_t dsti = ((size_t)img * kParts + p) * a.cap + k_in_part;
            a.peaks[dsti] = pk;
            if (a.band_subpix && a.subpix) {
                const double2 sp = __ldcg(reinterpret_cast<const double2 *>(a.band_subpix + srci * 2));
                a.subpix[dsti * 2] = sp.x;
                a.subpix[dsti * 2 + 1] = sp.y;
            }
        }
        if (j < a.line_stride) {
            ibp_peak o;
            o.x = pk.x;
            o.y = pk.y;
            o.score = pk.score;
            o.part = p;
            a.line[(size_t)img * a.line_stride + j] = o;
        }
    }

    // ---- the last image to be packed ranks the whole batch: heaviest image first ----
    // The limb kernel walks images in this order so that the longest grouping chains (a sequential tail of one warp
    // per image) start first and overlap the rest of the batch instead of trailing it.
    __threadfence();
    __syncwarp();
    if (lane == 0) {
        const unsigned t = atomicAdd(a.batch_ticket, 1u);
        last = (t == (unsigned)a.batch - 1);
        if (last) *a.batch_ticket = 0;
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    __threadfence();
    // counts into shared memory first (one round of loads), then an all-pairs rank from there
    int *cnt = scratch;  // band buffer, free again; launch guarantees >= batch ints
    for (int b = lane; b < a.batch; b += 32) cnt[b] = __ldcg(a.n_peaks + b);
    __syncwarp();
    for (int b = lane; b < a.batch; b += 32) {
        const int mine_n = cnt[b];
        int rank = 0;
        for (int u = 0; u < a.batch; u++) {
            const int other = cnt[u];
            rank += (other > mine_n) || (other == mine_n && u < b);
        }
        a.order[rank] = b;
    }
}

// Shared-memory carve-up common to both kernels, after the band buffers: per-warp patch buffer, found[], sorted[].
struct PeakScratch {
    float *hbuf_all;
    int hbuf_per_warp;
    int *found, *sorted;
    size_t end;
};
__device__ __forceinline__ PeakScratch peak_scratch(unsigned char *smem, size_t off, const PeakArgs &a, int nwarps) {
    PeakScratch s;
    const int side = 2 * a.win + 1;
    s.hbuf_per_warp = side * side * a.up;
    s.hbuf_all = reinterpret_cast<float *>(smem + off);  // per warp: side rows x side*up columns
    off += ((size_t)nwarps * s.hbuf_per_warp * 4 + 7) & ~(size_t)7;
    s.found = reinterpret_cast<int *>(smem + off);  // [cap] unordered pixel indices
    off += (size_t)a.cap * 4;
    s.sorted = reinterpret_cast<int *>(smem + off);  // [cap] row-major order
    off += (size_t)a.cap * 4;
    s.end = (off + 15) & ~(size_t)15;
    return s;
}

// rank sort of the found pixel indices (distinct) into row-major order == np.nonzero order; all threads
__device__ __forceinline__ void peaks_order(const int *found, int *sorted, int n, int tid, int nthr) {
    for (int i = tid; i < n; i += nthr) {
        const int mine = found[i];
        int rank = 0;
        for (int j = 0; j < n; j++) rank += (found[j] < mine);
        sorted[rank] = mine;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// General kernel: one CTA per (band, part, image); any element type, any alignment, with or without the flip twin.
// FLAVOUR is a template parameter so that the heatmap_nms instantiation does not carry the registers of the float64
// centroid code (and the other way round).
// ---------------------------------------------------------------------------------------------------------------
template <typename T, int FLAVOUR>
__global__ void __launch_bounds__(kPeakThreads, 5) peaks_kernel(PeakArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int bandi = blockIdx.x, part = blockIdx.y, img = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int nwarps = kPeakThreads / 32;
    const int w = a.w, h = a.h;

    const int r0 = bandi * a.rows_per_band, r1 = min(r0 + a.rows_per_band, h);
    const int l0 = max(r0 - a.halo, 0), l1 = min(r1 + a.halo, h);

    auto stamp = [&](int col) {
        if (a.timeline && tid == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            a.timeline[((((size_t)img * kParts + part) * a.nbands) + bandi) * 6 + col] = t;
        }
    };
    stamp(0);
    float *band = reinterpret_cast<float *>(smem_raw);
    const PeakScratch sc = peak_scratch(smem_raw, (size_t)a.band_bytes, a, nwarps);
    float *ring = reinterpret_cast<float *>(smem_raw + sc.end);  // twin band while staging (ring_bytes)
    __shared__ uint64_t bars[2];
    __shared__ int s_found;
    __shared__ float tab4[4][4];

    {
        const T *base = reinterpret_cast<const T *>(a.src) + (size_t)img * a.image_stride;
        const size_t hw = (size_t)h * w;
        const T *up = base + (size_t)(a.chan0 + part) * hw;
        // flipped image: channel 30 + flip_heat_ord[part] of its block (config/config.py:150-152)
        const T *twin = (a.twin_base >= 0) ? base + (size_t)(a.twin_base + c_flip_src[kLimbs + part]) * hw : nullptr;
        const int nrows = l1 - l0;
        // float32 + twin: both bands by the bulk-copy engine, the twin into its own buffer (one ring slot holds it all)
        const int slot_rows = (a.ring_bytes >= nrows * w * 4) ? nrows : 0;
        bool tma_flip = false;
        if (sizeof(T) == 4) tma_flip = flip_tma_ok(band, reinterpret_cast<const float *>(up), reinterpret_cast<const float *>(twin), ring, w, slot_rows);
        if (tid == 0) {
            mbar_init(&bars[0], 1);
            mbar_init(&bars[1], 1);
            fence_barrier_init();
            s_found = 0;
            if (tma_flip)
                flip_tma_issue<1>(band, reinterpret_cast<const float *>(up), reinterpret_cast<const float *>(twin), l0, nrows, w, ring, slot_rows, bars);
        }
        cubic_tab4_init(tab4);
        __syncthreads();
        if (tma_flip) {
            flip_tma_finish<1>(band, reinterpret_cast<const float *>(twin), l0, nrows, w, ring, slot_rows, bars);
        } else {
            uint32_t phase = 0;
            stage_rows_fused<T>(band, up, twin, l0, nrows, w, &bars[0], phase);
        }
    }
    stamp(1);
    const PlainBand B{band, l0, w};
    peaks_scan<FLAVOUR>(B, r0, r1, h, w, a.thr, a.cap, sc.found, &s_found, tid, kPeakThreads);
    __syncthreads();
    stamp(2);
    const int n_raw = s_found;
    const int n = min(n_raw, a.cap);
    peaks_order(sc.found, sc.sorted, n, tid, kPeakThreads);
    __syncthreads();

    const size_t slot_base = (((size_t)img * kParts + part) * a.nbands + bandi) * a.cap;
    peaks_refine<FLAVOUR>(B, a, n, sc.sorted, a.band_peaks + slot_base, slot_base, sc.hbuf_all + warp * sc.hbuf_per_warp, tab4, warp, nwarps);
    if (tid == 0) {
        a.band_count[((size_t)img * kParts + part) * a.nbands + bandi] = n;
        if (n_raw > a.cap) atomicOr(a.status + img, IBP_ST_PEAK_OVERFLOW);
    }

    stamp(3);
    // ---- last CTA of this image packs (warp 0 only; the other warps leave as soon as their writes are out) ----
    if (lane == 0 && (warp < n || tid == 0)) __threadfence();  // the lanes that wrote peaks / the count
    if (warp != 0) {
        asm volatile("bar.arrive 1, %0;" ::"n"(kPeakThreads) : "memory");
        return;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kPeakThreads) : "memory");
    stamp(4);
    peaks_tail(a, img, reinterpret_cast<int *>(smem_raw));  // band buffer: >= max(18*nbands+1, batch) ints by launch
    stamp(5);
}


// ---------------------------------------------------------------------------------------------------------------
// Streaming kernel (float32, 16-byte aligned rows): a persistent grid of warp-specialised CTAs over the
// (image, part, band) items, TWO sets of band buffers per CTA.
//   * warp 0, the control warp: claims items, keeps the bulk-copy engine one item ahead of the workers (the upright
//     band — and the flipped image's band, TWIN — of item k+1 lands in the other buffer set while item k is worked
//     on; item k+2 is requested the moment the workers release item k's set), publishes finished bands (one fence +
//     one ticket per item) and packs an image when its last band is in.  All of its latencies (claim, fence, ticket,
//     pack) overlap the workers' next item.
//   * warps 1..7, the workers: wait for the data, scan, order, refine, write the band list, release the buffer set.
//     They never fence and never wait for the control warp except for data.
// With TWIN the average is taken on read (FlipBand): no fold pass.
// Work distribution: item blockIdx.x first, then whatever the shared counter hands out — a CTA whose control warp is
// packing an image simply claims fewer bands (with a fixed schedule the laggard becomes the last CTA, hence the
// packer, of every image it touches).  Every CTA makes exactly one claim that fails, so a launch advances the
// counter by exactly `items`; the host tracks the base instead of resetting it.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kPeakStreamCtasPerSm = 3;
constexpr int kPeakWorkers = kPeakThreads - 32;

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int FLAVOUR, bool TWIN, typename T = float>
__global__ void __launch_bounds__(kPeakThreads, kPeakStreamCtasPerSm) peaks_stream_kernel(PeakArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int nworkers = kPeakWorkers / 32;
    constexpr int nplanes = TWIN ? 2 : 1;
    const int w = a.w, h = a.h;
    const size_t hw = (size_t)h * w;
    const int per_image = a.nbands * kParts;
    const int items = per_image * a.batch;

    // buffer set b: upright band at b*nplanes*band_bytes, twin band right after it
    auto buf_up = [&](int b) { return reinterpret_cast<T *>(smem_raw + (size_t)b * nplanes * a.band_bytes); };
    auto buf_tw = [&](int b) { return reinterpret_cast<T *>(smem_raw + ((size_t)b * nplanes + 1) * a.band_bytes); };
    const PeakScratch sc = peak_scratch(smem_raw, (size_t)2 * nplanes * a.band_bytes, a, nworkers);
    int *tail_scratch = reinterpret_cast<int *>(smem_raw + sc.end);  // max(18*nbands+1, batch) ints
    __shared__ uint64_t full[2];  // data of buffer set b has landed (or: no more items, see s_item)
    __shared__ uint64_t done[2];  // the workers are finished with buffer set b
    __shared__ int s_found[2];
    __shared__ int s_item[2];     // item in buffer set b, -1: none (the workers leave)
    __shared__ float tab4[4][4];

    struct Item {
        int img, part, bandi, r0, r1, l0, l1;
    };
    auto decode = [&](int item) {
        Item it;
        it.img = item / per_image;
        const int rem = item - it.img * per_image;
        it.part = rem / a.nbands;
        it.bandi = rem - it.part * a.nbands;
        it.r0 = it.bandi * a.rows_per_band;
        it.r1 = min(it.r0 + a.rows_per_band, h);
        it.l0 = max(it.r0 - a.halo, 0);
        it.l1 = min(it.r1 + a.halo, h);
        return it;
    };
    // one thread: publish `item` as the content of buffer set b and start its copies (or wake the workers up to leave)
    auto request = [&](int item, int b) {
        s_item[b] = item;
        s_found[b] = 0;
        if (item < 0) {
            mbar_arrive(&full[b]);
            return;
        }
        const Item it = decode(item);
        const T *base = reinterpret_cast<const T *>(a.src) + (size_t)it.img * a.image_stride;
        const uint32_t bytes = (uint32_t)(it.l1 - it.l0) * w * (uint32_t)sizeof(T);
        mbar_expect_tx(&full[b], bytes * nplanes);
        bulk_chunks(buf_up(b), base + (size_t)(a.chan0 + it.part) * hw + (size_t)it.l0 * w, bytes, &full[b]);
        if (TWIN)  // flipped image: channel 30 + flip_heat_ord[part] of its block (config/config.py:150-152)
            bulk_chunks(buf_tw(b), base + (size_t)(a.twin_base + c_flip_src[kLimbs + it.part]) * hw + (size_t)it.l0 * w, bytes, &full[b]);
    };
    auto claim_to_item = [&](unsigned t) {
        const unsigned rel = t - a.work_base + gridDim.x;
        return rel < (unsigned)items ? (int)rel : -1;
    };
    auto stamp = [&](int item, int col) {
        if (a.timeline) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            a.timeline[(size_t)item * 6 + col] = t;
        }
    };

    pdl_launch_dependents();  // the index kernel may be set up while this grid runs (it waits for our results itself)
    int cur = (int)blockIdx.x, nxt = -1;  // control warp's view (the grid never exceeds the item count)
    if (tid == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        mbar_init(&done[0], nworkers);
        mbar_init(&done[1], nworkers);
        fence_barrier_init();
        request(cur, 0);
        nxt = claim_to_item(atomicAdd(a.work_counter, 1u));
        request(nxt, 1);
    }
    cubic_tab4_init(tab4);
    __syncthreads();

    if (warp == 0) {
        // ------------------------------------------------------------------------------------------ control warp
        nxt = __shfl_sync(0xffffffffu, nxt, 0);
        for (int k = 0; cur >= 0; k++) {
            const int b = k & 1;
            // claim the item after next now, use the answer when this one is released (stop at the first failure)
            unsigned claimed = 0;
            const bool claiming = (lane == 0) && (nxt >= 0);
            if (claiming) claimed = atomicAdd(a.work_counter, 1u);
            mbar_wait(&done[b], (uint32_t)((k >> 1) & 1));
            int next2 = -1;
            if (lane == 0) {
                stamp(cur, 4);
                next2 = claiming ? claim_to_item(claimed) : -1;
                fence_proxy_async();  // the workers' generic-proxy reads of this set come before the refill
                request(next2, b);
                if (!a.defer_pack) __threadfence();  // the band list the workers wrote (observed through done[b]) before the ticket
            }
            __syncwarp();
            if (!a.defer_pack) peaks_tail(a, cur / per_image, tail_scratch);
            if (lane == 0) stamp(cur, 5);
            cur = nxt;
            nxt = __shfl_sync(0xffffffffu, next2, 0);
        }
        return;
    }

    // ---------------------------------------------------------------------------------------------------- workers
    const int wtid = tid - 32, wwarp = warp - 1;
    auto workers_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(kPeakWorkers) : "memory"); };
    for (int k = 0;; k++) {
        const int b = k & 1;
        unsigned long long t_wait = 0;
        if (a.timeline && wtid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_wait));
        mbar_wait(&full[b], (uint32_t)((k >> 1) & 1));
        const int item = *reinterpret_cast<volatile int *>(&s_item[b]);
        if (item < 0) break;
        if (wtid == 0) {
            if (a.timeline) a.timeline[(size_t)item * 6] = t_wait;
            stamp(item, 1);
        }
        const Item it = decode(item);
        typename std::conditional<TWIN, FlipBandT<T>, PlainBandT<T>>::type B;
        if constexpr (TWIN) {
            B.up = buf_up(b);
            B.tw = buf_tw(b);
        } else {
            B.p = buf_up(b);
        }
        B.l0 = it.l0;
        B.w = w;
        peaks_scan<FLAVOUR>(B, it.r0, it.r1, h, w, a.thr, a.cap, sc.found, &s_found[b], wtid, kPeakWorkers);
        workers_sync();
        if (wtid == 0) stamp(item, 2);
        const int n_raw = *reinterpret_cast<volatile int *>(&s_found[b]);
        const int n = min(n_raw, a.cap);
        peaks_order(sc.found, sc.sorted, n, wtid, kPeakWorkers);
        workers_sync();
        const size_t slot_base = (size_t)item * a.cap;  // == ((img*18 + part)*nbands + band) * cap
        peaks_refine<FLAVOUR>(B, a, n, sc.sorted, a.band_peaks + slot_base, slot_base, sc.hbuf_all + wwarp * sc.hbuf_per_warp, tab4, wwarp, nworkers,
                              (a.defer_pack && a.yext) ? a.yext + ((size_t)it.img * kParts + it.part) * 4 : nullptr);
        if (wtid == 0) {
            a.band_count[item] = n;
            if (n_raw > a.cap) atomicOr(a.status + it.img, IBP_ST_PEAK_OVERFLOW);
            if (a.defer_pack && n > 0) atomicAdd(a.n_raw + it.img, n);
            stamp(item, 3);
        }
        // release the buffer set (and, through the control warp's fence, the band list): one arrival per warp
        __syncwarp();
        if (lane == 0) mbar_arrive(&done[b]);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// After a streaming launch that left the band lists unpacked: one warp per image counts them — per-part totals
// (clamped to the capacity), the first running id of every part (parse_skeletons.py:172-173), the image total — and
// ranks its image among the batch by number of peaks (heaviest first, ties by index), which is the order the limb
// kernel walks the images in.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kGroupLag = 16;  // images between the limbs of an image and its grouping in the streaming limb kernel's queue

struct IndexArgs {
    const int *band_count;  // (batch, 18, nbands)
    const int *n_raw;       // (batch)
    int nbands, batch, cap, line_stride;
    int *part_count;        // (batch, 18)
    int *partbase;          // (batch, 19)
    int *n_peaks;           // (batch)
    int *order;             // (batch)
    int4 *item_meta;        // (batch*31) or null: work queue of the streaming limb kernel, images by rank (0 = most peaks); z = limb | (c0/4) << 8 | (c1/4) << 20:
                            // {image, nA | nB << 16, limb, r0 | r1 << 16} for the 30 limbs of an image — [r0, r1) = the
                            // feature rows any sample of the limb can touch — and {image, 0, 30, 0} = "group it",
                            // queued kGroupLag images after its limbs
    int *yext;              // (batch, 18, 4) from the peak kernel ({max -y, max y, max -x, max x} per part), reset here
    int w;                  // feature width
    int h, up;              // feature rows, up-sampling factor
    uint32_t *status;       // (batch)
    int group_lag;          // images between the limbs of an image and its "group it" item in the queue (>= 0)
};

__global__ void __launch_bounds__(32) index_bands_kernel(IndexArgs a) {
    const int img = blockIdx.x, lane = threadIdx.x;
    pdl_launch_dependents();  // the limb kernel's CTAs may take their places
    pdl_wait();               // the peak kernel's band lists
    const int items = kParts * a.nbands;
    const int *bc = a.band_count + (size_t)img * items;
    // every load first: the band counts of this image, the raw totals of the batch
    constexpr int kRounds = (kParts * 16 + 31) / 32;  // nbands <= 16 (kMaxDeferBands)
    int cv[kRounds];
#pragma unroll
    for (int r = 0; r < kRounds; r++) cv[r] = (r * 32 + lane < items) ? bc[r * 32 + lane] : 0;
    int ext_lo = 0, ext_hi = 0, ext_xlo = 0, ext_xhi = 0;  // y and x extent of part `lane`
    if (a.item_meta && lane < kParts) {
        const int *e = a.yext + ((size_t)img * kParts + lane) * 4;
        ext_lo = e[0];
        ext_hi = e[1];
        ext_xlo = e[2];
        ext_xhi = e[3];
    }
    const int mine = a.n_raw[img];
    int rank = 0;
    for (int u0 = 0; u0 < a.batch; u0 += 32) {
        const int u = u0 + lane;
        const int other = (u < a.batch) ? a.n_raw[u] : -1;
        rank += __popc(__ballot_sync(0xffffffffu, (other > mine) || (other == mine && u < img && u < a.batch)));
    }
    if (lane == 0) a.order[rank] = img;

    __shared__ int offs[kParts * 16 + 1];
    int total = 0;
#pragma unroll
    for (int r = 0; r < kRounds; r++) {
        const int i = r * 32 + lane;
        int incl = cv[r];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (i < items) offs[i] = total + incl - cv[r];
        total += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) offs[items] = total;
    __syncwarp();
    bool over = false;
    if (lane < kParts) {
        const int first = offs[lane * a.nbands];
        const int c = offs[(lane + 1) * a.nbands] - first;
        over = c > a.cap;
        a.part_count[img * kParts + lane] = min(c, a.cap);
        a.partbase[img * (kParts + 1) + lane] = first;
    }
    if (lane == 0) {
        a.partbase[img * (kParts + 1) + kParts] = total;
        a.n_peaks[img] = min(total, a.line_stride);
    }
    if (a.item_meta) {
        // y extent of each part's peaks (up-sampled pixels)
        __shared__ int ylo[kParts], yhi[kParts], xlo[kParts], xhi[kParts];
        if (lane < kParts) {
            int *e = a.yext + ((size_t)img * kParts + lane) * 4;
            ylo[lane] = -ext_lo;
            yhi[lane] = ext_hi;
            xlo[lane] = -ext_xlo;
            xhi[lane] = ext_xhi;
            e[0] = e[1] = e[2] = e[3] = INT_MIN;  // ready for the next launch
        }
        __syncwarp();
        if (lane < kLimbs) {
            const int pa = c_limb_a[lane], pb = c_limb_b[lane];
            const int na = min(offs[(pa + 1) * a.nbands] - offs[pa * a.nbands], a.cap);
            const int nb = min(offs[(pb + 1) * a.nbands] - offs[pb * a.nbands], a.cap);
            int r0 = 0, r1 = a.h, c0 = 0, c1 = a.w;
            if (a.up == 4 && na > 0 && nb > 0) {
                // samples lie between the two end points (pafprocess.cpp:319-320); the x4 bicubic of an up-sampled row Y
                // reads feature rows ((Y-2)>>2)-1 .. ((Y-2)>>2)+2 (ibp_device.cuh cubic_sample4); one more row either side
                const int lo = min(ylo[pa], ylo[pb]), hi = max(yhi[pa], yhi[pb]);
                r0 = max(0, ((lo - 2) >> 2) - 2);
                r1 = min(a.h, ((hi - 2) >> 2) + 4);
                if (r1 <= r0) { r0 = 0; r1 = a.h; }
                // the same for the columns, widened to whole 32-byte pieces (the rows of the box are staged one by one)
                const int xl = min(xlo[pa], xlo[pb]), xh = max(xhi[pa], xhi[pb]);
                c0 = max(0, ((xl - 2) >> 2) - 2) & ~7;
                c1 = min(a.w, (((xh - 2) >> 2) + 4 + 7) & ~7);
                if (c1 <= c0 || (a.w & 3) || a.w > 16380) { c0 = 0; c1 = a.w; }
            }
            // queue position: the 30 limbs of rank r follow those of rank r-1 and the "group it" items of ranks < r - lag
            a.item_meta[rank * kLimbs + lane + max(0, rank - a.group_lag)] = make_int4(img, na | (nb << 16), lane | ((c0 >> 2) << 8) | ((c1 >> 2) << 20), r0 | (r1 << 16));
        }
        // "group image of rank q" is queued kGroupLag images later (by then its limbs are finished and whoever claims it
        // does not have to wait), the last ones at the very end
        if (lane == kLimbs) {
            const int pos = (rank + a.group_lag <= a.batch - 1) ? (rank + a.group_lag + 1) * kLimbs + rank : a.batch * kLimbs + rank;
            a.item_meta[pos] = make_int4(img, 0, kLimbs, 0);
        }
    }
    // ids would no longer be a running counter: flag it; ibp_fetch turns the flag into IBP_ERR_CAPACITY
    if (__any_sync(0xffffffffu, over) && lane == 0) atomicOr(a.status + img, IBP_ST_PEAK_OVERFLOW);
}

}  // namespace ibp
